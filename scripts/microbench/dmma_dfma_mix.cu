// Do the fp64 tensor pipe (DMMA.8x8x4) and the fp64 SIMT pipe (DFMA) run concurrently on a B200 SM?  Half of the warps
// of every CTA issue DMMAs, the other half DFMAs; the sum of the two rates is compared with each alone.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void mix_kernel(double* out, int iters, int dmma_warps) {
    const int warp = threadIdx.x >> 5;
    double c[8][2];
    for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = (double)i;
    const double a = 1.0 + threadIdx.x * 1e-9, b = threadIdx.x * 1e-6;
    if (warp < dmma_warps) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                             : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
        }
    } else {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { c[i][0] = fma(c[i][0], a, b); c[i][1] = fma(c[i][1], a, b); }
        }
    }
    double s = 0.0;
    for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    double* out;
    cudaMalloc(&out, 148 * 1024 * 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000, warps = 16;
    for (int dw : {16, 0, 8, 12, 4}) {
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            mix_kernel<<<sms, warps * 32>>>(out, iters, dw);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep > 0 && ms < best) best = ms;
        }
        const double f_mma = (double)sms * dw * iters * 8 * 512.0, f_fma = (double)sms * (warps - dw) * 32 * iters * 16 * 2.0;
        printf("DMMA warps %2d / DFMA warps %2d: %.3f ms  DMMA %.1f + DFMA %.1f = %.1f TFLOP/s\n", dw, warps - dw, best,
               f_mma / (best * 1e-3) / 1e12, f_fma / (best * 1e-3) / 1e12, (f_mma + f_fma) / (best * 1e-3) / 1e12);
    }
    printf("cuda error: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
