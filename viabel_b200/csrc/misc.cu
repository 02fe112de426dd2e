#include "common.cuh"
extern "C" int vb_abi_version(void) { return 1; }
extern "C" int vb_compiled_sm(void) { return 100; }
extern "C" const char* vb_error_string(int status) {
    switch (status) {
        case VB_OK: return "ok";
        case VB_ERR_ARG: return "invalid argument";
        case VB_ERR_UNSUPPORTED: return "unsupported configuration";
        case VB_ERR_WORKSPACE: return "workspace too small";
        case VB_ERR_ALIGN: return "misaligned pointer";
        case VB_ERR_TOO_LARGE: return "problem larger than the fixed scratch of this entry point allows";
        default: return status > 0 ? cudaGetErrorString((cudaError_t)status) : "unknown error";
    }
}

// ------------------------------------------------------------------------------------------ measured fp64 MMA peak
// The roofline denominator of the fp64 tensor-core kernels (fr_chain.cu), measured instead of quoted: every SM runs
// eight warps of back-to-back DMMA.8x8x4 on register operands (the only fp64 MMA shape sm_100a has: m16n8k16 etc.
// compile to sequences of it), eight independent accumulators per warp.  B200: 37.1 TFLOP/s (DFMA: 34).
namespace vb {
__global__ void __launch_bounds__(256) dmma_peak_kernel(double* out, int iters) {
    double c[8][2];
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = 0.0;
    const double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-6;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace vb
extern "C" size_t vb_measure_dmma_peak_workspace_bytes(void) { return (size_t)1024 * 256 * sizeof(double); }
extern "C" int vb_measure_dmma_peak(double* h_out_tflops, void* d_ws, size_t ws_bytes, void* stream) {
    if (!h_out_tflops || !d_ws) return VB_ERR_ARG;
    if (ws_bytes < vb_measure_dmma_peak_workspace_bytes()) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const int sms = vb::device_sm_count();
    const int iters = 8192;
    cudaEvent_t e0, e1;
    VB_CUDA_TRY(cudaEventCreate(&e0));
    VB_CUDA_TRY(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        VB_CUDA_TRY(cudaEventRecord(e0, st));
        vb::dmma_peak_kernel<<<sms, 256, 0, st>>>((double*)d_ws, iters);
        VB_CUDA_TRY(cudaEventRecord(e1, st));
        VB_CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0.f;
        VB_CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    VB_LAUNCH_CHECK();
    *h_out_tflops = (double)sms * 8.0 * iters * 8.0 * 512.0 / ((double)best * 1e-3) / 1e12;
    return VB_OK;
}
