// Convergence diagnostics of the R-hat-monitored optimisers (reference: viabel/functions.py:8-77, used by
// rmsprop_IA_optimize_with_rhat / adam_IA_optimize_with_rhat, viabel/vb.py:486-522).
//
//   vb_rhat_windows      split R-hat of any number of iteration windows of a (n_chains, n_iters, K) fp64 history
//                        in one pass of two launches: every (window, half-chain, 32- or 64-parameter tile) is one CTA
//   vb_iterate_average   running mean of the iterates from `start` on (stochastic_iterate_averaging)
//
// Both are streaming reductions over the parameter history (HBM bound when the history is large: C3 keeps
// 131 840 parameters per iterate); the history is read twice by the R-hat pass (mean, then squared deviations
// about it -- the two-pass form the reference uses) and once by the averaging.
#include "common.cuh"

namespace vb {
namespace {

constexpr int kRhatTX = 32;      // parameters per CTA (one 256-byte row segment per warp load)
constexpr int kRhatTY = 8;       // iteration stripes per CTA (256 threads: 8 CTAs per SM overlap their load / reduce phases)

// ws layout: mean[n_windows][2C][K], s2[n_windows][2C][K]
// VEC = 2: two adjacent parameters per thread through 16-byte loads (K even, 16-byte aligned history): a warp
// reads 512 contiguous bytes per iteration instead of 256.
template <int VEC>
__global__ void __launch_bounds__(kRhatTX * kRhatTY)
rhat_halfchain_kernel(const double* __restrict__ chains, int n_chains, long long n_iters, long long K,
                      const long long* __restrict__ windows, double* __restrict__ out_mean,
                      double* __restrict__ out_s2) {
    __shared__ double red[VEC][kRhatTY][kRhatTX + 1];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const long long p = ((long long)blockIdx.x * kRhatTX + tx) * VEC;
    const int hc = blockIdx.y;                                  // half-chain 2c + h, as the reference's reshape orders them
    const int w = blockIdx.z;
    const long long begin = windows[2 * w], len = windows[2 * w + 1];
    const long long n = len / 2;                                // iterations per half-chain
    const int c = hc >> 1, h = hc & 1;
    const double* base = chains + ((size_t)c * n_iters + begin + (long long)h * n) * K;
    const bool live = p < K;                                    // VEC = 2 implies K even: p + 1 < K as well
    auto load = [&](long long t, double (&v)[VEC]) {
        if (VEC == 2) {
            const double2 d2 = *reinterpret_cast<const double2*>(base + (size_t)t * K + p);
            v[0] = d2.x; v[VEC - 1] = d2.y;
        } else {
            v[0] = base[(size_t)t * K + p];
        }
    };
    // pass 1: mean
    double s[VEC] = {};
    if (live) {
#pragma unroll 8
        for (long long t = ty; t < n; t += kRhatTY) {
            double v[VEC];
            load(t, v);
#pragma unroll
            for (int e = 0; e < VEC; ++e) s[e] += v[e];
        }
    }
#pragma unroll
    for (int e = 0; e < VEC; ++e) red[e][ty][tx] = s[e];
    __syncthreads();
    if (ty < VEC) {
        double a = 0.0;
#pragma unroll
        for (int j = 0; j < kRhatTY; ++j) a += red[ty][j][tx];  // fixed order
        red[ty][0][tx] = a / (double)n;
    }
    __syncthreads();
    double mean[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) mean[e] = red[e][0][tx];
    __syncthreads();
    // pass 2: squared deviations about the mean (functions.py:25); the second read is an L1 / L2 hit
    double q[VEC] = {};
    if (live) {
#pragma unroll 8
        for (long long t = ty; t < n; t += kRhatTY) {
            double v[VEC];
            load(t, v);
#pragma unroll
            for (int e = 0; e < VEC; ++e) {
                const double d = v[e] - mean[e];
                q[e] = fma(d, d, q[e]);
            }
        }
    }
#pragma unroll
    for (int e = 0; e < VEC; ++e) red[e][ty][tx] = q[e];
    __syncthreads();
    if (ty < VEC && live) {
        double a = 0.0;
#pragma unroll
        for (int j = 0; j < kRhatTY; ++j) a += red[ty][j][tx];
        const size_t o = ((size_t)w * (2 * n_chains) + hc) * K + p + ty;
        out_mean[o] = mean[ty];
        out_s2[o] = a / (double)(n - 1);                        // n = 1: 0/0 = NaN, dropped by the nanmean below
    }
}

// per (window, parameter): B, W, var_hat, R-hat (functions.py:23-31)
__global__ void rhat_finish_kernel(const double* __restrict__ mean, const double* __restrict__ s2, int n_chains,
                                   long long K, const long long* __restrict__ windows, int n_windows,
                                   double* __restrict__ out_var_hat, double* __restrict__ out_rhat) {
    const long long total = (long long)n_windows * K;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long w = e / K, p = e - w * K;
        const int m = 2 * n_chains;
        const double n = (double)(windows[2 * w + 1] / 2);
        const double* mu = mean + (size_t)w * m * K + p;
        const double* sv = s2 + (size_t)w * m * K + p;
        double grand = 0.0;
        for (int j = 0; j < m; ++j) grand += mu[(size_t)j * K];
        grand /= (double)m;
        double b = 0.0, wsum = 0.0;
        int cnt = 0;
        for (int j = 0; j < m; ++j) {
            const double d = mu[(size_t)j * K] - grand;
            b += d * d;
            const double v = sv[(size_t)j * K];
            if (v == v) { wsum += v; ++cnt; }                    // nanmean
        }
        const double B = n * b / (double)(m - 1);
        const double W = wsum / (double)cnt + 1e-8;             // jitter, functions.py:11,28
        const double var_hat = (n - 1.0) / n + B / (n * W);
        out_var_hat[e] = var_hat;
        out_rhat[e] = sqrt(var_hat);
    }
}

// one thread per column, sequential in the iteration (the summation order of numpy's cumsum along axis 0)
__global__ void iterate_average_kernel(const double* __restrict__ est, long long N, long long K, long long ld,
                                       long long start, double* __restrict__ out) {
    for (long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x; p < K;
         p += (long long)gridDim.x * blockDim.x) {
        double s = 0.0;
        long long t = start;
        for (; t + 8 <= N; t += 8) {                 // eight independent loads in flight, the adds stay in order
            double v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = est[(size_t)(t + u) * ld + p];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                s += v[u];
                out[(size_t)(t + u - start) * K + p] = s / (double)(t + u - start + 1);
            }
        }
        for (; t < N; ++t) {
            s += est[(size_t)t * ld + p];
            out[(size_t)(t - start) * K + p] = s / (double)(t - start + 1);
        }
    }
}

}  // namespace
}  // namespace vb

extern "C" size_t vb_rhat_workspace_bytes(int n_chains, long long n_params, int n_windows) {
    if (n_chains < 1 || n_params < 1 || n_windows < 1) return 0;
    return (size_t)2 * n_windows * (2 * (size_t)n_chains) * (size_t)n_params * sizeof(double) + 256;
}

extern "C" int vb_rhat_windows(const double* d_chains, int n_chains, long long n_iters, long long n_params,
                               const long long* d_windows, int n_windows, double* d_out_var_hat,
                               double* d_out_rhat, void* d_ws, size_t ws_bytes, void* stream) {
    using namespace vb;
    if (!d_chains || !d_windows || !d_out_var_hat || !d_out_rhat || n_chains < 1 || n_iters < 2 || n_params < 1 ||
        n_windows < 1 || n_windows > 65535 || 2 * n_chains > 65535)
        return VB_ERR_ARG;
    if (!d_ws || ws_bytes < vb_rhat_workspace_bytes(n_chains, n_params, n_windows)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    double* mean = (double*)(((uintptr_t)d_ws + 255) & ~(uintptr_t)255);
    double* s2 = mean + (size_t)n_windows * 2 * n_chains * n_params;
    const bool vec2 = (n_params % 2 == 0) && ((reinterpret_cast<uintptr_t>(d_chains) & 15u) == 0);
    const long long per_cta = (long long)kRhatTX * (vec2 ? 2 : 1);
    const dim3 grid((unsigned)((n_params + per_cta - 1) / per_cta), (unsigned)(2 * n_chains), (unsigned)n_windows);
    if (vec2)
        rhat_halfchain_kernel<2><<<grid, dim3(kRhatTX, kRhatTY), 0, st>>>(d_chains, n_chains, n_iters, n_params,
                                                                        d_windows, mean, s2);
    else
        rhat_halfchain_kernel<1><<<grid, dim3(kRhatTX, kRhatTY), 0, st>>>(d_chains, n_chains, n_iters, n_params,
                                                                        d_windows, mean, s2);
    VB_LAUNCH_CHECK();
    const long long total = (long long)n_windows * n_params;
    long long blocks = (total + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    rhat_finish_kernel<<<(unsigned)blocks, 256, 0, st>>>(mean, s2, n_chains, n_params, d_windows, n_windows,
                                                       d_out_var_hat, d_out_rhat);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" int vb_iterate_average(const double* d_est, long long n_iters, long long n_params, long long ld,
                                  long long start, double* d_out, void* stream) {
    using namespace vb;
    if (!d_est || !d_out || n_params < 1 || ld < n_params || start < 0 || start >= n_iters) return VB_ERR_ARG;
    long long blocks = (n_params + 127) / 128;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    iterate_average_kernel<<<(unsigned)blocks, 128, 0, (cudaStream_t)stream>>>(d_est, n_iters, n_params, ld, start, d_out);
    VB_LAUNCH_CHECK();
    return VB_OK;
}
