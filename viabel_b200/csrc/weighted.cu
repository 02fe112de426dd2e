// Importance-weighted moments: the consumer of PSIS output in the reference's notebooks
// (notebooks/experiments.py:73-89, improve_with_psis):
//     slw -= max(slw); wts = exp(slw); wts /= sum(wts)
//     approx_mean = sum_i wts_i x_i
//     approx_cov  = np.cov(samples, aweights=wts, ddof=0) = sum_i wts_i (x_i - m)(x_i - m)' / sum_i wts_i
// Three streaming passes over (n, d) samples + (n) log weights: max, weighted column sums, weighted
// scatter (fp64 GEMM with functor loaders, split-K over n, fixed-order sums).
#include "common.cuh"
#include "gemm_f64.cuh"

namespace vb {

static inline int grid_for_w(long long work_items, int threads, int max_blocks) {
    long long g = (work_items + threads - 1) / threads;
    if (g > max_blocks) g = max_blocks;
    if (g < 1) g = 1;
    return (int)g;
}

// w_i = exp(slw_i - M); block sums of w
__global__ void __launch_bounds__(256)
wm_weights_kernel(const double* __restrict__ slw, long long n, const double* __restrict__ M_ptr,
                  double* __restrict__ w, double* __restrict__ block_out) {
    __shared__ double scratch[32];
    const double M = *M_ptr;
    double s = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double v = exp(slw[i] - M);
        w[i] = v;
        s += v;
    }
    s = block_sum(s, scratch);
    if (threadIdx.x == 0) block_out[blockIdx.x] = s;
}
__global__ void __launch_bounds__(256)
wm_sum_blocks_kernel(const double* __restrict__ blocks, int G, double* __restrict__ out) {
    __shared__ double scratch[32];
    double s = 0.0;
    for (int g = threadIdx.x; g < G; g += blockDim.x) s += blocks[g];
    s = block_sum(s, scratch);
    if (threadIdx.x == 0) *out = s;
}
// column sums with row weights, chunked over rows: out[chunk][c] = sum_{r in chunk} w_r x[r, c]
template <typename T>
__global__ void wm_colsum_kernel(long long n, int D, const T* __restrict__ x, const double* __restrict__ w,
                                 long long rows_per_chunk, double* __restrict__ out) {
    __shared__ double sh[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const long long r0 = (long long)blockIdx.y * rows_per_chunk;
    long long r1 = r0 + rows_per_chunk;
    if (r1 > n) r1 = n;
    double a0 = 0.0;
    if (c < D)
        for (long long r = r0 + threadIdx.y; r < r1; r += 8) a0 = fma(w[r], (double)x[r * D + c], a0);
    sh[threadIdx.y][threadIdx.x] = a0;
    __syncthreads();
    if (threadIdx.y == 0 && c < D) {
        double s0 = 0.0;
#pragma unroll
        for (int y = 0; y < 8; ++y) s0 += sh[y][threadIdx.x];
        out[(size_t)blockIdx.y * D + c] = s0;
    }
}
// mean[c] = (sum over chunks) / sum_w
__global__ void wm_mean_kernel(const double* __restrict__ parts, int chunks, int D, const double* __restrict__ sum_w,
                               double* __restrict__ mean) {
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < D; c += gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < chunks; ++k) s += parts[(size_t)k * D + c];
        mean[c] = s / *sum_w;
    }
}
template <typename T>
struct WCenteredT {     // A(a, i) = w_i (x[i, a] - mean[a])
    static constexpr bool kContigCol = false;
    const T* x; long long ld; const double* m; const double* w;
    __device__ __forceinline__ double operator()(long long a, long long i) const {
        return w[i] * ((double)x[i * ld + a] - m[a]);
    }
};
template <typename T>
struct CenteredRowsN {  // B(i, b) = x[i, b] - mean[b]
    static constexpr bool kContigCol = true;
    const T* x; long long ld; const double* m;
    __device__ __forceinline__ double operator()(long long i, long long b) const { return (double)x[i * ld + b] - m[b]; }
};
struct WSplitStore {
    double* p; long long mn; long long ld;
    __device__ __forceinline__ void operator()(long long i, long long j, double acc, int split) const {
        p[(size_t)split * mn + i * ld + j] = acc;
    }
};
__global__ void wm_cov_kernel(const double* __restrict__ parts, int splits, long long mn, const double* __restrict__ sum_w,
                              double* __restrict__ cov) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < mn; e += (long long)gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < splits; ++k) s += parts[(size_t)k * mn + e];
        cov[e] = s / *sum_w;
    }
}

static int wm_chunks(long long n, int D) {
    const int col_tiles = (D + 31) / 32;
    long long want = (kNumSMs * 4 + col_tiles - 1) / col_tiles;
    const long long max_chunks = (n + 63) / 64;
    if (want > max_chunks) want = max_chunks;
    if (want < 1) want = 1;
    return (int)want;
}
static int wm_splits(long long n, int D) {
    const long long tiles = ((D + kGBM - 1) / kGBM) * (long long)((D + kGBN - 1) / kGBN);
    long long s = (kNumSMs * 2 + tiles - 1) / tiles;
    const long long max_s = (n + 255) / 256;
    if (s > max_s) s = max_s;
    if (s < 1) s = 1;
    return (int)s;
}
struct WmLayout { int G, chunks, splits; size_t off_M, off_sumw, off_blocks, off_w, off_colparts, off_parts, off_maxws, total; };
static WmLayout wm_layout(long long n, int D) {
    WmLayout L;
    L.G = grid_for_w(n, 256, kNumSMs * 8);
    L.chunks = wm_chunks(n, D);
    L.splits = wm_splits(n, D);
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o += align_up(bytes, 256); return at; };
    L.off_M = take(8);
    L.off_sumw = take(8);
    L.off_blocks = take((size_t)L.G * 8);
    L.off_w = take((size_t)n * 8);
    L.off_colparts = take((size_t)L.chunks * D * 8);
    L.off_parts = take((size_t)L.splits * D * D * 8);
    L.off_maxws = take(vb_max_workspace_bytes(n));
    L.total = o;
    return L;
}

}  // namespace vb

using namespace vb;

extern "C" size_t vb_weighted_moments_workspace_bytes(long long n, int dim) {
    if (n < 1 || dim < 1) return 0;
    return wm_layout(n, dim).total + 256;
}

extern "C" int vb_weighted_moments(int dtype, const void* d_x, const double* d_log_weights, long long n, int dim,
                                   double* d_out_mean, double* d_out_cov_or_null,
                                   void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_x || !d_log_weights || !d_out_mean || !d_ws || n < 1 || dim < 1) return VB_ERR_ARG;
    if (dtype != VB_F32 && dtype != VB_F64) return VB_ERR_ARG;
    if (ws_bytes < vb_weighted_moments_workspace_bytes(n, dim)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const WmLayout L = wm_layout(n, dim);
    char* ws = (char*)(((uintptr_t)d_ws + 255) & ~(uintptr_t)255);
    double* M_ptr = (double*)(ws + L.off_M);
    double* sum_w = (double*)(ws + L.off_sumw);
    double* blocks = (double*)(ws + L.off_blocks);
    double* w = (double*)(ws + L.off_w);
    double* colparts = (double*)(ws + L.off_colparts);
    double* parts = (double*)(ws + L.off_parts);
    int rc = vb_max(d_log_weights, n, M_ptr, ws + L.off_maxws, vb_max_workspace_bytes(n), stream);
    if (rc != VB_OK) return rc;
    wm_weights_kernel<<<L.G, 256, 0, st>>>(d_log_weights, n, M_ptr, w, blocks);
    VB_LAUNCH_CHECK();
    wm_sum_blocks_kernel<<<1, 256, 0, st>>>(blocks, L.G, sum_w);
    VB_LAUNCH_CHECK();
    const long long rpc = (n + L.chunks - 1) / L.chunks;
    const dim3 cg((dim + 31) / 32, L.chunks), cb(32, 8);
    if (dtype == VB_F64) wm_colsum_kernel<double><<<cg, cb, 0, st>>>(n, dim, (const double*)d_x, w, rpc, colparts);
    else wm_colsum_kernel<float><<<cg, cb, 0, st>>>(n, dim, (const float*)d_x, w, rpc, colparts);
    VB_LAUNCH_CHECK();
    wm_mean_kernel<<<grid_for_w(dim, 256, 8), 256, 0, st>>>(colparts, L.chunks, dim, sum_w, d_out_mean);
    VB_LAUNCH_CHECK();
    if (d_out_cov_or_null) {
        const long long mn = (long long)dim * dim;
        if (dtype == VB_F64)
            rc = launch_gemm_f64(dim, dim, n, L.splits, WCenteredT<double>{(const double*)d_x, dim, d_out_mean, w},
                                 CenteredRowsN<double>{(const double*)d_x, dim, d_out_mean}, WSplitStore{parts, mn, dim}, st);
        else
            rc = launch_gemm_f64(dim, dim, n, L.splits, WCenteredT<float>{(const float*)d_x, dim, d_out_mean, w},
                                 CenteredRowsN<float>{(const float*)d_x, dim, d_out_mean}, WSplitStore{parts, mn, dim}, st);
        if (rc != VB_OK) return rc;
        wm_cov_kernel<<<grid_for_w(mn, 256, kNumSMs * 4), 256, 0, st>>>(parts, L.splits, mn, sum_w, d_out_cov_or_null);
        VB_LAUNCH_CHECK();
    }
    return VB_OK;
}
