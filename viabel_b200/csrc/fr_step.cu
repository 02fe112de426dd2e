// Per-sample kernels of the full-rank Student-t family (viabel/vb.py:192-233,
// viabel/_distributions.py:8-38) — fp64 path.  The O(d^3) parameter chain is driven from
// the Python layer with vb_gemm and an eigendecomposition; see SURVEY.md Appendix A.2.
#include "common.cuh"
#include "gemm_f64.cuh"
#include "dense.cuh"

namespace vb {

static inline int grid_for(long long work_items, int threads, int max_blocks) {
    long long g = (work_items + threads - 1) / threads;
    if (g > max_blocks) g = max_blocks;
    if (g < 1) g = 1;
    return (int)g;
}

// x_i = mu + (z_i S) / sqrt(chi2_i / df)                 (vb.py:204-208)
template <typename T>
struct SampleStore {
    T* x; long long ld; const double* mu; const T* chi2; double df;
    __device__ __forceinline__ void operator()(long long i, long long j, double acc, int) const {
        x[i * ld + j] = (T)(mu[j] + acc / sqrt((double)chi2[i] / df));
    }
};

// rows: maha_i = |z_i|^2 / (chi2_i/df); lw_i = logp_i - (c - 1/2 (df+d) log1p(maha_i/df))
template <typename T>
__global__ void fr_rows_kernel(double df, long long n, int D, const T* __restrict__ z, const T* __restrict__ chi2,
                               const double* __restrict__ logp, double c_total, double* __restrict__ lw,
                               double* __restrict__ block_max_out) {
    __shared__ double scratch[32];
    const int lane = threadIdx.x & 31;
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    double m = -INFINITY;
    for (long long i = warp; i < n; i += nwarps) {
        double s = 0.0;
        for (int j = lane; j < D; j += 32) {
            const double v = (double)z[i * D + j];
            s = fma(v, v, s);
        }
        s = warp_sum(s);
        const double maha = s / ((double)chi2[i] / df);
        const double v = logp[i] - (c_total - 0.5 * (df + D) * log1p(maha / df));
        if (lane == 0) lw[i] = v;
        m = fmax(m, v);
    }
    m = block_max(m, scratch);
    if (threadIdx.x == 0) block_max_out[blockIdx.x] = m;
}
__global__ void fr_max_finish_kernel(const double* __restrict__ vals, int count, double* __restrict__ out) {
    __shared__ double scratch[32];
    double m = -INFINITY;
    for (int i = threadIdx.x; i < count; i += blockDim.x) m = fmax(m, vals[i]);
    m = block_max(m, scratch);
    if (threadIdx.x == 0) *out = m;
}
// w_i, and block sums of (w, logp, lw)
__global__ void fr_weights_kernel(int objective, double alpha, long long n, const double* __restrict__ lw,
                                  const double* __restrict__ logp, const double* __restrict__ M_ptr,
                                  double* __restrict__ w_out, double* __restrict__ block_out /* [grid][3] */) {
    __shared__ double scratch[32];
    const double M = *M_ptr;
    double sw = 0.0, slp = 0.0, slw = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const double l = lw[i];
        const double w = (objective == VB_OBJ_CHIVI) ? exp(alpha * (l - M)) : 1.0;
        w_out[i] = w;
        sw += w; slp += logp[i]; slw += l;
    }
    sw = block_sum(sw, scratch);
    slp = block_sum(slp, scratch);
    slw = block_sum(slw, scratch);
    if (threadIdx.x == 0) {
        block_out[3 * blockIdx.x] = sw; block_out[3 * blockIdx.x + 1] = slp; block_out[3 * blockIdx.x + 2] = slw;
    }
}
__global__ void fr_rows_finish_kernel(const double* __restrict__ M_ptr, const double* __restrict__ blocks, int G,
                                      double* __restrict__ out) {
    // one warp, lane-strided partial sums + a fixed butterfly (a serial loop over ~1000 blocks cost 30 us)
    if (blockIdx.x != 0 || threadIdx.x >= 32) return;
    double sw = 0, slp = 0, slw = 0;
    for (int g = threadIdx.x; g < G; g += 32) { sw += blocks[3 * g]; slp += blocks[3 * g + 1]; slw += blocks[3 * g + 2]; }
    sw = warp_sum(sw); slp = warp_sum(slp); slw = warp_sum(slw);
    if (threadIdx.x != 0) return;
    out[0] = *M_ptr; out[1] = sw; out[2] = slp; out[3] = slw;
}

// G_S = sum_i (w_i / s_i) z_i' g_i  :  A(a, i) = z[i, a] * w_i / s_i ,  B(i, b) = g[i, b]
template <typename T>
struct ScaledZt {
    static constexpr bool kContigCol = false;
    const T* z; long long ld; const double* w; const T* chi2; double df;
    __device__ __forceinline__ double operator()(long long a, long long i) const {
        return (double)z[i * ld + a] * (w[i] / sqrt((double)chi2[i] / df));
    }
};
struct SplitStoreFr {
    double* p; long long mn; long long ld;
    __device__ __forceinline__ void operator()(long long i, long long j, double acc, int split) const {
        p[(size_t)split * mn + i * ld + j] = acc;
    }
};
__global__ void fr_split_sum_kernel(const double* __restrict__ parts, int splits, long long mn, double* __restrict__ out) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < mn; e += (long long)gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < splits; ++k) s += parts[(size_t)k * mn + e];
        out[e] = s;
    }
}
// g_mu = sum_i w_i g_i : column sums with row weights, chunked over rows
template <typename T>
__global__ void fr_wcolsum_kernel(long long n, int D, const T* __restrict__ g, const double* __restrict__ w,
                                  long long rows_per_chunk, double* __restrict__ out /* [chunks][D] */) {
    __shared__ double sh[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const long long r0 = (long long)blockIdx.y * rows_per_chunk;
    long long r1 = r0 + rows_per_chunk;
    if (r1 > n) r1 = n;
    double a0 = 0.0;
    if (c < D)
        for (long long r = r0 + threadIdx.y; r < r1; r += 8) a0 = fma(w[r], (double)g[r * D + c], a0);
    sh[threadIdx.y][threadIdx.x] = a0;
    __syncthreads();
    if (threadIdx.y == 0 && c < D) {
        double s0 = 0.0;
#pragma unroll
        for (int y = 0; y < 8; ++y) s0 += sh[y][threadIdx.x];
        out[(size_t)blockIdx.y * D + c] = s0;
    }
}

// logdensity of arbitrary x: maha_i = |(x_i - mu) U|^2, U = u sqrt(pinv(s))     (_distributions.py:27-37)
template <typename T>
struct CenteredRowsFr {
    static constexpr bool kContigCol = true;
    const T* x; long long ld; const double* m;
    __device__ __forceinline__ double operator()(long long i, long long k) const { return (double)x[i * ld + k] - m[k]; }
};
struct PlainStore {
    double* c; long long ld;
    __device__ __forceinline__ void operator()(long long i, long long j, double acc, int) const { c[i * ld + j] = acc; }
};
__global__ void fr_logdensity_rows_kernel(long long n, int D, const double* __restrict__ Y, double c_total, double df,
                                          double* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = warp; i < n; i += nwarps) {
        double s = 0.0;
        for (int j = lane; j < D; j += 32) { const double v = Y[i * D + j]; s = fma(v, v, s); }
        s = warp_sum(s);
        // df = inf: the multivariate normal (_distributions.py:24-25), c_total carries its constant
        if (lane == 0) out[i] = isinf(df) ? c_total - 0.5 * s : c_total - 0.5 * (df + D) * log(1.0 + s / df);
    }
}

static int fr_splits(long long n, int D) {
    const long long tiles = ((D + kGBM - 1) / kGBM) * (long long)((D + kGBN - 1) / kGBN);
    long long s = (kNumSMs * 2 + tiles - 1) / tiles;
    const long long max_s = (n + 255) / 256;
    if (s > max_s) s = max_s;
    if (s < 1) s = 1;
    return (int)s;
}
static int fr_chunks(long long n, int D) {
    const int col_tiles = (D + 31) / 32;
    long long want = (kNumSMs * 4 + col_tiles - 1) / col_tiles;
    const long long max_chunks = (n + 63) / 64;
    if (want > max_chunks) want = max_chunks;
    if (want < 1) want = 1;
    return (int)want;
}

// plain GEMM epilogue: C = alpha * acc + beta * C
template <typename T>
struct AxpbyStore {
    T* c; long long ld; double alpha, beta;
    __device__ __forceinline__ void operator()(long long i, long long j, double acc, int) const {
        const double prev = (beta == 0.0) ? 0.0 : beta * (double)c[i * ld + j];
        c[i * ld + j] = (T)(alpha * acc + prev);
    }
};

}  // namespace vb

using namespace vb;

extern "C" size_t vb_fr_sample_workspace_bytes(int dtype, long long n, int dim) {
    if (n < 1 || dim < 1) return 0;
    return dtype == VB_F32 ? fr_sample_tf32_workspace_bytes(n, dim) : 0;
}

extern "C" int vb_fr_sample(int dtype, const double* d_mu, const void* d_scale, double df, long long n, int dim,
                            const void* d_z, const void* d_chi2, void* d_out_x, void* d_ws, size_t ws_bytes,
                            void* stream) {
    if (!d_mu || !d_scale || !d_z || !d_chi2 || !d_out_x || n < 1 || dim < 1 || !(df > 2.0)) return VB_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == VB_F64)
        return launch_gemm_f64(n, dim, dim, 1, RowMajor<double>{(const double*)d_z, dim},
                               RowMajor<double>{(const double*)d_scale, dim},
                               SampleStore<double>{(double*)d_out_x, dim, d_mu, (const double*)d_chi2, df}, st);
    if (dtype == VB_F32)
        return fr_sample_tf32(d_mu, (const float*)d_scale, df, n, dim, (const float*)d_z, (const float*)d_chi2,
                              (float*)d_out_x, d_ws, ws_bytes, st);
    return VB_ERR_ARG;
}

extern "C" size_t vb_fr_rows_workspace_bytes(long long n) {
    (void)n;
    return 256 + (size_t)kNumSMs * 8 * 8 + (size_t)kNumSMs * 4 * 24 + 256;
}

extern "C" int vb_fr_rows(int objective, double alpha, double df, int dtype, long long n, int dim,
                          const void* d_z, const void* d_chi2, const double* d_logp,
                          double logq_const_minus_entropy,
                          double* d_out_lw, double* d_out_w, double* d_out_rows,
                          void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_z || !d_chi2 || !d_logp || !d_out_lw || !d_out_w || !d_out_rows || !d_ws || n < 1 || dim < 1)
        return VB_ERR_ARG;
    if (ws_bytes < vb_fr_rows_workspace_bytes(n)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)d_ws;
    double* M_ptr = (double*)ws;
    double* bmax = (double*)(ws + 256);
    double* wb = bmax + (size_t)kNumSMs * 8;
    const int g_rows = grid_for(n * 32, 256, kNumSMs * 8);
    const int g_w = grid_for(n, 256, kNumSMs * 4);
    if (dtype == VB_F64)
        fr_rows_kernel<double><<<g_rows, 256, 0, st>>>(df, n, dim, (const double*)d_z, (const double*)d_chi2, d_logp,
                                                       logq_const_minus_entropy, d_out_lw, bmax);
    else if (dtype == VB_F32)
        fr_rows_kernel<float><<<g_rows, 256, 0, st>>>(df, n, dim, (const float*)d_z, (const float*)d_chi2, d_logp,
                                                      logq_const_minus_entropy, d_out_lw, bmax);
    else
        return VB_ERR_ARG;
    VB_LAUNCH_CHECK();
    fr_max_finish_kernel<<<1, 256, 0, st>>>(bmax, g_rows, M_ptr);
    VB_LAUNCH_CHECK();
    fr_weights_kernel<<<g_w, 256, 0, st>>>(objective, alpha, n, d_out_lw, d_logp, M_ptr, d_out_w, wb);
    VB_LAUNCH_CHECK();
    fr_rows_finish_kernel<<<1, 32, 0, st>>>(M_ptr, wb, g_w, d_out_rows);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" size_t vb_fr_weighted_workspace_bytes(long long n, int dim) {
    if (n < 1 || dim < 1) return 0;
    const size_t a = (size_t)fr_splits(n, dim) * dim * dim * 8;
    const size_t b = (size_t)fr_chunks(n, dim) * dim * 8;
    const size_t t = fr_weighted_tf32_workspace_bytes(n, dim) + b + 256;     // fp32 storage: tensor path
    return (a + b + 512) > t ? (a + b + 512) : t;
}

static int fr_weighted_impl(int dtype, double df, long long n, int dim, const void* d_z, const float* d_z_planes_or_null,
                            const void* d_chi2, const double* d_w, const void* d_grad,
                            double* d_out_gmu, double* d_out_GS, void* d_ws, size_t ws_bytes, void* stream);
extern "C" int vb_fr_weighted(int dtype, double df, long long n, int dim, const void* d_z,
                              const void* d_chi2, const double* d_w, const void* d_grad,
                              double* d_out_gmu, double* d_out_GS, void* d_ws, size_t ws_bytes, void* stream) {
    return fr_weighted_impl(dtype, df, n, dim, d_z, nullptr, d_chi2, d_w, d_grad, d_out_gmu, d_out_GS, d_ws, ws_bytes, stream);
}
// fp32 storage with the 3xTF32 operand planes of z already made (vb_base_draws_fr_split, or vb_split_f32 of z): the
// product reads them instead of splitting z again
extern "C" int vb_fr_weighted_split(double df, long long n, int dim, const float* d_z, const float* d_z_planes,
                                    const float* d_chi2, const double* d_w, const float* d_grad,
                                    double* d_out_gmu, double* d_out_GS, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_z_planes) return VB_ERR_ARG;
    if (reinterpret_cast<uintptr_t>(d_z_planes) & 15u) return VB_ERR_ALIGN;
    return fr_weighted_impl(VB_F32, df, n, dim, d_z, d_z_planes, d_chi2, d_w, d_grad, d_out_gmu, d_out_GS, d_ws, ws_bytes, stream);
}
static int fr_weighted_impl(int dtype, double df, long long n, int dim, const void* d_z, const float* d_z_planes_or_null,
                            const void* d_chi2, const double* d_w, const void* d_grad,
                            double* d_out_gmu, double* d_out_GS, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_z || !d_chi2 || !d_w || !d_grad || !d_out_gmu || !d_out_GS || !d_ws || n < 1 || dim < 1) return VB_ERR_ARG;
    if (dtype != VB_F32 && dtype != VB_F64) return VB_ERR_ARG;
    if (ws_bytes < vb_fr_weighted_workspace_bytes(n, dim)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const int splits = fr_splits(n, dim);
    const int chunks = fr_chunks(n, dim);
    const long long mn = (long long)dim * dim;
    double* parts = (double*)d_ws;
    double* colparts = parts + (size_t)splits * mn;
    const long long rpc = (n + chunks - 1) / chunks;
    int rc;
    if (dtype == VB_F64) {
        rc = launch_gemm_f64(dim, dim, n, splits,
                             ScaledZt<double>{(const double*)d_z, dim, d_w, (const double*)d_chi2, df},
                             RowMajor<double>{(const double*)d_grad, dim}, SplitStoreFr{parts, mn, dim}, st);
        if (rc != VB_OK) return rc;
        fr_wcolsum_kernel<double><<<dim3((dim + 31) / 32, chunks), dim3(32, 8), 0, st>>>(n, dim, (const double*)d_grad,
                                                                                       d_w, rpc, colparts);
    } else {
        // fp32 storage: G_S on the tensor path (3xTF32, split-K), g_mu by the fp64-accumulating column sum
        double* colparts32 = (double*)((char*)d_ws + ((fr_weighted_tf32_workspace_bytes(n, dim) + 255) & ~(size_t)255));
        const bool fused = fr_weighted_colsum_fused(n, dim);
        rc = fr_weighted_GS_tf32(df, n, dim, (const float*)d_z, (const float*)d_chi2, d_w, (const float*)d_grad,
                                 d_out_GS, d_ws, fr_weighted_tf32_workspace_bytes(n, dim), st, fused ? d_out_gmu : nullptr,
                                 d_z_planes_or_null);
        if (rc != VB_OK) return rc;
        if (fused) return VB_OK;
        fr_wcolsum_kernel<float><<<dim3((dim + 31) / 32, chunks), dim3(32, 8), 0, st>>>(n, dim, (const float*)d_grad,
                                                                                      d_w, rpc, colparts32);
        VB_LAUNCH_CHECK();
        fr_split_sum_kernel<<<grid_for(dim, 256, 8), 256, 0, st>>>(colparts32, chunks, dim, d_out_gmu);
        VB_LAUNCH_CHECK();
        return VB_OK;
    }
    VB_LAUNCH_CHECK();
    fr_split_sum_kernel<<<grid_for(mn, 256, kNumSMs * 4), 256, 0, st>>>(parts, splits, mn, d_out_GS);
    VB_LAUNCH_CHECK();
    fr_split_sum_kernel<<<grid_for(dim, 256, 8), 256, 0, st>>>(colparts, chunks, dim, d_out_gmu);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" size_t vb_gemm_workspace_bytes(int dtype, long long m, long long n, long long k) {
    if (dtype == VB_F32) return gemm_tf32x3_workspace_bytes(m, n, k);
    return 0;
}

extern "C" int vb_gemm(int dtype, int transa, int transb, long long m, long long n, long long k,
                       double alpha, const void* d_A, long long lda, const void* d_B, long long ldb,
                       double beta, void* d_C, long long ldc, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_A || !d_B || !d_C || m < 1 || n < 1 || k < 1) return VB_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == VB_F32)
        return gemm_tf32x3(transa, transb, m, n, k, (float)alpha, (const float*)d_A, lda, (const float*)d_B, ldb,
                           (float)beta, (float*)d_C, ldc, d_ws, ws_bytes, st);
    if (dtype != VB_F64) return VB_ERR_ARG;
    const double* A = (const double*)d_A;
    const double* B = (const double*)d_B;
    AxpbyStore<double> epi{(double*)d_C, ldc, alpha, beta};
    if (!transa && !transb) return launch_gemm_f64(m, n, k, 1, RowMajor<double>{A, lda}, RowMajor<double>{B, ldb}, epi, st);
    if (!transa && transb) return launch_gemm_f64(m, n, k, 1, RowMajor<double>{A, lda}, ColMajor<double>{B, ldb}, epi, st);
    if (transa && !transb) return launch_gemm_f64(m, n, k, 1, ColMajor<double>{A, lda}, RowMajor<double>{B, ldb}, epi, st);
    return launch_gemm_f64(m, n, k, 1, ColMajor<double>{A, lda}, ColMajor<double>{B, ldb}, epi, st);
}

extern "C" size_t vb_fr_logdensity_workspace_bytes(long long n, int dim) {
    if (n < 1 || dim < 1) return 0;
    return (size_t)n * dim * 8 + 256;
}

extern "C" int vb_fr_logdensity(int dtype, const double* d_mu, const double* d_U, double log_pdet, double df,
                                long long n, int dim, const void* d_x, double* d_out_logq,
                                void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_mu || !d_U || !d_x || !d_out_logq || !d_ws || n < 1 || dim < 1 || !(df > 0.0)) return VB_ERR_ARG;
    if (ws_bytes < vb_fr_logdensity_workspace_bytes(n, dim)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    double* Y = (double*)d_ws;
    int rc;
    if (dtype == VB_F64)
        rc = launch_gemm_f64(n, dim, dim, 1, CenteredRowsFr<double>{(const double*)d_x, dim, d_mu},
                             RowMajor<double>{d_U, dim}, PlainStore{Y, dim}, st);
    else if (dtype == VB_F32)
        rc = launch_gemm_f64(n, dim, dim, 1, CenteredRowsFr<float>{(const float*)d_x, dim, d_mu},
                             RowMajor<double>{d_U, dim}, PlainStore{Y, dim}, st);
    else
        return VB_ERR_ARG;
    if (rc != VB_OK) return rc;
    // lgamma((df+d)/2) - lgamma(df/2) - d/2 log(pi df) - 1/2 log pdet      (_distributions.py:33-34)
    const double c_total = isinf(df)
        ? -0.5 * dim * 1.8378770664093453 - 0.5 * log_pdet                       // mvn.logpdf: -d/2 log(2 pi) - 1/2 log pdet
        : lgamma(0.5 * (df + dim)) - lgamma(0.5 * df) - 0.5 * dim * log(3.14159265358979323846 * df) - 0.5 * log_pdet;
    fr_logdensity_rows_kernel<<<grid_for(n * 32, 256, kNumSMs * 8), 256, 0, st>>>(n, dim, Y, c_total, df, d_out_logq);
    VB_LAUNCH_CHECK();
    return VB_OK;
}
