// Streaming reductions behind viabel/bounds.py.
//   divergence_bound / mean_and_check_mc_error (bounds.py:142-192): ONE pass over lw with
//     online max rescaling -> (max, sum w, sum w^2, n, mean lw, M2 lw), w = exp(alpha (lw - max)).
//   wasserstein_bounds / np.cov (bounds.py:130-135, 55-56): column sums, then central
//     2nd/4th moments and the scatter matrix in a second pass.
// HBM-bound: 8 B/weight, 16 d B/sample.  All cross-block merges are in fixed order.
#include <cstdlib>
#include <mutex>
#include "common.cuh"
#include "gemm_f64.cuh"

namespace vb {

static inline int grid_for(long long work_items, int threads, int max_blocks) {
    long long g = (work_items + threads - 1) / threads;
    if (g > max_blocks) g = max_blocks;
    if (g < 1) g = 1;
    return (int)g;
}

constexpr int kDivThreads = 256;

struct DivAcc {
    double M, sw, sw2, cnt, sd, sd2;
    __device__ __forceinline__ void init() { M = -INFINITY; sw = sw2 = cnt = sd = sd2 = 0.0; }
    __device__ __forceinline__ void raise_to(double m, double alpha) {      // rare: a new running maximum
        const double f = rescale_factor(alpha, M, m);
        sw *= f;
        sw2 *= f * f;
        M = m;
    }
    // branch-free accumulation of one element that is known to be <= M
    template <int MIXED = 0>
    __device__ __forceinline__ void add_below(double x, double alpha, double K, const double* table) {
        const double w = exp_nonpos_flush_v<MIXED>(alpha * (x - M), table);
        sw += w;
        sw2 = fma(w, w, sw2);
        const double d = x - K;
        sd += d;
        sd2 = fma(d, d, sd2);
    }
    __device__ __forceinline__ void add(double x, double alpha, double K, const double* table) {
        if (x > M) raise_to(x, alpha);
        add_below(x, alpha, K, table);
        cnt += 1.0;
    }
    __device__ __forceinline__ void rescale_to(double Mn, double alpha) {
        const double f = rescale_factor(alpha, M, Mn);
        sw *= f;
        sw2 *= f * f;
        M = Mn;
    }
};

template <int MIXED>
__global__ void __launch_bounds__(kDivThreads)
divergence_blocks_kernel(const double* __restrict__ lw, long long n, double alpha,
                         double* __restrict__ block_out /* [grid][6] */) {
    __shared__ double red[kDivThreads / 32][6];
    __shared__ double s_table[64];
    if (threadIdx.x < 64) s_table[threadIdx.x] = kExp2Table[threadIdx.x];
    __syncthreads();
    const double K = lw[0];
    DivAcc a, b;                        // two independent accumulator sets (same running max)
    a.init();
    b.init();
    const long long gtid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long gthreads = (long long)gridDim.x * blockDim.x;
    const bool aligned = (reinterpret_cast<uintptr_t>(lw) & 15u) == 0;
    if (aligned) {
        const long long n2 = n >> 1;
        const double2* p = reinterpret_cast<const double2*>(lw);
        long long i = gtid;
        // 4 independent 16-byte loads in flight per thread; one (rare) branch per 8 elements
        for (; i + 3 * gthreads < n2; i += 4 * gthreads) {
            const double2 v0 = ld_stream2(p + i), v1 = ld_stream2(p + i + gthreads);
            const double2 v2 = ld_stream2(p + i + 2 * gthreads), v3 = ld_stream2(p + i + 3 * gthreads);
            const double m8 = fmax(fmax(fmax(v0.x, v0.y), fmax(v1.x, v1.y)), fmax(fmax(v2.x, v2.y), fmax(v3.x, v3.y)));
            if (m8 > a.M) { a.raise_to(m8, alpha); b.raise_to(m8, alpha); }
            a.add_below<MIXED>(v0.x, alpha, K, s_table); b.add_below<MIXED>(v0.y, alpha, K, s_table);
            a.add_below<MIXED>(v1.x, alpha, K, s_table); b.add_below<MIXED>(v1.y, alpha, K, s_table);
            a.add_below<MIXED>(v2.x, alpha, K, s_table); b.add_below<MIXED>(v2.y, alpha, K, s_table);
            a.add_below<MIXED>(v3.x, alpha, K, s_table); b.add_below<MIXED>(v3.y, alpha, K, s_table);
            a.cnt += 8.0;
        }
        for (; i < n2; i += gthreads) {
            const double2 v = ld_stream2(p + i);
            const double m2 = fmax(v.x, v.y);
            if (m2 > a.M) { a.raise_to(m2, alpha); b.raise_to(m2, alpha); }
            a.add_below<MIXED>(v.x, alpha, K, s_table); b.add_below<MIXED>(v.y, alpha, K, s_table);
            a.cnt += 2.0;
        }
        if ((n & 1) && gtid == 0) {
            const double x = lw[n - 1];
            if (x > a.M) { a.raise_to(x, alpha); b.raise_to(x, alpha); }
            a.add_below<MIXED>(x, alpha, K, s_table);
            a.cnt += 1.0;
        }
    } else {
        for (long long i = gtid; i < n; i += gthreads) {
            const double x = lw[i];
            if (x > a.M) { a.raise_to(x, alpha); b.raise_to(x, alpha); }
            a.add_below(x, alpha, K, s_table);
            a.cnt += 1.0;
        }
    }
    a.sw += b.sw; a.sw2 += b.sw2; a.sd += b.sd; a.sd2 += b.sd2;
    // warp, then block merge
    const double Mw = warp_max(a.M);
    a.rescale_to(Mw, alpha);
    a.sw = warp_sum(a.sw); a.sw2 = warp_sum(a.sw2); a.cnt = warp_sum(a.cnt);
    a.sd = warp_sum(a.sd); a.sd2 = warp_sum(a.sd2);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        red[warp][0] = a.M; red[warp][1] = a.sw; red[warp][2] = a.sw2;
        red[warp][3] = a.cnt; red[warp][4] = a.sd; red[warp][5] = a.sd2;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double M = -INFINITY;
        for (int w = 0; w < kDivThreads / 32; ++w) M = fmax(M, red[w][0]);
        double o[6] = {M, 0, 0, 0, 0, 0};
        for (int w = 0; w < kDivThreads / 32; ++w) {
            const double f = rescale_factor(alpha, red[w][0], M);
            o[1] += red[w][1] * f; o[2] += red[w][2] * f * f;
            o[3] += red[w][3]; o[4] += red[w][4]; o[5] += red[w][5];
        }
        for (int k = 0; k < 6; ++k) block_out[(size_t)blockIdx.x * 6 + k] = o[k];
    }
}

// partial layout shared with the multi-rank merge: [max, sum w, sum w^2, count, mean lw, M2 lw]
__device__ void divergence_finalize(const double* parts, int R, double alpha, double* out) {
    double M = -INFINITY;
    for (int r = 0; r < R; ++r) M = fmax(M, parts[r * 6]);
    double sw = 0, sw2 = 0, cnt = 0, mean = 0, M2 = 0;
    for (int r = 0; r < R; ++r) {
        const double* p = parts + r * 6;
        const double f = rescale_factor(alpha, p[0], M);
        sw += p[1] * f;
        sw2 += p[2] * f * f;
        // Chan et al. pairwise merge of (count, mean, M2)
        const double nb = p[3];
        if (nb > 0) {
            const double tot = cnt + nb, delta = p[4] - mean;
            mean += delta * nb / tot;
            M2 += p[5] + delta * delta * cnt * nb / tot;
            cnt = tot;
        }
    }
    const double mw = sw / cnt;
    double var_w = sw2 / cnt - mw * mw;
    if (var_w < 0) var_w = 0;
    double var_lw = M2 / cnt;
    if (var_lw < 0) var_lw = 0;
    out[0] = M;
    out[1] = mw;
    out[2] = sqrt(var_w) / sqrt(cnt);
    out[3] = mean;
    out[4] = sqrt(var_lw) / sqrt(cnt);
    out[5] = log(mw) / alpha + M;     // CUBO, bounds.py:173
    out[6] = cnt;
    out[7] = 0.0;
}

__global__ void __launch_bounds__(256)
divergence_finish_kernel(const double* __restrict__ lw, const double* __restrict__ blocks, int G,
                         double alpha, double* out_partials, double* out) {
    __shared__ double scratch[32];
    __shared__ double p[6];
    const double K = lw[0];
    double M = -INFINITY;
    for (int b = threadIdx.x; b < G; b += blockDim.x) M = fmax(M, blocks[(size_t)b * 6]);
    M = block_max(M, scratch);
    double acc[5] = {0, 0, 0, 0, 0};
    for (int b = threadIdx.x; b < G; b += blockDim.x) {          // fixed thread->block map, fixed tree below
        const double* q = blocks + (size_t)b * 6;
        const double f = rescale_factor(alpha, q[0], M);
        acc[0] += q[1] * f; acc[1] += q[2] * f * f; acc[2] += q[3]; acc[3] += q[4]; acc[4] += q[5];
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) acc[k] = block_sum(acc[k], scratch);
    if (threadIdx.x != 0) return;
    const double sd = acc[3], sd2 = acc[4];
    p[0] = M; p[1] = acc[0]; p[2] = acc[1]; p[3] = acc[2];
    p[4] = K + sd / p[3];
    p[5] = sd2 - sd * sd / p[3];
    if (out_partials) for (int k = 0; k < 6; ++k) out_partials[k] = p[k];
    if (out) divergence_finalize(p, 1, alpha, out);
}

__global__ void divergence_merge_kernel(const double* __restrict__ parts, int R, double alpha, double* out) {
    if (threadIdx.x == 0) divergence_finalize(parts, R, alpha, out);
}

// ------------------------------------------------------------------------------------------ max
__global__ void __launch_bounds__(256)
max_blocks_kernel(const double* __restrict__ x, long long n, double* __restrict__ block_out) {
    __shared__ double scratch[32];
    double m = -INFINITY;
    const long long gtid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long gthreads = (long long)gridDim.x * blockDim.x;
    if ((reinterpret_cast<uintptr_t>(x) & 15u) == 0) {
        const long long n2 = n >> 1;
        const double2* p = reinterpret_cast<const double2*>(x);
        long long i = gtid;
        for (; i + 3 * gthreads < n2; i += 4 * gthreads) {
            const double2 v0 = ld_stream2(p + i), v1 = ld_stream2(p + i + gthreads);
            const double2 v2 = ld_stream2(p + i + 2 * gthreads), v3 = ld_stream2(p + i + 3 * gthreads);
            m = fmax(m, fmax(fmax(fmax(v0.x, v0.y), fmax(v1.x, v1.y)), fmax(fmax(v2.x, v2.y), fmax(v3.x, v3.y))));
        }
        for (; i < n2; i += gthreads) { const double2 v = ld_stream2(p + i); m = fmax(m, fmax(v.x, v.y)); }
        if ((n & 1) && gtid == 0) m = fmax(m, x[n - 1]);
    } else {
        for (long long i = gtid; i < n; i += gthreads) m = fmax(m, x[i]);
    }
    m = block_max(m, scratch);
    if (threadIdx.x == 0) block_out[blockIdx.x] = m;
}
__global__ void max_finish_kernel2(const double* __restrict__ vals, int count, double* __restrict__ out) {
    __shared__ double scratch[32];
    double m = -INFINITY;
    for (int i = threadIdx.x; i < count; i += blockDim.x) m = fmax(m, vals[i]);
    m = block_max(m, scratch);
    if (threadIdx.x == 0) *out = m;
}

// ------------------------------------------------------------------------------------------ column statistics
// Thread g always touches column g % D (stride is a multiple of D), so it keeps scalar
// accumulators; per-thread partials are merged per column in a fixed tree.
//   mode 0: sum x            mode 1: sum c^2 and sum c^4, c = x - mean[col]
template <typename T>
__global__ void colstat_kernel(const T* __restrict__ x, long long total, int D, const double* __restrict__ mean,
                               long long active_threads, double* __restrict__ out /* [active][2] */) {
    const long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (g >= active_threads) return;
    const int col = (int)(g % D);
    double a0 = 0.0, a1 = 0.0;
    if (mean == nullptr) {
        for (long long e = g; e < total; e += active_threads) a0 += (double)x[e];
    } else {
        const double mu = mean[col];
        for (long long e = g; e < total; e += active_threads) {
            const double c = (double)x[e] - mu, c2 = c * c;
            a0 += c2;
            a1 = fma(c2, c2, a1);
        }
    }
    out[2 * g] = a0;
    out[2 * g + 1] = a1;
}
// one block per column: out_cols[col][2]
__global__ void colstat_finish_kernel(const double* __restrict__ parts, long long active_threads, int D,
                                      double* __restrict__ out_cols) {
    __shared__ double scratch[32];
    const int col = blockIdx.x;
    double a0 = 0.0, a1 = 0.0;
    for (long long g = col + (long long)threadIdx.x * D; g < active_threads; g += (long long)blockDim.x * D) {
        a0 += parts[2 * g];
        a1 += parts[2 * g + 1];
    }
    a0 = block_sum(a0, scratch);
    a1 = block_sum(a1, scratch);
    if (threadIdx.x == 0) { out_cols[2 * col] = a0; out_cols[2 * col + 1] = a1; }
}
__global__ void colsum_write_kernel(const double* __restrict__ cols, int D, double* __restrict__ out_sum) {
    for (int j = threadIdx.x; j < D; j += blockDim.x) out_sum[j] = cols[2 * j];
}
__global__ void central_write_kernel(const double* __restrict__ cols, int D, double* __restrict__ out) {
    if (threadIdx.x == 0) {
        double s2 = 0, s4 = 0;
        for (int j = 0; j < D; ++j) { s2 += cols[2 * j]; s4 += cols[2 * j + 1]; }
        out[0] = s2; out[1] = s4;
    }
}

// scatter matrix: S = sum_i c_i c_i' through the fp64 GEMM with centering loaders and split-K
template <typename T>
struct CenteredT {      // A(a, i) = x[i, a] - mean[a]      (first index = column, second = sample)
    static constexpr bool kContigCol = false;
    const T* x; long long ld; const double* m;
    __device__ __forceinline__ double operator()(long long a, long long i) const { return (double)x[i * ld + a] - m[a]; }
};
template <typename T>
struct CenteredN {      // B(i, b) = x[i, b] - mean[b]
    static constexpr bool kContigCol = true;
    const T* x; long long ld; const double* m;
    __device__ __forceinline__ double operator()(long long i, long long b) const { return (double)x[i * ld + b] - m[b]; }
};
struct SplitStore {
    double* p; long long mn; long long ld;
    __device__ __forceinline__ void operator()(long long i, long long j, double acc, int split) const {
        p[(size_t)split * mn + i * ld + j] = acc;
    }
};
__global__ void split_sum_kernel(const double* __restrict__ parts, int splits, long long mn, double* __restrict__ out) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < mn; e += (long long)gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < splits; ++k) s += parts[(size_t)k * mn + e];
        out[e] = s;
    }
}

static long long colstat_threads(long long total, int D) {
    long long t = (long long)kNumSMs * 8 * 256;
    if (t > total) t = total;
    t = t / D * D;
    if (t < D) t = D;
    return t;
}
static int scatter_splits(long long n, int D) {
    const long long tiles = ((D + kGBM - 1) / kGBM) * (long long)((D + kGBN - 1) / kGBN);
    long long s = (kNumSMs * 2 + tiles - 1) / tiles;
    const long long max_s = (n + 255) / 256;
    if (s > max_s) s = max_s;
    if (s < 1) s = 1;
    return (int)s;
}

}  // namespace vb

using namespace vb;

extern "C" size_t vb_bounds_divergence_workspace_bytes(long long n) {
    (void)n;
    return (size_t)kNumSMs * 8 * 6 * sizeof(double) + 64;
}

static int run_divergence(const double* lw, long long n, double alpha, double* out_partials, double* out,
                          void* ws, size_t ws_bytes, cudaStream_t st) {
    if (!lw || n < 1 || !ws || !(alpha > 0.0)) return VB_ERR_ARG;
    if (ws_bytes < vb_bounds_divergence_workspace_bytes(n)) return VB_ERR_WORKSPACE;
    const int grid = grid_for((n + 1) / 2, kDivThreads, kNumSMs * 8);
    // VB_EXP_MIXED=1 (experiment): series tail of the exp in fp32 (7 instead of 12 fp64 operations per weight)
    static const bool mixed = getenv("VB_EXP_MIXED") != nullptr && atoi(getenv("VB_EXP_MIXED")) != 0;
    if (mixed) divergence_blocks_kernel<1><<<grid, kDivThreads, 0, st>>>(lw, n, alpha, (double*)ws);
    else divergence_blocks_kernel<0><<<grid, kDivThreads, 0, st>>>(lw, n, alpha, (double*)ws);
    VB_LAUNCH_CHECK();
    divergence_finish_kernel<<<1, 256, 0, st>>>(lw, (const double*)ws, grid, alpha, out_partials, out);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" int vb_bounds_divergence(const double* d_lw, long long n, double alpha, double* d_out,
                                    void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_out) return VB_ERR_ARG;
    return run_divergence(d_lw, n, alpha, nullptr, d_out, d_ws, ws_bytes, (cudaStream_t)stream);
}

extern "C" int vb_bounds_divergence_partials(const double* d_lw, long long n, double alpha,
                                             double* d_out_partials, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_out_partials) return VB_ERR_ARG;
    return run_divergence(d_lw, n, alpha, d_out_partials, nullptr, d_ws, ws_bytes, (cudaStream_t)stream);
}

extern "C" int vb_bounds_divergence_merge(const double* d_partials, int n_ranks, double alpha,
                                          double* d_out, void* stream) {
    if (!d_partials || !d_out || n_ranks < 1) return VB_ERR_ARG;
    divergence_merge_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(d_partials, n_ranks, alpha, d_out);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ------------------------------------------------------------------------------------------ no-allocation variant
// vb_bounds_divergence with its scratch (the per-CTA partials) in a file-scope device array instead of a caller
// workspace: for callers that cannot allocate (CUDA graph capture, serving loops).  The array is part of the
// library image: every GPU that loads the library holds its 56 832 bytes until the process exits.  In host code
// the symbol below is only a stand-in; its device address is resolved once per device with cudaGetSymbolAddress.
constexpr int kDivStaticBlocks = kNumSMs * 8;                    // the grid cap of run_divergence
static __device__ __align__(256) double g_div_static_scratch[kDivStaticBlocks * 6];
constexpr int kDivStaticMaxDevices = 64;
static double* g_div_static_ptr[kDivStaticMaxDevices];
static std::mutex g_div_static_mutex;

static int div_static_scratch(double** out) {
    int dev = -1;
    VB_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kDivStaticMaxDevices) return VB_ERR_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_div_static_mutex);
    if (!g_div_static_ptr[dev]) {
        void* p = nullptr;
        VB_CUDA_TRY(cudaGetSymbolAddress(&p, g_div_static_scratch));      // device `dev` is current
        if (!p) return VB_ERR_UNSUPPORTED;
        g_div_static_ptr[dev] = (double*)p;
    }
    *out = g_div_static_ptr[dev];
    return VB_OK;
}

extern "C" long long vb_bounds_divergence_static_max_n(void) { return VB_BOUNDS_DIVERGENCE_STATIC_MAX_N; }

extern "C" int vb_bounds_divergence_static(const double* d_lw, long long n, double alpha, double* d_out, void* stream) {
    if (!d_lw || !d_out || n < 1 || !(alpha > 0.0)) return VB_ERR_ARG;
    if (n > VB_BOUNDS_DIVERGENCE_STATIC_MAX_N) return VB_ERR_TOO_LARGE;
    double* scratch = nullptr;
    const int rc = div_static_scratch(&scratch);
    if (rc != VB_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = grid_for((n + 1) / 2, kDivThreads, kDivStaticBlocks);
    // nothing to clear: the first kernel overwrites all six doubles of each of its `grid` slots and the second
    // reads exactly those slots; what an earlier call left beyond them is never read
    divergence_blocks_kernel<0><<<grid, kDivThreads, 0, st>>>(d_lw, n, alpha, scratch);
    VB_LAUNCH_CHECK();
    divergence_finish_kernel<<<1, 256, 0, st>>>(d_lw, scratch, grid, alpha, nullptr, d_out);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" size_t vb_max_workspace_bytes(long long n) {
    (void)n;
    return (size_t)kNumSMs * 8 * sizeof(double) + 64;
}

extern "C" int vb_max(const double* d_x, long long n, double* d_out, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_x || !d_out || !d_ws || n < 1) return VB_ERR_ARG;
    if (ws_bytes < vb_max_workspace_bytes(n)) return VB_ERR_WORKSPACE;
    const int grid = grid_for((n + 1) / 2, 256, kNumSMs * 8);
    max_blocks_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_x, n, (double*)d_ws);
    VB_LAUNCH_CHECK();
    max_finish_kernel2<<<1, 256, 0, (cudaStream_t)stream>>>((const double*)d_ws, grid, d_out);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" size_t vb_bounds_moments_workspace_bytes(long long n, int dim) {
    if (n < 1 || dim < 1) return 0;
    const long long t = colstat_threads(n * dim, dim);
    size_t bytes = (size_t)t * 16 + (size_t)dim * 16 + 256;
    const size_t sc = (size_t)scatter_splits(n, dim) * dim * dim * 8 + 256;
    return bytes > sc ? bytes : sc;
}

template <typename T>
static int run_colstat(const T* x, long long n, int D, const double* mean, double* cols_out, char* ws,
                       cudaStream_t st) {
    const long long total = n * D;
    const long long t = colstat_threads(total, D);
    double* parts = (double*)ws;
    const int grid = (int)((t + 255) / 256);
    colstat_kernel<T><<<grid, 256, 0, st>>>(x, total, D, mean, t, parts);
    VB_LAUNCH_CHECK();
    colstat_finish_kernel<<<D, 256, 0, st>>>(parts, t, D, cols_out);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" int vb_bounds_colsum(int dtype, const void* d_x, long long n, int dim, double* d_out_sum,
                                void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_x || !d_out_sum || !d_ws || n < 1 || dim < 1) return VB_ERR_ARG;
    if (ws_bytes < vb_bounds_moments_workspace_bytes(n, dim)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)d_ws;
    const long long t = colstat_threads(n * dim, dim);
    double* cols = (double*)(ws + (size_t)t * 16);
    int rc;
    if (dtype == VB_F64) rc = run_colstat<double>((const double*)d_x, n, dim, nullptr, cols, ws, st);
    else if (dtype == VB_F32) rc = run_colstat<float>((const float*)d_x, n, dim, nullptr, cols, ws, st);
    else return VB_ERR_ARG;
    if (rc != VB_OK) return rc;
    colsum_write_kernel<<<1, 256, 0, st>>>(cols, dim, d_out_sum);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" int vb_bounds_central(int dtype, const void* d_x, long long n, int dim, const double* d_mean,
                                 double* d_out, double* d_out_scatter_or_null,
                                 void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_x || !d_mean || !d_out || !d_ws || n < 1 || dim < 1) return VB_ERR_ARG;
    if (dtype != VB_F32 && dtype != VB_F64) return VB_ERR_ARG;
    if (ws_bytes < vb_bounds_moments_workspace_bytes(n, dim)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)d_ws;
    const long long t = colstat_threads(n * dim, dim);
    double* cols = (double*)(ws + (size_t)t * 16);
    int rc;
    if (dtype == VB_F64) rc = run_colstat<double>((const double*)d_x, n, dim, d_mean, cols, ws, st);
    else rc = run_colstat<float>((const float*)d_x, n, dim, d_mean, cols, ws, st);
    if (rc != VB_OK) return rc;
    central_write_kernel<<<1, 32, 0, st>>>(cols, dim, d_out);
    VB_LAUNCH_CHECK();
    if (d_out_scatter_or_null) {
        const int splits = scatter_splits(n, dim);
        double* parts = (double*)ws;
        const long long mn = (long long)dim * dim;
        if (dtype == VB_F64)
            rc = launch_gemm_f64(dim, dim, n, splits, CenteredT<double>{(const double*)d_x, dim, d_mean},
                                 CenteredN<double>{(const double*)d_x, dim, d_mean}, SplitStore{parts, mn, dim}, st);
        else
            rc = launch_gemm_f64(dim, dim, n, splits, CenteredT<float>{(const float*)d_x, dim, d_mean},
                                 CenteredN<float>{(const float*)d_x, dim, d_mean}, SplitStore{parts, mn, dim}, st);
        if (rc != VB_OK) return rc;
        split_sum_kernel<<<grid_for(mn, 256, kNumSMs * 4), 256, 0, st>>>(parts, splits, mn, d_out_scatter_or_null);
        VB_LAUNCH_CHECK();
    }
    return VB_OK;
}
