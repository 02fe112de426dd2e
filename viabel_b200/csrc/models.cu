// Model plugin handles and the standalone log p / grad log p op (plugin contract of
// viabel/vb.py:314-321: x (n, d) -> log p (n), grad (n, d) on the unconstrained space).
#include <new>
#include "common.cuh"
#include "models.cuh"
#include "gemm_f64.cuh"
#include "dense.cuh"

namespace vb {

static inline int grid_for(long long work_items, int threads, int max_blocks) {
    long long g = (work_items + threads - 1) / threads;
    if (g > max_blocks) g = max_blocks;
    if (g < 1) g = 1;
    return (int)g;
}

// ------------------------------------------------------------------------------------------ elementwise plugins
template <typename T, class Model>
__global__ void model_rows_kernel(const __grid_constant__ typename Model::P mp, long long n,
                                  const T* __restrict__ x, double* __restrict__ logp, T* __restrict__ grad) {
    constexpr int D = Model::D;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        T xv[D], g[D], lp;
#pragma unroll
        for (int j = 0; j < D; ++j) xv[j] = x[i * D + j];
        Model::template eval<T>(mp, xv, lp, g);
        logp[i] = (double)lp;
        if (grad) {
#pragma unroll
            for (int j = 0; j < D; ++j) grad[i * D + j] = g[j];
        }
    }
}

template <class Model>
static int run_elementwise(const typename Model::P& mp, int dtype, const void* x, long long n, double* logp,
                           void* grad, cudaStream_t st) {
    const int grid = grid_for(n, 256, kNumSMs * 8);
    if (dtype == VB_F64)
        model_rows_kernel<double, Model><<<grid, 256, 0, st>>>(mp, n, (const double*)x, logp, (double*)grad);
    else if (dtype == VB_F32)
        model_rows_kernel<float, Model><<<grid, 256, 0, st>>>(mp, n, (const float*)x, logp, (float*)grad);
    else
        return VB_ERR_ARG;
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ------------------------------------------------------------------------------------------ correlated Gaussian
// G = -(X - m) P ;  log p_i = 1/2 sum_j G_ij (X_ij - m_j)
template <typename T>
struct CenteredRows {                     // A(i, k) = X[i, k] - m[k]
    static constexpr bool kContigCol = true;
    const T* x; long long ld; const double* m;
    __device__ __forceinline__ double operator()(long long i, long long k) const {
        return (double)x[i * ld + k] - m[k];
    }
};
template <typename T>
struct StoreNeg {                         // C(i, j) = -acc
    T* c; long long ld;
    __device__ __forceinline__ void operator()(long long i, long long j, double acc, int) const {
        c[i * ld + j] = (T)(-acc);
    }
};
template <typename T>
__global__ void gauss_logp_kernel(long long n, int D, const T* __restrict__ x, const double* __restrict__ m,
                                  const T* __restrict__ g, double* __restrict__ logp) {
    const int lane = threadIdx.x & 31;
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = warp; i < n; i += nwarps) {
        double s = 0.0;
        for (int j = lane; j < D; j += 32) s += (double)g[i * D + j] * ((double)x[i * D + j] - m[j]);
        s = warp_sum(s);
        if (lane == 0) logp[i] = 0.5 * s;
    }
}

// ------------------------------------------------------------------------------------------ Student-t regression (fp64 path)
// r = y - X b ;  log p = -|b|^2 / (2 sd^2) - (df+1)/2 sum_r log1p(r^2/df) ;
// grad = -b / sd^2 + psi' X ,  psi = (df+1) r / (df + r^2).      
template <typename TB, typename TX>
struct ResidualStore {                    // R(s, r) = y_r - acc
    double* R; long long ld; const TX* y; long long r0;
    __device__ __forceinline__ void operator()(long long s, long long r, double acc, int) const {
        R[s * ld + r] = (double)y[r0 + r] - acc;
    }
};
struct PsiLoad {                          // A(s, r) = (df+1) R / (df + R^2)
    static constexpr bool kContigCol = true;
    const double* R; long long ld; double df;
    __device__ __forceinline__ double operator()(long long s, long long r) const {
        const double v = R[s * ld + r];
        return (df + 1.0) * v / (df + v * v);
    }
};
template <typename T>
struct AccumStore {                       // C(i, j) += acc  (single owner per element: deterministic)
    T* c; long long ld;
    __device__ __forceinline__ void operator()(long long i, long long j, double acc, int) const {
        c[i * ld + j] = (T)((double)c[i * ld + j] + acc);
    }
};
template <typename T>
__global__ void regression_init_kernel(long long n, int D, const T* __restrict__ b, double inv_var,
                                       double* __restrict__ logp, double* __restrict__ grad_acc) {
    const int lane = threadIdx.x & 31;
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = warp; i < n; i += nwarps) {
        double s = 0.0;
        for (int j = lane; j < D; j += 32) {
            const double v = (double)b[i * D + j];
            s += v * v;
            grad_acc[i * D + j] = -v * inv_var;
        }
        s = warp_sum(s);
        if (lane == 0) logp[i] = -0.5 * s * inv_var;
    }
}
__global__ void regression_rowsum_kernel(long long n, long long cols, const double* __restrict__ R, long long ld,
                                         double df, double* __restrict__ logp) {
    const int lane = threadIdx.x & 31;
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = warp; i < n; i += nwarps) {
        double s = 0.0;
        for (long long c = lane; c < cols; c += 32) {
            const double v = R[i * ld + c];
            s += log1p(v * v / df);
        }
        s = warp_sum(s);
        if (lane == 0) logp[i] -= 0.5 * (df + 1.0) * s;
    }
}
template <typename T>
__global__ void cast_out_kernel(long long total, const double* __restrict__ src, T* __restrict__ dst) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x)
        dst[e] = (T)src[e];
}

static long long regression_chunk_rows(long long n, long long n_rows) {
    // residual scratch of at most ~1 GiB, at least one GEMM tile wide
    long long c = (1ll << 27) / (n > 0 ? n : 1);
    c = c / kGBN * kGBN;
    if (c < kGBN) c = kGBN;
    if (c > n_rows) c = n_rows;
    return c;
}

template <typename TB, typename TX>
static int run_regression_f64(const ModelHost* m, const TB* b, long long n, double* logp, TB* grad_out,
                              char* ws, cudaStream_t st) {
    const int D = m->dim;
    const long long N = m->n_rows;
    const long long chunk = regression_chunk_rows(n, N);
    double* grad_acc = (double*)ws;
    double* R = (double*)(ws + align_up((size_t)n * D * 8, 256));
    const TX* X = (const TX*)m->d_X;
    const TX* y = (const TX*)m->d_y;
    const double inv_var = 1.0 / (m->prior_sd * m->prior_sd);
    regression_init_kernel<TB><<<grid_for(n * 32, 256, kNumSMs * 8), 256, 0, st>>>(n, D, b, inv_var, logp, grad_acc);
    VB_LAUNCH_CHECK();
    for (long long r0 = 0; r0 < N; r0 += chunk) {
        const long long cols = (r0 + chunk <= N) ? chunk : (N - r0);
        // R (n x cols) = y - B X_chunk'
        int rc = launch_gemm_f64(n, cols, D, 1, RowMajor<TB>{b, D}, ColMajor<TX>{X + r0 * D, D},
                                 ResidualStore<TB, TX>{R, cols, y, r0}, st);
        if (rc != VB_OK) return rc;
        regression_rowsum_kernel<<<grid_for(n * 32, 256, kNumSMs * 8), 256, 0, st>>>(n, cols, R, cols, m->df, logp);
        VB_LAUNCH_CHECK();
        if (grad_out) {
            // grad_acc (n x D) += psi(R) X_chunk
            rc = launch_gemm_f64(n, D, cols, 1, PsiLoad{R, cols, m->df}, RowMajor<TX>{X + r0 * D, D},
                                 AccumStore<double>{grad_acc, D}, st);
            if (rc != VB_OK) return rc;
        }
    }
    if (grad_out) {
        cast_out_kernel<TB><<<grid_for(n * D, 256, kNumSMs * 8), 256, 0, st>>>(n * D, grad_acc, grad_out);
        VB_LAUNCH_CHECK();
    }
    return VB_OK;
}

}  // namespace vb

using namespace vb;

static int new_model(int kind, int dim, vb_model** out, ModelHost** mh) {
    if (!out) return VB_ERR_ARG;
    ModelHost* m = new (std::nothrow) ModelHost();
    if (!m) return VB_ERR_ARG;
    m->kind = kind;
    m->dim = dim;
    *mh = m;
    *out = reinterpret_cast<vb_model*>(m);
    return VB_OK;
}

extern "C" int vb_model_create_funnel(double log_sigma_stdev, vb_model** out) {
    if (!(log_sigma_stdev > 0.0)) return VB_ERR_ARG;
    ModelHost* m;
    int rc = new_model(VB_MODEL_FUNNEL, 2, out, &m);
    if (rc) return rc;
    m->p0 = log_sigma_stdev;
    return VB_OK;
}

extern "C" int vb_model_create_mixture(double loc, vb_model** out) {
    ModelHost* m;
    int rc = new_model(VB_MODEL_MIXTURE, 1, out, &m);
    if (rc) return rc;
    m->p0 = loc;
    return VB_OK;
}

extern "C" int vb_model_create_eight_schools(const double* h_y, const double* h_sigma, int n_schools,
                                             vb_model** out) {
    if (!h_y || !h_sigma || n_schools < 1) return VB_ERR_ARG;
    if (n_schools != kMaxSchools) return VB_ERR_UNSUPPORTED;
    for (int j = 0; j < n_schools; ++j)
        if (!(h_sigma[j] > 0.0)) return VB_ERR_ARG;
    ModelHost* m;
    int rc = new_model(VB_MODEL_EIGHT_SCHOOLS, n_schools + 2, out, &m);
    if (rc) return rc;
    m->n_schools = n_schools;
    for (int j = 0; j < n_schools; ++j) { m->y[j] = h_y[j]; m->sigma[j] = h_sigma[j]; }
    return VB_OK;
}

extern "C" int vb_model_create_regression(const void* d_X, const void* d_y, long long n_rows, int dim,
                                          double df, double prior_sd, int dtype, vb_model** out) {
    if (!d_X || !d_y || n_rows < 1 || dim < 1 || !(df > 0.0) || !(prior_sd > 0.0)) return VB_ERR_ARG;
    if (dtype != VB_F32 && dtype != VB_F64) return VB_ERR_ARG;
    ModelHost* m;
    int rc = new_model(VB_MODEL_REGRESSION, dim, out, &m);
    if (rc) return rc;
    m->d_X = d_X; m->d_y = d_y; m->n_rows = n_rows; m->df = df; m->prior_sd = prior_sd; m->data_dtype = dtype;
    m->d_X_planes = nullptr;
    m->d_X_planes16 = nullptr;
    return VB_OK;
}

extern "C" int vb_model_regression_attach_split_f16(vb_model* model, const void* d_X_planes16) {
    if (!model) return VB_ERR_ARG;
    ModelHost* m = reinterpret_cast<ModelHost*>(model);
    if (m->kind != VB_MODEL_REGRESSION || m->data_dtype != VB_F32) return VB_ERR_UNSUPPORTED;
    if (reinterpret_cast<uintptr_t>(d_X_planes16) & 255u) return VB_ERR_ALIGN;
    m->d_X_planes16 = d_X_planes16;
    return VB_OK;
}

extern "C" int vb_model_regression_attach_split(vb_model* model, const float* d_X_planes) {
    if (!model) return VB_ERR_ARG;
    ModelHost* m = reinterpret_cast<ModelHost*>(model);
    if (m->kind != VB_MODEL_REGRESSION || m->data_dtype != VB_F32) return VB_ERR_UNSUPPORTED;
    m->d_X_planes = d_X_planes;
    return VB_OK;
}

extern "C" int vb_model_create_gauss(const double* d_mean, const double* d_precision, int dim, vb_model** out) {
    if (!d_mean || !d_precision || dim < 1) return VB_ERR_ARG;
    ModelHost* m;
    int rc = new_model(VB_MODEL_GAUSS, dim, out, &m);
    if (rc) return rc;
    m->d_mean = d_mean; m->d_prec = d_precision;
    m->d_gauss_cache = nullptr;
    return VB_OK;
}
extern "C" size_t vb_model_gauss_cache_bytes(int dim) { return dim < 1 ? 0 : gauss_cache_bytes(dim); }
extern "C" int vb_model_gauss_attach_cache(vb_model* model, float* d_cache, void* stream) {
    if (!model || !d_cache) return VB_ERR_ARG;
    ModelHost* m = reinterpret_cast<ModelHost*>(model);
    if (m->kind != VB_MODEL_GAUSS) return VB_ERR_UNSUPPORTED;
    if (reinterpret_cast<uintptr_t>(d_cache) & 255u) return VB_ERR_ALIGN;
    const int rc = gauss_fill_cache(m->d_mean, m->d_prec, m->dim, d_cache, (cudaStream_t)stream);
    if (rc != VB_OK) return rc;
    m->d_gauss_cache = d_cache;
    return VB_OK;
}

extern "C" int vb_model_kind(const vb_model* m) { return m ? reinterpret_cast<const ModelHost*>(m)->kind : VB_ERR_ARG; }
extern "C" int vb_model_dim(const vb_model* m) { return m ? reinterpret_cast<const ModelHost*>(m)->dim : VB_ERR_ARG; }
extern "C" void vb_model_destroy(vb_model* m) { delete reinterpret_cast<ModelHost*>(m); }

extern "C" size_t vb_model_logp_grad_workspace_bytes(const vb_model* model, long long n, int dtype) {
    if (!model || n < 1) return 0;
    const ModelHost* m = reinterpret_cast<const ModelHost*>(model);
    if (m->kind == VB_MODEL_REGRESSION) {
        if (dtype == VB_F32 && m->data_dtype == VB_F32 && regression_fused_ok(m))
            return regression_fused_workspace_bytes(n, m->n_rows);
        if (dtype == VB_F32 && m->data_dtype == VB_F32)
            return regression_tf32_workspace_bytes(n, m->n_rows, m->dim, m->d_X_planes != nullptr);
        const long long chunk = regression_chunk_rows(n, m->n_rows);
        return align_up((size_t)n * m->dim * 8, 256) + (size_t)n * chunk * 8 + 256;
    }
    if (m->kind == VB_MODEL_GAUSS)
        return dtype == VB_F32 ? gauss_f32_workspace_bytes(n, m->dim) : (size_t)n * m->dim * 8 + 256;    // G when the caller does not want it
    return 256;
}

extern "C" int vb_model_logp_grad(const vb_model* model, int dtype, const void* d_x, long long n,
                                  double* d_out_logp, void* d_out_grad_or_null,
                                  void* d_ws, size_t ws_bytes, void* stream) {
    if (!model || !d_x || !d_out_logp || n < 1) return VB_ERR_ARG;
    if (dtype != VB_F32 && dtype != VB_F64) return VB_ERR_ARG;
    const ModelHost* m = reinterpret_cast<const ModelHost*>(model);
    cudaStream_t st = (cudaStream_t)stream;
    switch (m->kind) {
        case VB_MODEL_FUNNEL: {
            const Funnel::P p = Funnel::make(m->p0);
            return run_elementwise<Funnel>(p, dtype, d_x, n, d_out_logp, d_out_grad_or_null, st);
        }
        case VB_MODEL_MIXTURE: {
            Mixture::P p{m->p0};
            return run_elementwise<Mixture>(p, dtype, d_x, n, d_out_logp, d_out_grad_or_null, st);
        }
        case VB_MODEL_EIGHT_SCHOOLS: {
            EightSchools<8>::P p;
            for (int j = 0; j < 8; ++j) { p.y[j] = m->y[j]; p.inv_sigma[j] = 1.0 / m->sigma[j]; }
            return run_elementwise<EightSchools<8>>(p, dtype, d_x, n, d_out_logp, d_out_grad_or_null, st);
        }
        case VB_MODEL_GAUSS: {
            const int D = m->dim;
            // G is needed for log p even when the caller does not want it: use the workspace then.
            if (dtype == VB_F64) {
                double* G = (double*)d_out_grad_or_null;
                if (!G) {
                    if (!d_ws || ws_bytes < (size_t)n * D * 8) return VB_ERR_WORKSPACE;
                    G = (double*)d_ws;
                }
                int rc = launch_gemm_f64(n, D, D, 1, CenteredRows<double>{(const double*)d_x, D, m->d_mean},
                                         RowMajor<double>{m->d_prec, D}, StoreNeg<double>{G, D}, st);
                if (rc != VB_OK) return rc;
                gauss_logp_kernel<double><<<grid_for(n * 32, 256, kNumSMs * 8), 256, 0, st>>>(
                    n, D, (const double*)d_x, m->d_mean, G, d_out_logp);
                VB_LAUNCH_CHECK();
                return VB_OK;
            }
            return gauss_logp_grad_f32(m->d_mean, m->d_prec, D, (const float*)d_x, n, d_out_logp,
                                       (float*)d_out_grad_or_null, d_ws, ws_bytes, st, m->d_gauss_cache);
        }
        case VB_MODEL_REGRESSION: {
            const size_t need = vb_model_logp_grad_workspace_bytes(model, n, dtype);
            if (!d_ws || ws_bytes < need) return VB_ERR_WORKSPACE;
            if (dtype == VB_F32 && m->data_dtype == VB_F32 && regression_fused_ok(m))
                return regression_logp_grad_fused(m, (const float*)d_x, n, d_out_logp, (float*)d_out_grad_or_null,
                                                  d_ws, ws_bytes, st);
            if (dtype == VB_F32 && m->data_dtype == VB_F32)
                return regression_logp_grad_tf32(m, (const float*)d_x, n, d_out_logp, (float*)d_out_grad_or_null,
                                                 d_ws, ws_bytes, st);
            if (dtype == VB_F64 && m->data_dtype == VB_F64)
                return run_regression_f64<double, double>(m, (const double*)d_x, n, d_out_logp,
                                                          (double*)d_out_grad_or_null, (char*)d_ws, st);
            if (dtype == VB_F64 && m->data_dtype == VB_F32)
                return run_regression_f64<double, float>(m, (const double*)d_x, n, d_out_logp,
                                                         (double*)d_out_grad_or_null, (char*)d_ws, st);
            return run_regression_f64<float, double>(m, (const float*)d_x, n, d_out_logp,
                                                     (float*)d_out_grad_or_null, (char*)d_ws, st);
        }
        default:
            return VB_ERR_UNSUPPORTED;
    }
}
