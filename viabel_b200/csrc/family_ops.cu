// Variational-family kernels outside the fused step: base draws, sample, logdensity, and
// the generic (unfused) mean-field reductions used with GEMM-shaped model plugins.
// Reference: viabel/vb.py:54-65, 148-162, 202-208.
#include "common.cuh"
#include "philox.cuh"
#include "tc05.cuh"

namespace vb {

static inline int grid_for(long long work_items, int threads, int max_blocks) {
    long long g = (work_items + threads - 1) / threads;
    if (g > max_blocks) g = max_blocks;
    if (g < 1) g = 1;
    return (int)g;
}

// ------------------------------------------------------------------------------------------ base draws
template <typename T>
__global__ void base_draws_kernel(int family, float df, long long n, int D, PhiloxStream rng,
                                  unsigned long long index0, T* __restrict__ out) {
    if (family == VB_FAMILY_MF_T) {          // one thread per (sample, coordinate)
        const long long total = n * D;
        for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
             e += (long long)gridDim.x * blockDim.x) {
            const long long i = e / D;
            const int j = (int)(e - i * D);
            out[e] = (T)draw_student_t(rng, index0 + (uint64_t)i, (uint32_t)j, df);
        }
    } else {                                 // one thread per (sample, quad of coordinates)
        const int Q = (D + 3) / 4;
        const long long total = n * Q;
        for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
             e += (long long)gridDim.x * blockDim.x) {
            const long long i = e / Q;
            const int q = (int)(e - i * Q);
            float z[4];
            draw_normal_pair(rng, index0 + (uint64_t)i, (uint32_t)q, z[0], z[1], z[2], z[3]);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (4 * q + k < D) out[i * D + 4 * q + k] = (T)z[k];
        }
    }
}

// Standard normals of the full-rank family (fp32) written together with their 3xTF32 operand planes (hi, lo; row pitch
// pad4(D), the layout of split_kernel / vb_split_f32): the weighted outer product Z' diag(w) G then needs no split pass
// over z (134 MB read + 268 MB written per step at C3).  Same variates and same split as the two separate kernels.
__global__ void base_draws_split_kernel(long long n, int D, PhiloxStream rng, unsigned long long index0,
                                        float* __restrict__ out, float* __restrict__ planes, long long ldd, long long plane) {
    const int Q = (D + 3) / 4;
    const long long total = n * Q;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long i = e / Q;
        const int q = (int)(e - i * Q);
        float z[4], hi[4], lo[4];
        draw_normal_pair(rng, index0 + (uint64_t)i, (uint32_t)q, z[0], z[1], z[2], z[3]);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            hi[k] = lo[k] = 0.f;
            if (4 * q + k < D) tc::split_tf32(z[k], hi[k], lo[k]);
        }
        // 4 q + k < ldd always (Q = ldd / 4): padding columns of the planes get zeros; 16-byte stores where aligned
        float* ph = planes + i * ldd + 4 * q;
        *reinterpret_cast<float4*>(ph) = make_float4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<float4*>(ph + plane) = make_float4(lo[0], lo[1], lo[2], lo[3]);
        if ((D & 3) == 0) {
            *reinterpret_cast<float4*>(out + i * D + 4 * q) = make_float4(z[0], z[1], z[2], z[3]);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (4 * q + k < D) out[i * D + 4 * q + k] = z[k];
        }
    }
}

template <typename T>
__global__ void base_draws_chi2_kernel(float df, long long n, PhiloxStream rng, unsigned long long index0,
                                       T* __restrict__ out) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x)
        out[i] = (T)draw_chi2(rng, index0 + (uint64_t)i, df);
}

// ------------------------------------------------------------------------------------------ sample
template <typename T>
__global__ void mf_sample_kernel(const double* __restrict__ var_param, long long n, int D,
                                 const T* __restrict__ eps, T* __restrict__ out) {
    const long long total = n * D;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const int j = (int)(e % D);
        out[e] = (T)(var_param[j] + exp(var_param[D + j]) * (double)eps[e]);
    }
}

// D*c0 - sum(l): the lambda-dependent constant of log q for both mean-field families.
__device__ __forceinline__ double mf_logq0_block(int family, double df, const double* var_param, int D,
                                                 double* scratch) {
    double sl = 0.0;
    for (int j = threadIdx.x; j < D; j += blockDim.x) sl += var_param[D + j];
    sl = block_sum(sl, scratch);
    double c0;
    if (family == VB_FAMILY_MF_GAUSS) c0 = -0.9189385332046727;
    else c0 = lgamma(0.5 * (df + 1.0)) - lgamma(0.5 * df) - 0.5 * log(df * 3.14159265358979323846);
    return D * c0 - sl;
}

// ------------------------------------------------------------------------------------------ logdensity (one warp per row)
// log q(x; lambda) = D c0 - sum l - 1/2 sum y^2                      (Gaussian, vb.py:63-65)
//                  = D c0 - sum l - (df+1)/2 sum log1p(y^2/df)       (Student t, vb.py:158-162)
template <typename T>
__global__ void mf_logdensity_kernel(int family, double df, const double* __restrict__ var_param,
                                     long long n, int D, const T* __restrict__ x, double* __restrict__ out) {
    __shared__ double scratch[32];
    const double c_total = mf_logq0_block(family, df, var_param, D, scratch);
    const double half_nu = (family == VB_FAMILY_MF_GAUSS) ? 0.5 : 0.5 * (df + 1.0);
    const int lane = threadIdx.x & 31;
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = warp; i < n; i += nwarps) {
        double s = 0.0;
        for (int j = lane; j < D; j += 32) {
            const double y = ((double)x[i * D + j] - var_param[j]) * exp(-var_param[D + j]);
            s += (family == VB_FAMILY_MF_GAUSS) ? y * y : log1p(y * y / df);
        }
        s = warp_sum(s);
        if (lane == 0) out[i] = c_total - half_nu * s;
    }
}

// ------------------------------------------------------------------------------------------ generic reductions
// rows: lw_i = logp_i - log q_i with q evaluated through the base draw (y == eps); block maxima.
template <typename T>
__global__ void mf_rows_kernel(int family, double df, const double* __restrict__ var_param, long long n, int D,
                               const T* __restrict__ eps, const double* __restrict__ logp,
                               double* __restrict__ lw, double* __restrict__ block_max_out) {
    __shared__ double scratch[32];
    const double logq0 = mf_logq0_block(family, df, var_param, D, scratch);
    const double half_nu = (family == VB_FAMILY_MF_GAUSS) ? 0.5 : 0.5 * (df + 1.0);
    const int lane = threadIdx.x & 31;
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    double m = -INFINITY;
    for (long long i = warp; i < n; i += nwarps) {
        double s = 0.0;
        for (int j = lane; j < D; j += 32) {
            const double e = (double)eps[i * D + j];
            s += (family == VB_FAMILY_MF_GAUSS) ? e * e : log1p(e * e / df);
        }
        s = warp_sum(s);
        const double v = logp[i] - (logq0 - half_nu * s);
        if (lane == 0) lw[i] = v;
        m = fmax(m, v);
    }
    m = block_max(m, scratch);
    if (threadIdx.x == 0) block_max_out[blockIdx.x] = m;
}

__global__ void max_finish_kernel(const double* __restrict__ vals, int count, double* __restrict__ out) {
    __shared__ double scratch[32];
    double m = -INFINITY;
    for (int i = threadIdx.x; i < count; i += blockDim.x) m = fmax(m, vals[i]);
    m = block_max(m, scratch);
    if (threadIdx.x == 0) *out = m;
}

// weights: w_i = exp(alpha (lw_i - M)) (CHIVI) or 1 (KLVI); block sums of (w, second).
__global__ void weights_kernel(int objective, double alpha, long long n, const double* __restrict__ lw,
                               const double* __restrict__ second, const double* __restrict__ M_ptr,
                               double* __restrict__ w_out, double* __restrict__ block_out /* [grid][2] */) {
    __shared__ double scratch[32];
    const double M = (objective == VB_OBJ_CHIVI) ? *M_ptr : 0.0;
    double sw = 0.0, s2 = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const double w = (objective == VB_OBJ_CHIVI) ? exp(alpha * (lw[i] - M)) : 1.0;
        w_out[i] = w;
        sw += w;
        s2 += second[i];
    }
    sw = block_sum(sw, scratch);
    s2 = block_sum(s2, scratch);
    if (threadIdx.x == 0) { block_out[2 * blockIdx.x] = sw; block_out[2 * blockIdx.x + 1] = s2; }
}

// column sums over a chunk of rows: sum_i w_i g_ij and sum_i w_i g_ij eps_ij.
// block (32, 8): x -> column inside a 32-wide tile, y -> row phase.  grid (col_tiles, row_chunks).
template <typename T>
__global__ void weighted_colsum_kernel(long long n, int D, const T* __restrict__ g, const T* __restrict__ eps_or_null,
                                       const double* __restrict__ w, long long rows_per_chunk,
                                       double* __restrict__ out /* [chunks][2][D] */) {
    __shared__ double sh[2][8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const long long r0 = (long long)blockIdx.y * rows_per_chunk;
    long long r1 = r0 + rows_per_chunk;
    if (r1 > n) r1 = n;
    double a0 = 0.0, a1 = 0.0;
    if (c < D) {
        for (long long r = r0 + threadIdx.y; r < r1; r += 8) {
            const double wg = w[r] * (double)g[r * D + c];
            a0 += wg;
            if (eps_or_null) a1 += wg * (double)eps_or_null[r * D + c];
        }
    }
    sh[0][threadIdx.y][threadIdx.x] = a0;
    sh[1][threadIdx.y][threadIdx.x] = a1;
    __syncthreads();
    if (threadIdx.y == 0 && c < D) {
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int y = 0; y < 8; ++y) { s0 += sh[0][y][threadIdx.x]; s1 += sh[1][y][threadIdx.x]; }
        out[((size_t)blockIdx.y * 2 + 0) * D + c] = s0;
        out[((size_t)blockIdx.y * 2 + 1) * D + c] = s1;
    }
}

// merge chunk partials in order into the VB_MF_NPART layout, optionally finalise (Appendix A.1).
__global__ void mf_reduce_finish_kernel(int family, int objective, double alpha, const double* __restrict__ var_param,
                                        int D, long long n, const double* __restrict__ M_ptr,
                                        const double* __restrict__ wblocks, int n_wblocks,
                                        const double* __restrict__ colparts, int n_chunks,
                                        double* __restrict__ out_partials, double* out_value, double* out_grad) {
    extern __shared__ double merged_dyn[];
    const int NP = 2 * D + 4;
    for (int j = threadIdx.x; j < NP; j += blockDim.x) {
        double v = 0.0;
        if (j == 0) v = (objective != VB_OBJ_KLVI) ? *M_ptr : 0.0;
        else if (j == 1) { for (int b = 0; b < n_wblocks; ++b) v += wblocks[2 * b]; }
        else if (j == 2) { for (int b = 0; b < n_wblocks; ++b) v += wblocks[2 * b + 1]; }
        else if (j == 3) v = (double)n;
        else {
            const int which = (j - 4) / D, c = (j - 4) % D;
            for (int k = 0; k < n_chunks; ++k) v += colparts[((size_t)k * 2 + which) * D + c];
        }
        merged_dyn[j] = v;
        out_partials[j] = v;
    }
    __syncthreads();
    if (out_value == nullptr) return;
    const double n_total = (double)n;
    for (int j = threadIdx.x; j < NP; j += blockDim.x) {
        if (j < D) {
            const double sig = exp(var_param[D + j]);
            const double gm = merged_dyn[4 + j], gl = merged_dyn[4 + D + j] * sig;
            if (objective == VB_OBJ_KLVI || objective == VB_OBJ_KLVI_PD) {
                out_grad[j] = -gm / n_total;
                out_grad[D + j] = -(1.0 + gl / n_total);
            } else {
                out_grad[j] = alpha * gm / n_total;
                out_grad[D + j] = alpha * (gl + merged_dyn[1]) / n_total;
            }
        }
        if (j == 0) {
            if (objective == VB_OBJ_KLVI) {
                double H = 0.0;
                for (int k = 0; k < D; ++k) H += var_param[D + k];
                if (family == VB_FAMILY_MF_GAUSS) H += 0.5 * D * (1.0 + 1.8378770664093453);
                *out_value = -(H + merged_dyn[2] / n_total);
            } else if (objective == VB_OBJ_KLVI_PD) {
                *out_value = -(merged_dyn[2] / n_total);          // -(mean log p - mean log q), vb.py:268-278
            } else {
                *out_value = log(merged_dyn[1] / n_total) / alpha + merged_dyn[0];
            }
        }
    }
}

// workspace layout of vb_mf_reduce:
//   [M, pad] | lw (n) | w (n) | row-block maxima (g_rows) | weight blocks (2*g_w) | column partials (chunks*2*D)
struct ReducePlan {
    int g_rows, g_w, chunks, col_tiles;
    long long rows_per_chunk;
    size_t off_lw, off_w, off_bmax, off_wb, off_col, total;
};
static ReducePlan plan_reduce(long long n, int D) {
    ReducePlan p;
    p.g_rows = grid_for(n * 32, 256, kNumSMs * 8);
    p.g_w = grid_for(n, 256, kNumSMs * 4);
    p.col_tiles = (D + 31) / 32;
    long long want = (kNumSMs * 4 + p.col_tiles - 1) / p.col_tiles;
    const long long max_chunks = (n + 63) / 64;
    if (want > max_chunks) want = max_chunks;
    if (want < 1) want = 1;
    p.chunks = (int)want;
    p.rows_per_chunk = (n + p.chunks - 1) / p.chunks;
    size_t o = 2 * sizeof(double);
    p.off_lw = o;   o += (size_t)n * 8;
    p.off_w = o;    o += (size_t)n * 8;
    p.off_bmax = o; o += (size_t)p.g_rows * 8;
    p.off_wb = o;   o += (size_t)p.g_w * 16;
    p.off_col = o;  o += (size_t)p.chunks * 2 * D * 8;
    p.total = o;
    return p;
}

template <typename T>
static int run_mf_reduce(int family, double df, int objective, double alpha, const double* var_param,
                         long long n, int D, const T* eps, const double* logp, const T* grad,
                         double* out_value, double* out_grad, double* out_partials, char* ws,
                         const ReducePlan& p, cudaStream_t st) {
    double* M_ptr = (double*)ws;
    double* lw = (double*)(ws + p.off_lw);
    double* w = (double*)(ws + p.off_w);
    double* bmax = (double*)(ws + p.off_bmax);
    double* wb = (double*)(ws + p.off_wb);
    double* col = (double*)(ws + p.off_col);
    // path-derivative KLVI = the CHIVI passes with alpha = 0 (unit weights, sum of lw), finalised as KLVI
    const int pass_obj = (objective == VB_OBJ_KLVI_PD) ? (int)VB_OBJ_CHIVI : objective;
    const double pass_alpha = (objective == VB_OBJ_KLVI_PD) ? 0.0 : alpha;
    if (pass_obj == VB_OBJ_CHIVI) {
        mf_rows_kernel<T><<<p.g_rows, 256, 0, st>>>(family, df, var_param, n, D, eps, logp, lw, bmax);
        VB_LAUNCH_CHECK();
        max_finish_kernel<<<1, 256, 0, st>>>(bmax, p.g_rows, M_ptr);
        VB_LAUNCH_CHECK();
    }
    weights_kernel<<<p.g_w, 256, 0, st>>>(pass_obj, pass_alpha, n, lw, pass_obj == VB_OBJ_CHIVI ? lw : logp, M_ptr,
                                         w, wb);
    VB_LAUNCH_CHECK();
    weighted_colsum_kernel<T><<<dim3(p.col_tiles, p.chunks), dim3(32, 8), 0, st>>>(n, D, grad, eps, w,
                                                                                 p.rows_per_chunk, col);
    VB_LAUNCH_CHECK();
    mf_reduce_finish_kernel<<<1, 256, (2 * D + 4) * sizeof(double), st>>>(
        family, objective, alpha, var_param, D, n, M_ptr, wb, p.g_w, col, p.chunks, out_partials, out_value, out_grad);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

}  // namespace vb

using namespace vb;

extern "C" int vb_base_draws(int family, double df, int dtype, long long n, int dim,
                             unsigned long long seed, unsigned int call_offset,
                             unsigned long long sample_index0,
                             void* d_out_eps, void* d_out_chi2_or_null, void* stream) {
    if (!d_out_eps || n < 1 || dim < 1) return VB_ERR_ARG;
    if (family != VB_FAMILY_MF_GAUSS && !(df > 2.0)) return VB_ERR_ARG;
    if (family == VB_FAMILY_FR_T && !d_out_chi2_or_null) return VB_ERR_ARG;
    if (dtype != VB_F32 && dtype != VB_F64) return VB_ERR_ARG;
    PhiloxStream rng{(uint32_t)seed, (uint32_t)(seed >> 32), call_offset};
    cudaStream_t st = (cudaStream_t)stream;
    const int fam_eps = (family == VB_FAMILY_MF_T) ? VB_FAMILY_MF_T : VB_FAMILY_MF_GAUSS;
    const long long items = (fam_eps == VB_FAMILY_MF_T) ? n * dim : n * ((dim + 3) / 4);
    const int grid = grid_for(items, 256, kNumSMs * 8);
    if (dtype == VB_F64)
        base_draws_kernel<double><<<grid, 256, 0, st>>>(fam_eps, (float)df, n, dim, rng, sample_index0, (double*)d_out_eps);
    else
        base_draws_kernel<float><<<grid, 256, 0, st>>>(fam_eps, (float)df, n, dim, rng, sample_index0, (float*)d_out_eps);
    VB_LAUNCH_CHECK();
    if (family == VB_FAMILY_FR_T) {
        const int g2 = grid_for(n, 256, kNumSMs * 8);
        if (dtype == VB_F64)
            base_draws_chi2_kernel<double><<<g2, 256, 0, st>>>((float)df, n, rng, sample_index0, (double*)d_out_chi2_or_null);
        else
            base_draws_chi2_kernel<float><<<g2, 256, 0, st>>>((float)df, n, rng, sample_index0, (float*)d_out_chi2_or_null);
        VB_LAUNCH_CHECK();
    }
    return VB_OK;
}

// vb_base_draws for the full-rank family in fp32 storage that also writes the 3xTF32 operand planes of z
// (vb_split_bytes(n, dim) bytes; consumed by vb_fr_weighted_split)
extern "C" int vb_base_draws_fr_split(double df, long long n, int dim, unsigned long long seed, unsigned int call_offset,
                                      unsigned long long sample_index0, float* d_out_z, float* d_out_chi2,
                                      float* d_out_z_planes, void* stream) {
    if (!d_out_z || !d_out_chi2 || !d_out_z_planes || n < 1 || dim < 1 || !(df > 2.0)) return VB_ERR_ARG;
    if ((reinterpret_cast<uintptr_t>(d_out_z_planes) & 15u) || ((dim & 3) == 0 && (reinterpret_cast<uintptr_t>(d_out_z) & 15u)))
        return VB_ERR_ALIGN;
    PhiloxStream rng{(uint32_t)seed, (uint32_t)(seed >> 32), call_offset};
    cudaStream_t st = (cudaStream_t)stream;
    const long long ldd = ((long long)dim + 3) / 4 * 4;
    base_draws_split_kernel<<<grid_for(n * (ldd / 4), 256, kNumSMs * 8), 256, 0, st>>>(n, dim, rng, sample_index0, d_out_z,
                                                                                     d_out_z_planes, ldd, n * ldd);
    VB_LAUNCH_CHECK();
    base_draws_chi2_kernel<float><<<grid_for(n, 256, kNumSMs * 8), 256, 0, st>>>((float)df, n, rng, sample_index0, d_out_chi2);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" int vb_mf_sample(int dtype, const double* d_var_param, long long n, int dim,
                            const void* d_eps, void* d_out_x, void* stream) {
    if (!d_var_param || !d_eps || !d_out_x || n < 1 || dim < 1) return VB_ERR_ARG;
    const int grid = grid_for(n * dim, 256, kNumSMs * 8);
    if (dtype == VB_F64)
        mf_sample_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>(d_var_param, n, dim, (const double*)d_eps, (double*)d_out_x);
    else if (dtype == VB_F32)
        mf_sample_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(d_var_param, n, dim, (const float*)d_eps, (float*)d_out_x);
    else
        return VB_ERR_ARG;
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" int vb_mf_logdensity(int family, double df, int dtype, const double* d_var_param,
                                long long n, int dim, const void* d_x, double* d_out_logq, void* stream) {
    if (!d_var_param || !d_x || !d_out_logq || n < 1 || dim < 1) return VB_ERR_ARG;
    if (family != VB_FAMILY_MF_GAUSS && family != VB_FAMILY_MF_T) return VB_ERR_ARG;
    if (family == VB_FAMILY_MF_T && !(df > 2.0)) return VB_ERR_ARG;
    const int grid = grid_for(n * 32, 256, kNumSMs * 8);
    if (dtype == VB_F64)
        mf_logdensity_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>(family, df, d_var_param, n, dim, (const double*)d_x, d_out_logq);
    else if (dtype == VB_F32)
        mf_logdensity_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(family, df, d_var_param, n, dim, (const float*)d_x, d_out_logq);
    else
        return VB_ERR_ARG;
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" size_t vb_mf_reduce_workspace_bytes(long long n, int dim) {
    if (n < 1 || dim < 1) return 0;
    return plan_reduce(n, dim).total;
}

extern "C" int vb_mf_reduce(int family, double df, int objective, double alpha, int dtype,
                            const double* d_var_param, long long n, int dim,
                            const void* d_eps, const double* d_logp, const void* d_grad,
                            double* d_out_value_or_null, double* d_out_grad_or_null,
                            double* d_out_partials, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_var_param || !d_eps || !d_logp || !d_grad || !d_out_partials || !d_ws || n < 1 || dim < 1)
        return VB_ERR_ARG;
    if (family != VB_FAMILY_MF_GAUSS && family != VB_FAMILY_MF_T) return VB_ERR_ARG;
    if (family == VB_FAMILY_MF_T && !(df > 2.0)) return VB_ERR_ARG;
    if ((d_out_value_or_null == nullptr) != (d_out_grad_or_null == nullptr)) return VB_ERR_ARG;
    const ReducePlan p = plan_reduce(n, dim);
    if (ws_bytes < p.total) return VB_ERR_WORKSPACE;
    if (dtype == VB_F64)
        return run_mf_reduce<double>(family, df, objective, alpha, d_var_param, n, dim, (const double*)d_eps,
                                     d_logp, (const double*)d_grad, d_out_value_or_null, d_out_grad_or_null,
                                     d_out_partials, (char*)d_ws, p, (cudaStream_t)stream);
    if (dtype == VB_F32)
        return run_mf_reduce<float>(family, df, objective, alpha, d_var_param, n, dim, (const float*)d_eps,
                                    d_logp, (const float*)d_grad, d_out_value_or_null, d_out_grad_or_null,
                                    d_out_partials, (char*)d_ws, p, (cudaStream_t)stream);
    return VB_ERR_ARG;
}
