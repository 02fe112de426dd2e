// Host side of the 3xTF32 tcgen05 path: tensor-map creation, operand splitting, the plain GEMM
// entry (vb_gemm with dtype f32) and the fused pipelines built on the kernel in gemm_tf32x3.cuh.
#include <cstdlib>
#include "gemm_tf32x3.cuh"
#include "dense.cuh"
#include "models.cuh"

namespace vb {

static inline int grid_for(long long work_items, int threads, int max_blocks) {
    long long g = (work_items + threads - 1) / threads;
    if (g > max_blocks) g = max_blocks;
    if (g < 1) g = 1;
    return (int)g;
}

EncodeTiledFn get_encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

int tf32x3_num_sms() { return device_sm_count(); }

int make_split_tmap(CUtensorMap* out, const float* base, long long rows, long long cols, long long ld,
                    long long plane_stride_elems, int box_rows, bool mn_major) {
    EncodeTiledFn enc = get_encode_tiled();
    if (!enc) return VB_ERR_UNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(base) & 15u) || (ld % 4) || (plane_stride_elems % 4)) return VB_ERR_ALIGN;
    const cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, 2};
    const cuuint64_t strides[2] = {(cuuint64_t)ld * 4, (cuuint64_t)plane_stride_elems * 4};
    const cuuint32_t box[3] = {32, (cuuint32_t)box_rows, 2};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE,
                           mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? VB_OK : VB_ERR_ARG;
}

int make_raw_tmap(CUtensorMap* out, const float* base, long long rows, long long cols, long long ld, int box_rows,
                  bool atom32) {
    EncodeTiledFn enc = get_encode_tiled();
    if (!enc) return VB_ERR_UNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(base) & 15u) || (ld % 4)) return VB_ERR_ALIGN;
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {32, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE,
                           atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? VB_OK : VB_ERR_ARG;
}

int make_f16_planes_tmap(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld,
                         long long plane_stride_elems, int box_rows) {
    EncodeTiledFn enc = get_encode_tiled();
    if (!enc) return VB_ERR_UNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(base) & 15u) || (ld % 8) || (plane_stride_elems % 8)) return VB_ERR_ALIGN;
    const cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, 2};
    const cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)plane_stride_elems * 2};
    const cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 2};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? VB_OK : VB_ERR_ARG;
}

// ------------------------------------------------------------------------------------------ operand splitting
// src (rows, cols) with leading dimension lds -> dst planes (rows, ldd): hi at dst, lo at dst + plane
__global__ void split_kernel(const float* __restrict__ src, long long rows, long long cols, long long lds,
                             float* __restrict__ dst, long long ldd, long long plane, float scale) {
    const long long total = rows * ldd;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long r = e / ldd, c = e - r * ldd;
        float hi = 0.f, lo = 0.f;
        if (c < cols) tc::split_tf32(scale * src[r * lds + c], hi, lo);
        dst[e] = hi;
        dst[e + plane] = lo;
    }
}

static inline long long pad4(long long x) { return (x + 3) / 4 * 4; }

// ------------------------------------------------------------------------------------------ plain GEMM epilogue
// 256-bit global accesses (sm_100: STG/LDG.E.ENL2.256); addresses 32-byte aligned.  Each thread of an
// epilogue warp owns one row, so a warp-wide store touches 32 different lines: with 256-bit accesses
// every request is at least a full 32-byte sector (scalar stores were the bottleneck of every epilogue).
__device__ __forceinline__ void st_global_v8(float* dst, const float (&a)[8]) {
    asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 :: "l"(dst), "f"(a[0]), "f"(a[1]), "f"(a[2]), "f"(a[3]), "f"(a[4]), "f"(a[5]), "f"(a[6]), "f"(a[7])
                 : "memory");
}
__device__ __forceinline__ void ld_global_v8(const float* src, float (&a)[8]) {
    asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=f"(a[0]), "=f"(a[1]), "=f"(a[2]), "=f"(a[3]), "=f"(a[4]), "=f"(a[5]), "=f"(a[6]), "=f"(a[7])
                 : "l"(src));
}
__device__ __forceinline__ bool aligned32(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 31u) == 0; }
// p[0..ncols) = w[0..ncols): four 256-bit stores for a full, aligned block, scalar otherwise
__device__ __forceinline__ void store_row32(float* p, const float (&w)[32], int ncols) {
    if (ncols == 32 && aligned32(p)) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
            const float t[8] = {w[j], w[j + 1], w[j + 2], w[j + 3], w[j + 4], w[j + 5], w[j + 6], w[j + 7]};
            st_global_v8(p + j, t);
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (j < ncols) p[j] = w[j];
    }
}

struct AxpbyEpi {
    struct Pre {};
    __device__ __forceinline__ Pre prefetch(long long, long long, int) const { return Pre{}; }
    float* c; long long ldc; float alpha, beta;
    __device__ __forceinline__ void operator()(long long row, long long col0, float (&v)[32], int ncols, int, const Pre&) const {
        float* p = c + row * ldc + col0;
        if (beta == 0.f) {
            float w[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) w[j] = alpha * v[j];
            store_row32(p, w, ncols);
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (j < ncols) p[j] = alpha * v[j] + beta * p[j];
        }
    }
};

size_t gemm_tf32x3_workspace_bytes(long long m, long long n, long long k) {
    // split copies of A and B (either orientation needs rows*pad4(cols) <= (dim)*(pad4(other)) )
    const size_t a = (size_t)2 * (size_t)(m > k ? m : k) * (size_t)pad4(m > k ? k : m) * 4 + 1024;
    const size_t a2 = (size_t)2 * (size_t)m * pad4(k) * 4, a3 = (size_t)2 * (size_t)k * pad4(m) * 4;
    const size_t b2 = (size_t)2 * (size_t)k * pad4(n) * 4, b3 = (size_t)2 * (size_t)n * pad4(k) * 4;
    (void)a;
    return (a2 > a3 ? a2 : a3) + (b2 > b3 ? b2 : b3) + 2048;
}

int gemm_tf32x3(int transa, int transb, long long m, long long n, long long k, float alpha,
                const float* A, long long lda, const float* B, long long ldb, float beta,
                float* C, long long ldc, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (!ws || ws_bytes < gemm_tf32x3_workspace_bytes(m, n, k)) return VB_ERR_WORKSPACE;
    // stored shapes
    const long long ar = transa ? k : m, ac = transa ? m : k;
    const long long br = transb ? n : k, bc = transb ? k : n;
    const long long alds = pad4(ac), blds = pad4(bc);
    char* w = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
    float* As = (float*)w;
    float* Bs = (float*)(w + (((size_t)2 * ar * alds * 4 + 255) & ~(size_t)255));
    split_kernel<<<grid_for(ar * alds, 256, kNumSMs * 8), 256, 0, st>>>(A, ar, ac, lda, As, alds, ar * alds, 1.f);
    VB_LAUNCH_CHECK();
    split_kernel<<<grid_for(br * blds, 256, kNumSMs * 8), 256, 0, st>>>(B, br, bc, ldb, Bs, blds, br * blds, 1.f);
    VB_LAUNCH_CHECK();
    const SplitMat a{As, ar, ac, alds, ar * alds}, b{Bs, br, bc, blds, br * blds};
    const GemmShape shape = plan_gemm(m, n, k, 1);
    const AxpbyEpi epi{C, ldc, alpha, beta};
    // A is K-major when stored (m, k); B is K-major when stored (n, k)
    if (!transa && transb) return launch_tf32x3<false, false>(a, b, shape, epi, st);
    if (!transa && !transb) return launch_tf32x3<false, true>(a, b, shape, epi, st);
    if (transa && transb) return launch_tf32x3<true, false>(a, b, shape, epi, st);
    return launch_tf32x3<true, true>(a, b, shape, epi, st);
}

// ------------------------------------------------------------------------------------------ full-rank sample
// x_i = mu + (z_i S) / sqrt(chi2_i / df)            (viabel/vb.py:204-208), S given as stored (k, n)
struct FrSampleEpi {
    struct Pre {};
    __device__ __forceinline__ Pre prefetch(long long, long long, int) const { return Pre{}; }
    float* x; long long ld; const double* mu; const float* chi2; float df;
    __device__ __forceinline__ void operator()(long long row, long long col0, float (&v)[32], int ncols, int, const Pre&) const {
        const float s = rsqrtf(chi2[row] / df);
        float* p = x + row * ld + col0;
        float w[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) w[j] = (j < ncols) ? (float)mu[col0 + j] + v[j] * s : 0.f;
        store_row32(p, w, ncols);
    }
};

static size_t split_bytes(long long rows, long long cols) {
    return ((size_t)2 * rows * pad4(cols) * 4 + 255) & ~(size_t)255;
}
static int split_into(const float* src, long long rows, long long cols, long long lds, float* dst, cudaStream_t st,
                      SplitMat* out) {
    const long long ldd = pad4(cols);
    split_kernel<<<grid_for(rows * ldd, 256, kNumSMs * 8), 256, 0, st>>>(src, rows, cols, lds, dst, ldd, rows * ldd, 1.f);
    VB_LAUNCH_CHECK();
    *out = SplitMat{dst, rows, cols, ldd, rows * ldd};
    return VB_OK;
}

size_t fr_sample_tf32_workspace_bytes(long long n, int dim) {
    return split_bytes(n, dim) + split_bytes(dim, dim) + 512;
}

int fr_sample_tf32(const double* mu, const float* scale, double df, long long n, int dim, const float* z,
                   const float* chi2, float* out_x, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (!ws || ws_bytes < fr_sample_tf32_workspace_bytes(n, dim)) return VB_ERR_WORKSPACE;
    char* w = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
    SplitMat zs, ss;
    int rc = split_into(scale, dim, dim, dim, (float*)(w + split_bytes(n, dim)), st, &ss);
    if (rc != VB_OK) return rc;
    const GemmShape shape = plan_gemm(n, dim, dim, 1);
    const FrSampleEpi epi{out_x, dim, mu, chi2, (float)df};
    if (dim % 4 == 0 && (reinterpret_cast<uintptr_t>(z) & 15u) == 0)    // z read as plain fp32, split on chip into TMEM
        return launch_tf32x3<false, true, FrSampleEpi, 4, false, true, false>(SplitMat{z, n, dim, dim, 0}, ss, shape, epi, st);
    rc = split_into(z, n, dim, dim, (float*)w, st, &zs);
    if (rc != VB_OK) return rc;
    return launch_tf32x3<false, true>(zs, ss, shape, epi, st);
}

// ------------------------------------------------------------------------------------------ correlated Gaussian
// G = -(X - m) P = -X P + (m P);  log p_i = 1/2 sum_j G_ij (X_ij - m_j)
__global__ void gauss_prep_kernel(const double* __restrict__ mean, const double* __restrict__ prec, int D,
                                  float* __restrict__ prec_f32, float* __restrict__ bias) {
    // bias_j = sum_k m_k P_kj (fp64), prec_f32 = (float) P
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < D; j += gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < D; ++k) s = fma(mean[k], prec[(size_t)k * D + j], s);
        bias[j] = (float)s;
    }
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < (long long)D * D;
         e += (long long)gridDim.x * blockDim.x)
        prec_f32[e] = (float)prec[e];
}
struct GaussEpi {
    float* g; long long ld; const float* bias; const float* x; const double* mean; float* rowpart; int n_chunks;
    struct Pre { float x[32]; };
    __device__ __forceinline__ Pre prefetch(long long row, long long col0, int ncols) const {
        Pre p;
        const float* xp = x + row * ld + col0;
        if (ncols == 32 && aligned32(xp)) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
                float t[8];
                ld_global_v8(xp + j, t);
#pragma unroll
                for (int q = 0; q < 8; ++q) p.x[j + q] = t[q];
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) p.x[j] = (j < ncols) ? xp[j] : 0.f;
        }
        return p;
    }
    __device__ __forceinline__ void operator()(long long row, long long col0, float (&v)[32], int ncols, int,
                                               const Pre& pre) const {
        float w[32];
        float dot = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const bool ok = j < ncols;
            const float gv = ok ? bias[col0 + j] - v[j] : 0.f;
            w[j] = gv;
            dot = fmaf(gv, ok ? pre.x[j] - (float)mean[col0 + j] : 0.f, dot);
        }
        store_row32(g + row * ld + col0, w, ncols);
        rowpart[row * n_chunks + (col0 >> 5)] = dot;
    }
};
__global__ void rowpart_sum_kernel(const float* __restrict__ rowpart, long long n, int n_chunks, double scale,
                                   double* __restrict__ out) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int c = 0; c < n_chunks; ++c) s += (double)rowpart[i * n_chunks + c];
        out[i] = scale * s;
    }
}

size_t gauss_f32_workspace_bytes(long long n, int dim) {
    const size_t chunks = (dim + 31) / 32;
    return split_bytes(n, dim) + split_bytes(dim, dim) + (((size_t)dim * dim * 4 + 255) & ~(size_t)255) +
           (((size_t)dim * 4 + 255) & ~(size_t)255) + (((size_t)n * chunks * 4 + 255) & ~(size_t)255) +
           (((size_t)n * dim * 4 + 255) & ~(size_t)255) + 512;
}

size_t gauss_cache_bytes(int dim) { return split_bytes(dim, dim) + (((size_t)dim * 4 + 255) & ~(size_t)255) + (((size_t)dim * dim * 4 + 255) & ~(size_t)255); }
// model constants of the fp32 path, made once: tf32 planes of float(P), bias = mean' P, (scratch: float(P))
int gauss_fill_cache(const double* mean, const double* prec, int dim, float* cache, cudaStream_t st) {
    char* w = (char*)cache;
    float* planes = (float*)w;                  w += split_bytes(dim, dim);
    float* bias = (float*)w;                    w += ((size_t)dim * 4 + 255) & ~(size_t)255;
    float* p32 = (float*)w;
    gauss_prep_kernel<<<grid_for((long long)dim * dim, 256, kNumSMs * 2), 256, 0, st>>>(mean, prec, dim, p32, bias);
    VB_LAUNCH_CHECK();
    SplitMat ps;
    return split_into(p32, dim, dim, dim, planes, st, &ps);
}
int gauss_logp_grad_f32(const double* mean, const double* prec, int dim, const float* x, long long n,
                        double* out_logp, float* out_grad_or_null, void* ws, size_t ws_bytes, cudaStream_t st,
                        const float* cache) {
    if (!ws || ws_bytes < gauss_f32_workspace_bytes(n, dim)) return VB_ERR_WORKSPACE;
    const int chunks = (dim + 31) / 32;
    char* w = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
    float* xs_buf = (float*)w;                  w += split_bytes(n, dim);
    float* ps_buf = (float*)w;                  w += split_bytes(dim, dim);
    float* p32 = (float*)w;                     w += ((size_t)dim * dim * 4 + 255) & ~(size_t)255;
    float* bias = (float*)w;                    w += ((size_t)dim * 4 + 255) & ~(size_t)255;
    float* rowpart = (float*)w;                 w += ((size_t)n * chunks * 4 + 255) & ~(size_t)255;
    float* g = out_grad_or_null ? out_grad_or_null : (float*)w;
    SplitMat xs, ps;
    int rc = VB_OK;
    if (cache) {
        const long long ldd = pad4(dim);
        ps = SplitMat{cache, dim, dim, ldd, (long long)dim * ldd};
        bias = const_cast<float*>(cache) + split_bytes(dim, dim) / 4;
    } else {
        gauss_prep_kernel<<<grid_for((long long)dim * dim, 256, kNumSMs * 2), 256, 0, st>>>(mean, prec, dim, p32, bias);
        VB_LAUNCH_CHECK();
        rc = split_into(p32, dim, dim, dim, ps_buf, st, &ps);
        if (rc != VB_OK) return rc;
    }
    const GemmShape shape = plan_gemm(n, dim, dim, 1);
    const GaussEpi epi{g, dim, bias, x, mean, rowpart, chunks};
    if (dim % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15u) == 0) {  // x read as plain fp32, split on chip into TMEM
        rc = launch_tf32x3<false, true, GaussEpi, 4, false, true, false>(SplitMat{x, n, dim, dim, 0}, ps, shape, epi, st);
    } else {
        rc = split_into(x, n, dim, dim, xs_buf, st, &xs);
        if (rc != VB_OK) return rc;
        rc = launch_tf32x3<false, true>(xs, ps, shape, epi, st);
    }
    if (rc != VB_OK) return rc;
    rowpart_sum_kernel<<<grid_for(n, 256, kNumSMs * 8), 256, 0, st>>>(rowpart, n, chunks, 0.5, out_logp);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ------------------------------------------------------------------------------------------ G_S = Ztilde' diag(w) G
// A(a, i) = z[i, a] (MN-major), B(i, b) = (w_i / s_i) g[i, b] (MN-major), K = n, split-K partials in fp32,
// reduced in fp64.
// Also returns the weighted column sums sum_i w_i g[i, c] (the gradient w.r.t. the mean) when `colparts` is given:
// the launch then has a thread count that is a multiple of ldd, so a thread always meets the same column and keeps
// its sum in a register; colparts[block][thread] are merged in fixed order by colparts_sum_kernel.
__global__ void split_rowscaled_kernel(const float* __restrict__ src, long long rows, long long cols,
                                       const double* __restrict__ w, const float* __restrict__ chi2, float df,
                                       float* __restrict__ dst, long long ldd, long long plane,
                                       double* __restrict__ colparts) {
    const long long total = rows * ldd;
    double acc = 0.0;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long r = e / ldd, c = e - r * ldd;
        float hi = 0.f, lo = 0.f;
        if (c < cols) {
            const double wr = w[r];
            const float gv = src[r * cols + c];
            const float sc = (float)(wr / sqrt((double)chi2[r] / (double)df));
            tc::split_tf32(sc * gv, hi, lo);
            acc = fma(wr, (double)gv, acc);
        }
        dst[e] = hi;
        dst[e + plane] = lo;
    }
    if (colparts) colparts[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
// The same with four columns per thread (cols a multiple of 4, 16-byte aligned rows) and the thread count a multiple of
// ldd / 4: a thread keeps its four columns for the whole kernel, its row index advances by a constant (no division per
// element) and the row scale (an fp64 square root and division) is taken once per four elements.  The scalar version
// spent 220 us on 134 MB (C3); this one is bound by its 0.4 GB of traffic.  colparts[4 t + q] belongs to column
// (4 t + q) mod ldd, as colparts_sum_kernel expects.
__global__ void __launch_bounds__(256)
split_rowscaled_vec4_kernel(const float* __restrict__ src, long long rows, long long cols,
                            const double* __restrict__ w, const float* __restrict__ chi2, float df,
                            float* __restrict__ dst, long long ldd, long long plane, double* __restrict__ colparts) {
    const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    const long long v0 = 4 * t;
    const long long c = v0 % ldd, r0 = v0 / ldd, rstep = 4 * nthreads / ldd;
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    const bool live = c < cols;                           // padding columns of dst are written as zeros
    for (long long r = r0; r < rows; r += rstep) {
        float4 hi = make_float4(0.f, 0.f, 0.f, 0.f), lo = hi;
        if (live) {
            const double wr = w[r];
            const float4 g = *reinterpret_cast<const float4*>(src + r * cols + c);
            const float sc = (float)(wr / sqrt((double)chi2[r] / (double)df));
            tc::split_tf32(sc * g.x, hi.x, lo.x);
            tc::split_tf32(sc * g.y, hi.y, lo.y);
            tc::split_tf32(sc * g.z, hi.z, lo.z);
            tc::split_tf32(sc * g.w, hi.w, lo.w);
            acc[0] = fma(wr, (double)g.x, acc[0]);
            acc[1] = fma(wr, (double)g.y, acc[1]);
            acc[2] = fma(wr, (double)g.z, acc[2]);
            acc[3] = fma(wr, (double)g.w, acc[3]);
        }
        *reinterpret_cast<float4*>(dst + r * ldd + c) = hi;
        *reinterpret_cast<float4*>(dst + r * ldd + c + plane) = lo;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) colparts[4 * t + q] = acc[q];
}
// out[c] = sum over the threads t with (t mod ldd) == c of colparts[t], ascending t (fixed order)
__global__ void colparts_sum_kernel(const double* __restrict__ colparts, long long n_threads, long long ldd, int cols,
                                    double* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long c = warp; c < cols; c += nwarps) {
        double s = 0.0;
        for (long long t = c + (long long)lane * ldd; t < n_threads; t += 32 * ldd) s += colparts[t];
        s = warp_sum(s);
        if (lane == 0) out[c] = s;
    }
}
struct SplitKEpi {
    struct Pre {};
    __device__ __forceinline__ Pre prefetch(long long, long long, int) const { return Pre{}; }
    float* parts; long long mn; long long ld;
    __device__ __forceinline__ void operator()(long long row, long long col0, float (&v)[32], int ncols, int split, const Pre&) const {
        store_row32(parts + (size_t)split * mn + row * ld + col0, v, ncols);
    }
};
// Split-K partials accumulated in fp64 across launches: every (tile, split) work item owns its slice of
// acc[split][M][N], so the read-modify-write needs no atomics and the summation order is fixed.
struct SplitKAccEpi {
    struct Pre {};
    __device__ __forceinline__ Pre prefetch(long long, long long, int) const { return Pre{}; }
    double* acc; long long mn; long long ld;
    __device__ __forceinline__ void operator()(long long row, long long col0, float (&v)[32], int ncols, int split, const Pre&) const {
        double* p = acc + (size_t)split * mn + row * ld + col0;
        if (ncols == 32 && ((reinterpret_cast<uintptr_t>(p) & 15u) == 0)) {
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
                double2 t = *reinterpret_cast<double2*>(p + j);
                t.x += (double)v[j]; t.y += (double)v[j + 1];
                *reinterpret_cast<double2*>(p + j) = t;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (j < ncols) p[j] += (double)v[j];
        }
    }
};
__global__ void splitk_sum_f64_kernel(const float* __restrict__ parts, int splits, long long mn, double* __restrict__ out) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < mn; e += (long long)gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < splits; ++k) s += (double)parts[(size_t)k * mn + e];
        out[e] = s;
    }
}

size_t fr_weighted_tf32_workspace_bytes(long long n, int dim) {
    const int splits = plan_gemm(dim, dim, n, auto_splits(dim, dim, n)).splits;
    return 2 * split_bytes(n, dim) + (((size_t)splits * dim * dim * 4 + 255) & ~(size_t)255) +
           (size_t)kNumSMs * 8 * 256 * 8 + 512;                    // + per-thread column-sum partials
}
bool fr_weighted_colsum_fused(long long n, int dim) {
    const long long ldd = pad4(dim);
    return ((long long)grid_for(n * ldd, 256, kNumSMs * 8) * 256) % ldd == 0;
}

int fr_weighted_GS_tf32(double df, long long n, int dim, const float* z, const float* chi2, const double* w,
                        const float* grad, double* out_GS, void* ws, size_t ws_bytes, cudaStream_t st,
                        double* out_gmu_or_null, const float* z_planes_or_null) {
    if (!ws || ws_bytes < fr_weighted_tf32_workspace_bytes(n, dim)) return VB_ERR_WORKSPACE;
    char* wp = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
    float* zs_buf = (float*)wp;     wp += split_bytes(n, dim);
    float* gs_buf = (float*)wp;     wp += split_bytes(n, dim);
    float* parts = (float*)wp;
    SplitMat zs;
    int rc = VB_OK;
    const long long ldd = pad4(dim);
    if (z_planes_or_null) {                              // made with the draws (vb_base_draws_fr_split): no split pass
        zs = SplitMat{z_planes_or_null, n, dim, ldd, n * ldd};
    } else {
        rc = split_into(z, n, dim, dim, zs_buf, st, &zs);
        if (rc != VB_OK) return rc;
    }
    // the weighted column sums ride along when the thread count is a multiple of the padded row length
    const int sgrid = grid_for(n * ldd, 256, kNumSMs * 8);
    const long long sthreads = (long long)sgrid * 256;
    const bool fuse = out_gmu_or_null != nullptr && sthreads % ldd == 0;
    double* colparts = fuse ? (double*)(wp + (((size_t)plan_gemm(dim, dim, n, auto_splits(dim, dim, n)).splits * dim * dim * 4 + 255) & ~(size_t)255)) : nullptr;
    const bool vec4 = fuse && dim % 4 == 0 && sgrid % 4 == 0 && ((reinterpret_cast<uintptr_t>(grad) | reinterpret_cast<uintptr_t>(gs_buf)) & 15u) == 0 &&
                      ((n * ldd) % 4 == 0);
    if (vec4)
        split_rowscaled_vec4_kernel<<<sgrid / 4, 256, 0, st>>>(grad, n, dim, w, chi2, (float)df, gs_buf, ldd, n * ldd, colparts);
    else
        split_rowscaled_kernel<<<sgrid, 256, 0, st>>>(grad, n, dim, w, chi2, (float)df, gs_buf, ldd, n * ldd, colparts);
    VB_LAUNCH_CHECK();
    if (fuse) {
        colparts_sum_kernel<<<grid_for((long long)dim * 32, 256, kNumSMs), 256, 0, st>>>(colparts, sthreads, ldd, dim, out_gmu_or_null);
        VB_LAUNCH_CHECK();
    }
    const SplitMat gs{gs_buf, n, dim, ldd, n * ldd};
    const GemmShape shape = plan_gemm(dim, dim, n, auto_splits(dim, dim, n));
    const long long mn = (long long)dim * dim;
    rc = launch_tf32x3<true, true>(zs, gs, shape, SplitKEpi{parts, mn, dim}, st);
    if (rc != VB_OK) return rc;
    splitk_sum_f64_kernel<<<grid_for(mn, 256, kNumSMs * 4), 256, 0, st>>>(parts, shape.splits, mn, out_GS);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

}  // namespace vb

// ------------------------------------------------------------------------------------------ split-operand C ABI
extern "C" size_t vb_split_bytes(long long rows, long long cols) {
    if (rows < 1 || cols < 1) return 0;
    return vb::split_bytes(rows, cols);
}
extern "C" int vb_split_f32(const float* d_src, long long rows, long long cols, long long ld,
                            float* d_dst_planes, void* stream) {
    if (!d_src || !d_dst_planes || rows < 1 || cols < 1 || ld < cols) return VB_ERR_ARG;
    vb::SplitMat out;
    return vb::split_into(d_src, rows, cols, ld, d_dst_planes, (cudaStream_t)stream, &out);
}
extern "C" int vb_gemm_tf32x3_split(int a_mn_major, int b_mn_major, long long m, long long n, long long k,
                                    const float* d_A_planes, const float* d_B_planes,
                                    float* d_C, long long ldc, void* stream) {
    using namespace vb;
    if (!d_A_planes || !d_B_planes || !d_C || m < 1 || n < 1 || k < 1 || ldc < n) return VB_ERR_ARG;
    const long long ar = a_mn_major ? k : m, ac = a_mn_major ? m : k;
    const long long br = b_mn_major ? k : n, bc = b_mn_major ? n : k;
    const SplitMat a{d_A_planes, ar, ac, pad4(ac), ar * pad4(ac)}, b{d_B_planes, br, bc, pad4(bc), br * pad4(bc)};
    const GemmShape shape = plan_gemm(m, n, k, 1);
    const AxpbyEpi epi{d_C, ldc, 1.f, 0.f};
    cudaStream_t st = (cudaStream_t)stream;
    if (!a_mn_major && !b_mn_major) return launch_tf32x3<false, false>(a, b, shape, epi, st);
    if (!a_mn_major && b_mn_major) return launch_tf32x3<false, true>(a, b, shape, epi, st);
    if (a_mn_major && !b_mn_major) return launch_tf32x3<true, false>(a, b, shape, epi, st);
    return launch_tf32x3<true, true>(a, b, shape, epi, st);
}

namespace vb {
// ------------------------------------------------------------------------------------------ Student-t regression
// log p(b) = -|b|^2 / (2 sd^2) - (df+1)/2 sum_r log1p(r_r^2/df),  r = y - X b
// grad     = -b / sd^2 + psi' X,  psi_r = (df+1) r_r / (df + r_r^2)                  
// Batched over S samples and tiled over chunks of C design-matrix rows:
//   GEMM1  acc (S x C) = B (S x d) X_chunk'  with the fused epilogue below: r, the log1p partial sums and
//          psi written straight into split (hi, lo) planes -- the S x N residual matrix is never stored
//   GEMM2  G (S x d) += Psi (S x C) X_chunk  (split-K over the chunk, partials summed into fp64)
// The same split copy of X serves both products: K-major B operand in GEMM1, MN-major B operand in GEMM2.
__device__ __noinline__ float log1p_sum4_slow(float r0, float r1, float r2, float r3, float inv_df) {
    return (log1pf(r0 * r0 * inv_df) + log1pf(r1 * r1 * inv_df)) + (log1pf(r2 * r2 * inv_df) + log1pf(r3 * r3 * inv_df));
}
struct PsiEpi {
    static constexpr bool kStaged = true;
    alignas(64) CUtensorMap tmap;                    // Psi / c as a plain fp32 (S, cols) matrix, 32 x 32 store box
    const float* __restrict__ y;                     // y + r0
    float* __restrict__ rowpart; int n_blocks;       // (S, C/32) partial sums of log1p(r^2/df)
    float inv_df;
    // y is indexed by the column: the kernel stages the tile's 128-float segment of it in shared memory (bulk copy
    // issued when the accumulator stage frees up), so the epilogue reads it with broadcast LDS instead of waiting
    // for L2 in the middle of the arithmetic.
    static constexpr bool kColVec = true;
    __device__ __forceinline__ const float* colvec() const { return y; }
    struct Pre {};
    __device__ __forceinline__ Pre prefetch(long long, long long, int) const { return Pre{}; }
    // Four elements at a time: u_i = 1 + r_i^2/df, sum_i log1p(r_i^2/df) = log(u_0 u_1 u_2 u_3) (one log per four
    // elements; the rare product overflow falls back to log1pf per element), psi_i / c = r_i / u_i with the
    // approximate reciprocal (1 ulp; u >= 1) -- the constant c = (df+1)/df is applied once to the accumulated
    // gradient.  Psi is stored as plain fp32; the consumer GEMM splits it into tf32 hi / lo in shared memory.
    static __device__ __forceinline__ float log_ge1(float p) {
        // p in [1, 3e38): p = 2^e m, m in [2/3, 4/3); log m = f - f^2/2 + f^3 Q(f), f = m - 1, Q = degree-6 minimax
        // fit (max abs error 3.4e-8 over the interval, fitted in scripts/fit_log_poly.py)
        const uint32_t bits = __float_as_uint(p);
        const uint32_t e = (bits - 0x3f2aaaabu) & 0xff800000u;
        const float f = __uint_as_float(bits - e) - 1.f;
        float q = 0.1440092772245407f;
        q = fmaf(q, f, -0.15269480645656586f);
        q = fmaf(q, f, 0.13915356993675232f);
        q = fmaf(q, f, -0.1644672304391861f);
        q = fmaf(q, f, 0.20016367733478546f);
        q = fmaf(q, f, -0.2500529885292053f);
        q = fmaf(q, f, 0.33333104848861694f);
        q = fmaf(q, f, -0.5f);
        q = fmaf(f, q * f, f);
        return fmaf((float)(int)e, 8.262958294867817e-08f, q);     // e * 2^-23 * ln 2
    }
    // packed fp32x2 arithmetic (sm_100 FADD2 / FMUL2 / FFMA2): the epilogue is issue-bound, not pipe-bound
    static __device__ __forceinline__ float2 log_ge1_x2(float2 p) {
        const uint32_t bx = __float_as_uint(p.x), by = __float_as_uint(p.y);
        const uint32_t ex = (bx - 0x3f2aaaabu) & 0xff800000u, ey = (by - 0x3f2aaaabu) & 0xff800000u;
        const float2 f = __fadd2_rn(make_float2(__uint_as_float(bx - ex), __uint_as_float(by - ey)), make_float2(-1.f, -1.f));
        float2 q = make_float2(0.1440092772245407f, 0.1440092772245407f);
        q = __ffma2_rn(q, f, make_float2(-0.15269480645656586f, -0.15269480645656586f));
        q = __ffma2_rn(q, f, make_float2(0.13915356993675232f, 0.13915356993675232f));
        q = __ffma2_rn(q, f, make_float2(-0.1644672304391861f, -0.1644672304391861f));
        q = __ffma2_rn(q, f, make_float2(0.20016367733478546f, 0.20016367733478546f));
        q = __ffma2_rn(q, f, make_float2(-0.2500529885292053f, -0.2500529885292053f));
        q = __ffma2_rn(q, f, make_float2(0.33333104848861694f, 0.33333104848861694f));
        q = __ffma2_rn(q, f, make_float2(-0.5f, -0.5f));
        q = __ffma2_rn(f, __fmul2_rn(q, f), f);
        return __ffma2_rn(make_float2((float)(int)ex, (float)(int)ey), make_float2(8.262958294867817e-08f, 8.262958294867817e-08f), q);
    }
    // eight elements: two products of four, one packed log
    __device__ __forceinline__ void oct(const float* yv, const float* av, float2& acc, float4& o0, float4& o1) const {
        const float2 neg1 = make_float2(-1.f, -1.f), one = make_float2(1.f, 1.f), idf = make_float2(inv_df, inv_df);
        float2 r[4], u[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            r[i] = __ffma2_rn(make_float2(av[2 * i], av[2 * i + 1]), neg1, make_float2(yv[2 * i], yv[2 * i + 1]));
            u[i] = __ffma2_rn(__fmul2_rn(r[i], r[i]), idf, one);
        }
        const float2 pa = __fmul2_rn(u[0], u[1]), pb = __fmul2_rn(u[2], u[3]);
        const float2 p = __fmul2_rn(make_float2(pa.x, pb.x), make_float2(pa.y, pb.y));
        if (p.x < 3.0e38f && p.y < 3.0e38f) {
            acc = __fadd2_rn(acc, log_ge1_x2(p));
        } else {
            acc.x += log1p_sum4_slow(r[0].x, r[0].y, r[1].x, r[1].y, inv_df);
            acc.y += log1p_sum4_slow(r[2].x, r[2].y, r[3].x, r[3].y, inv_df);
        }
        float2 o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float ix, iy;
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ix) : "f"(u[i].x));
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iy) : "f"(u[i].y));
            o[i] = __fmul2_rn(r[i], make_float2(ix, iy));
        }
        o0 = make_float4(o[0].x, o[0].y, o[1].x, o[1].y);
        o1 = make_float4(o[2].x, o[2].y, o[3].x, o[3].y);
    }
    // Thread = one row of the warp's 32 x 32 block.  The block is written in the 128-byte-swizzled layout the
    // TMA store expects: 16-byte unit j of row r lives at r*128 + ((j ^ (r & 7)) << 4) (conflict-free STS.128).
    __device__ __forceinline__ void staged(long long row, long long col0, float (&v)[32], int ncols, int,
                                           const Pre& pre, unsigned char* staging, int lane, const float* ysm) const {
        float2 acc = make_float2(0.f, 0.f);
        // running sum of this (row, column block of the chunk) over the chunks: the slot belongs to this thread for
        // the whole launch, so a plain read-modify-write keeps the order fixed (chunk by chunk)
        float* slot = rowpart + row * n_blocks + (col0 >> 5);
        const float prev = *slot;
        unsigned char* my = staging + lane * 128;
        const float* yc = y + col0;
        if (ncols == 32) {
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
                const float4 t0 = *reinterpret_cast<const float4*>(ysm + 4 * j);
                const float4 t1 = *reinterpret_cast<const float4*>(ysm + 4 * j + 4);
                const float yy[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
                float4 o0, o1;
                oct(yy, v + 4 * j, acc, o0, o1);
                *reinterpret_cast<float4*>(my + ((j ^ (lane & 7)) << 4)) = o0;
                *reinterpret_cast<float4*>(my + (((j + 1) ^ (lane & 7)) << 4)) = o1;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
                float yy[8], aa[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) {      // masked columns: r = 0 -> u = 1 -> no contribution; the store clips them
                    const bool on = 4 * j + q < ncols;
                    yy[q] = on ? __ldg(yc + 4 * j + q) : 0.f;
                    aa[q] = on ? v[4 * j + q] : 0.f;
                }
                float4 o0, o1;
                oct(yy, aa, acc, o0, o1);
                *reinterpret_cast<float4*>(my + ((j ^ (lane & 7)) << 4)) = o0;
                *reinterpret_cast<float4*>(my + (((j + 1) ^ (lane & 7)) << 4)) = o1;
            }
        }
        *slot = prev + (acc.x + acc.y);
    }
};
// logp[s] -= (df+1)/2 * sum_blocks rowpart[s][.]   (fp64 accumulation across blocks and chunks)
__global__ void reg_rowpart_acc_kernel(const float* __restrict__ rowpart, long long S, int n_blocks, int used_blocks,
                                       double scale, double* __restrict__ logp) {
    const int lane = threadIdx.x & 31;
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long s = warp; s < S; s += nwarps) {
        double a = 0.0;
        for (int b = lane; b < used_blocks; b += 32) a += (double)rowpart[s * n_blocks + b];
        a = warp_sum(a);
        if (lane == 0) logp[s] -= scale * a;
    }
}
// logp[s] = -|b_s|^2 / (2 sd^2);  gacc[s, j] = -b_sj / sd^2
__global__ void reg_init_kernel(const float* __restrict__ b, long long S, int D, double inv_var,
                                double* __restrict__ logp, double* __restrict__ gacc) {
    const int lane = threadIdx.x & 31;
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long s = warp; s < S; s += nwarps) {
        double a = 0.0;
        for (int j = lane; j < D; j += 32) {
            const double v = (double)b[s * D + j];
            a = fma(v, v, a);
            gacc[s * D + j] = -v * inv_var;
        }
        a = warp_sum(a);
        if (lane == 0) logp[s] = -0.5 * a * inv_var;
    }
}
// grad[e] = prior part + scale * sum_k acc[k][e]
__global__ void reg_final_kernel(const double* __restrict__ prior, const double* __restrict__ acc, int splits, long long total,
                                 double scale, float* __restrict__ dst) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        double a = 0.0;
        for (int k = 0; k < splits; ++k) a += acc[(size_t)k * total + e];
        dst[e] = (float)fma(scale, a, prior[e]);
    }
}

static long long reg_chunk_rows(long long S, long long N) {
    // Psi (S x chunk fp32) of at most 1 GiB per chunk.  Every chunk costs two kernel boundaries (pipeline fill,
    // tail imbalance, launch latency: ~20 us), so larger chunks are faster even though Psi then streams through HBM:
    // measured at S = 4096, N = 2^20: 14.9 ms with 16384-row chunks, 13.5 with 32768, 13.15 with 65536, 12.9 with
    // 262144.  VB_REG_CHUNK overrides (tuning).
    long long forced = 0;
    { const char* e = getenv("VB_REG_CHUNK"); if (e) forced = atoll(e); }
    long long c = forced > 0 ? forced : (1ll << 28) / (S > 0 ? S : 1);
    c = c / 128 * 128;
    if (c < 128) c = 128;
    if (c > 262144) c = 262144;
    if (c > N) c = (N + 31) / 32 * 32;
    return c;
}
// at most 256 k tiles (8192 rows) per TMEM accumulation chain of GEMM2: the truncating fp32 accumulate of the
// tensor core costs about 1e-9 of the gradient per row of the chain (measured 1.1e-5 at 10 000 rows); stated
// tolerance 2e-5
static long long reg_max_chain() {
    const char* e = getenv("VB_REG_CHAIN");        // tuning / experiments
    return e ? atoll(e) : 256;
}
#define kRegMaxChain reg_max_chain()
struct RegLayout { long long C; int splits; size_t off_xs, off_bs, off_psi, off_rowpart, off_parts, off_gacc, total; };
static RegLayout reg_layout(long long S, long long N, int D, bool have_planes) {
    RegLayout L;
    L.C = reg_chunk_rows(S, N);
    L.splits = plan_gemm(S, D, L.C, auto_splits(S, D, L.C, kRegMaxChain)).splits;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o += (bytes + 255) & ~(size_t)255; return at; };
    const bool resident = D <= kTResKTiles * kTBK;     // no pre-split operands needed
    L.off_xs = take(have_planes ? 0 : split_bytes(N, D));     // X planes attached to the model: nothing to stage
    L.off_bs = take(resident ? 0 : split_bytes(S, D));
    L.off_psi = take((size_t)S * L.C * 4);
    L.off_rowpart = take((size_t)S * (L.C / 32 + 1) * 4);
    L.off_parts = take((size_t)L.splits * S * D * 8);
    L.off_gacc = take((size_t)S * D * 8);
    L.total = o;
    return L;
}

size_t regression_tf32_workspace_bytes(long long n, long long n_rows, int dim, bool have_planes) {
    return reg_layout(n, n_rows, dim, have_planes).total + 512;
}

int regression_logp_grad_tf32(const ModelHost* m, const float* b, long long S, double* out_logp,
                              float* out_grad_or_null, void* ws, size_t ws_bytes, cudaStream_t st) {
    const int D = m->dim;
    const long long N = m->n_rows;
    if (D % 4) return VB_ERR_ALIGN;                       // TMA rows must be 16-byte multiples
    const RegLayout L = reg_layout(S, N, D, m->d_X_planes != nullptr);
    if (!ws || ws_bytes < L.total + 512) return VB_ERR_WORKSPACE;
    char* w = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
    float* psi = (float*)(w + L.off_psi);
    float* rowpart = (float*)(w + L.off_rowpart);
    double* parts = (double*)(w + L.off_parts);
    double* gacc = (double*)(w + L.off_gacc);
    const float* X = (const float*)m->d_X;
    const float* y = (const float*)m->d_y;
    // d <= 128: both GEMMs read X and the coefficient draws as plain fp32 and split them on chip; wider models fall
    // back to pre-split operands for GEMM1
    const bool resident = D <= kTResKTiles * kTBK;
    // VB_REG_RAW: bit 0 = GEMM1 reads X as plain fp32 and splits it in shared memory, bit 1 = same for GEMM2.
    // Default 0: measured on B200 the pre-split X planes are faster (14.0 vs 16.1 ms / step at N = 2^20, d = 128):
    // halving the L2 -> SM traffic does not pay for the extra shared-memory pass of the converter warps.
    int raw_mask = 0;
    { const char* e = getenv("VB_REG_RAW"); if (e) raw_mask = atoi(e); }
    const bool raw1 = resident && (raw_mask & 1), raw2 = (raw_mask & 2) != 0;
    SplitMat xs{}, bs{};
    int rc = VB_OK;
    if (m->d_X_planes != nullptr) {
        // the design matrix is model data: its tf32 planes were made once (vb_model_regression_attach_split)
        xs = SplitMat{m->d_X_planes, N, D, pad4(D), N * pad4(D)};
    } else if (!raw1 || !raw2) {
        rc = split_into(X, N, D, D, (float*)(w + L.off_xs), st, &xs);
        if (rc != VB_OK) return rc;
    }
    if (!resident) {
        rc = split_into(b, S, D, D, (float*)(w + L.off_bs), st, &bs);
        if (rc != VB_OK) return rc;
    }
    const double inv_var = 1.0 / (m->prior_sd * m->prior_sd);
    reg_init_kernel<<<grid_for(S * 32, 256, kNumSMs * 8), 256, 0, st>>>(b, S, D, inv_var, out_logp, gacc);
    VB_LAUNCH_CHECK();
    const int n_blocks = (int)(L.C / 32 + 1);
    const long long mn = S * D;
    VB_CUDA_TRY(cudaMemsetAsync(rowpart, 0, (size_t)S * n_blocks * 4, st));
    if (out_grad_or_null) VB_CUDA_TRY(cudaMemsetAsync(parts, 0, (size_t)L.splits * mn * 8, st));
    for (long long r0 = 0; r0 < N; r0 += L.C) {
        const long long cols = (r0 + L.C <= N) ? L.C : (N - r0);
        const SplitMat xraw{X + r0 * D, cols, D, D, 0};
        // GEMM1 + fused elementwise epilogue; Psi / c leaves through a TMA store as one fp32 plane (S, cols), ld = C
        PsiEpi e1;
        rc = make_raw_tmap(&e1.tmap, psi, S, cols, L.C, 32);
        if (rc != VB_OK) return rc;
        e1.y = y + r0; e1.rowpart = rowpart; e1.n_blocks = n_blocks; e1.inv_df = (float)(1.0 / m->df);
        // d <= 128: the 128-sample tile of draws stays in tensor memory while the CTA walks the chunk's row tiles
        const SplitMat xc{xs.base + r0 * xs.ld, cols, D, xs.ld, xs.plane_stride};
        if (raw1) {
            rc = launch_tf32x3<false, false, PsiEpi, 16, true, false, true>(SplitMat{b, S, D, D, 0}, xraw, plan_gemm(S, cols, D, 1), e1, st);
        } else if (resident) {
            rc = launch_tf32x3<false, false, PsiEpi, 16, true, false, false>(SplitMat{b, S, D, D, 0}, xc, plan_gemm(S, cols, D, 1), e1, st);
        } else {
            rc = launch_tf32x3<false, false, PsiEpi, 8>(bs, xc, plan_gemm(S, cols, D, 1), e1, st);
        }
        if (rc != VB_OK) return rc;
        if (out_grad_or_null) {
            const SplitMat ps{psi, S, cols, L.C, 0};                 // raw fp32: split in shared memory by the GEMM
            const GemmShape shape = plan_gemm(S, D, cols, auto_splits(S, D, cols, kRegMaxChain));
            if (shape.splits > L.splits) return VB_ERR_WORKSPACE;
            rc = raw2 ? launch_tf32x3<false, true, SplitKAccEpi, 4, false, true, true>(ps, xraw, shape, SplitKAccEpi{parts, mn, D}, st)
                      : launch_tf32x3<false, true, SplitKAccEpi, 4, false, true, false>(ps, xc, shape, SplitKAccEpi{parts, mn, D}, st);
            if (rc != VB_OK) return rc;
        }
    }
    // one reduction per call instead of one per chunk: log p from the per-(row, block) running sums, the gradient
    // from the per-split fp64 accumulators
    reg_rowpart_acc_kernel<<<grid_for(S * 32, 256, kNumSMs * 8), 256, 0, st>>>(
        rowpart, S, n_blocks, (int)(((N < L.C ? N : L.C) + 31) / 32), 0.5 * (m->df + 1.0), out_logp);
    VB_LAUNCH_CHECK();
    if (out_grad_or_null) {
        reg_final_kernel<<<grid_for(mn, 256, kNumSMs * 4), 256, 0, st>>>(gacc, parts, L.splits, mn, (m->df + 1.0) / m->df,
                                                                        out_grad_or_null);
        VB_LAUNCH_CHECK();
    }
    return VB_OK;
}

}  // namespace vb
