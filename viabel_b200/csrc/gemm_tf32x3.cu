// Host side of the 3xTF32 tcgen05 path: tensor-map creation, operand splitting, the plain GEMM
// entry (vb_gemm with dtype f32) and the fused pipelines built on the kernel in gemm_tf32x3.cuh.
#include "gemm_tf32x3.cuh"
#include "dense.cuh"

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

int tf32x3_num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0, v = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0)
            n = v;
        else
            n = kNumSMs;
    }
    return n;
}

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
struct AxpbyEpi {
    float* c; long long ldc; float alpha, beta;
    __device__ __forceinline__ void operator()(long long row, long long col0, float (&v)[32], int ncols, int) const {
        float* p = c + row * ldc + col0;
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (j < ncols) p[j] = alpha * v[j] + (beta == 0.f ? 0.f : beta * p[j]);
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
    float* x; long long ld; const double* mu; const float* chi2; float df;
    __device__ __forceinline__ void operator()(long long row, long long col0, float (&v)[32], int ncols, int) const {
        const float s = rsqrtf(chi2[row] / df);
        float* p = x + row * ld + col0;
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (j < ncols) p[j] = (float)mu[col0 + j] + v[j] * s;
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
    int rc = split_into(z, n, dim, dim, (float*)w, st, &zs);
    if (rc != VB_OK) return rc;
    rc = split_into(scale, dim, dim, dim, (float*)(w + split_bytes(n, dim)), st, &ss);
    if (rc != VB_OK) return rc;
    const GemmShape shape = plan_gemm(n, dim, dim, 1);
    return launch_tf32x3<false, true>(zs, ss, shape, FrSampleEpi{out_x, dim, mu, chi2, (float)df}, st);
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
    __device__ __forceinline__ void operator()(long long row, long long col0, float (&v)[32], int ncols, int) const {
        float* gp = g + row * ld + col0;
        const float* xp = x + row * ld + col0;
        float dot = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (j < ncols) {
                const float gv = bias[col0 + j] - v[j];
                gp[j] = gv;
                dot = fmaf(gv, xp[j] - (float)mean[col0 + j], dot);
            }
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

int gauss_logp_grad_f32(const double* mean, const double* prec, int dim, const float* x, long long n,
                        double* out_logp, float* out_grad_or_null, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (!ws || ws_bytes < gauss_f32_workspace_bytes(n, dim)) return VB_ERR_WORKSPACE;
    const int chunks = (dim + 31) / 32;
    char* w = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
    float* xs_buf = (float*)w;                  w += split_bytes(n, dim);
    float* ps_buf = (float*)w;                  w += split_bytes(dim, dim);
    float* p32 = (float*)w;                     w += ((size_t)dim * dim * 4 + 255) & ~(size_t)255;
    float* bias = (float*)w;                    w += ((size_t)dim * 4 + 255) & ~(size_t)255;
    float* rowpart = (float*)w;                 w += ((size_t)n * chunks * 4 + 255) & ~(size_t)255;
    float* g = out_grad_or_null ? out_grad_or_null : (float*)w;
    gauss_prep_kernel<<<grid_for((long long)dim * dim, 256, kNumSMs * 2), 256, 0, st>>>(mean, prec, dim, p32, bias);
    VB_LAUNCH_CHECK();
    SplitMat xs, ps;
    int rc = split_into(x, n, dim, dim, xs_buf, st, &xs);
    if (rc != VB_OK) return rc;
    rc = split_into(p32, dim, dim, dim, ps_buf, st, &ps);
    if (rc != VB_OK) return rc;
    const GemmShape shape = plan_gemm(n, dim, dim, 1);
    rc = launch_tf32x3<false, true>(xs, ps, shape, GaussEpi{g, dim, bias, x, mean, rowpart, chunks}, st);
    if (rc != VB_OK) return rc;
    rowpart_sum_kernel<<<grid_for(n, 256, kNumSMs * 8), 256, 0, st>>>(rowpart, n, chunks, 0.5, out_logp);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ------------------------------------------------------------------------------------------ G_S = Ztilde' diag(w) G
// A(a, i) = z[i, a] (MN-major), B(i, b) = (w_i / s_i) g[i, b] (MN-major), K = n, split-K partials in fp32,
// reduced in fp64.
__global__ void split_rowscaled_kernel(const float* __restrict__ src, long long rows, long long cols,
                                       const double* __restrict__ w, const float* __restrict__ chi2, float df,
                                       float* __restrict__ dst, long long ldd, long long plane) {
    const long long total = rows * ldd;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long r = e / ldd, c = e - r * ldd;
        float hi = 0.f, lo = 0.f;
        if (c < cols) {
            const float sc = (float)(w[r] / sqrt((double)chi2[r] / (double)df));
            tc::split_tf32(sc * src[r * cols + c], hi, lo);
        }
        dst[e] = hi;
        dst[e + plane] = lo;
    }
}
struct SplitKEpi {
    float* parts; long long mn; long long ld;
    __device__ __forceinline__ void operator()(long long row, long long col0, float (&v)[32], int ncols, int split) const {
        float* p = parts + (size_t)split * mn + row * ld + col0;
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (j < ncols) p[j] = v[j];
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
    return 2 * split_bytes(n, dim) + (((size_t)splits * dim * dim * 4 + 255) & ~(size_t)255) + 512;
}

int fr_weighted_GS_tf32(double df, long long n, int dim, const float* z, const float* chi2, const double* w,
                        const float* grad, double* out_GS, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (!ws || ws_bytes < fr_weighted_tf32_workspace_bytes(n, dim)) return VB_ERR_WORKSPACE;
    char* wp = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
    float* zs_buf = (float*)wp;     wp += split_bytes(n, dim);
    float* gs_buf = (float*)wp;     wp += split_bytes(n, dim);
    float* parts = (float*)wp;
    SplitMat zs;
    int rc = split_into(z, n, dim, dim, zs_buf, st, &zs);
    if (rc != VB_OK) return rc;
    const long long ldd = pad4(dim);
    split_rowscaled_kernel<<<grid_for(n * ldd, 256, kNumSMs * 8), 256, 0, st>>>(grad, n, dim, w, chi2, (float)df,
                                                                               gs_buf, ldd, n * ldd);
    VB_LAUNCH_CHECK();
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
// ------------------------------------------------------------------------------------------ not yet on the tensor path
size_t regression_tf32_workspace_bytes(long long, long long, int) { return 0; }
int regression_logp_grad_tf32(const ModelHost*, const float*, long long, double*, float*, void*, size_t,
                              cudaStream_t) { return VB_ERR_UNSUPPORTED; }

}  // namespace vb
