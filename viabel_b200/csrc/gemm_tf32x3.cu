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

// ------------------------------------------------------------------------------------------ not yet on the tensor path
int fr_sample_tf32(const double*, const float*, double, long long, int, const float*, const float*, float*, void*,
                   size_t, cudaStream_t) { return VB_ERR_UNSUPPORTED; }
int gauss_logp_grad_f32(const double*, const double*, int, const float*, long long, double*, float*, void*, size_t,
                        cudaStream_t) { return VB_ERR_UNSUPPORTED; }
size_t regression_tf32_workspace_bytes(long long, long long, int) { return 0; }
int regression_logp_grad_tf32(const ModelHost*, const float*, long long, double*, float*, void*, size_t,
                              cudaStream_t) { return VB_ERR_UNSUPPORTED; }

}  // namespace vb
