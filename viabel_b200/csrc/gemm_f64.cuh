// fp64 SIMT GEMM with functor operands — the parity-grade dense path (rtol 1e-10 vs the
// oracle).  C(i, j) = sum_k A(i, k) B(k, j) over k in [k0, k1); operands are read through
// functors so prologues (centering, row scaling, dtype promotion) and epilogues (bias, row
// scaling, fused elementwise maps, split-K partial stores) fuse without extra passes.
//
// Tile 128 x 64 x 16, 256 threads, 8 x 4 micro-tile, register prefetch of the next k-slab.
// DFMA peak on B200 is ~37 TFLOP/s; this kernel is the validation instrument, the
// throughput path is the 3xTF32 tcgen05 kernel in gemm_tf32x3.cu.
#pragma once
#include "common.cuh"

namespace vb {

constexpr int kGBM = 128, kGBN = 64, kGBK = 16, kGThreads = 256;

// Loader contract:  double operator()(long long row, long long k) const   (A)   /   (k, col) (B)
//                   static constexpr bool kContigCol — true when the SECOND index is the contiguous one
//                   (only steers the thread->element mapping of the tile loads for coalescing)
// Epilogue contract: void operator()(long long i, long long j, double acc, int split) const
template <class ALoad, class BLoad, class Epi>
__global__ void __launch_bounds__(kGThreads)
gemm_f64_kernel(long long M, long long N, long long K, long long k_per_split, ALoad A, BLoad B, Epi epi) {
    __shared__ double As[kGBK][kGBM + 4];
    __shared__ double Bs[kGBK][kGBN + 4];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;           // 16 x 16 threads; thread owns rows ty*8.., cols tx*4..
    const long long i0 = (long long)blockIdx.y * kGBM, j0 = (long long)blockIdx.x * kGBN;
    const int split = blockIdx.z;
    const long long kb = (long long)split * k_per_split;
    long long ke = kb + k_per_split;
    if (ke > K) ke = K;

    double acc[8][4];
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;

    // per-thread staging: A tile 128x16 = 2048 elems -> 8 per thread; B tile 16x64 = 1024 -> 4 per thread
    double ra[8], rb[4];
    auto load_tiles = [&](long long kk0) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int idx = tid + e * kGThreads;
            int r, c;
            if (ALoad::kContigCol) { r = idx / kGBK; c = idx % kGBK; } else { r = idx % kGBM; c = idx / kGBM; }
            const long long gi = i0 + r, gk = kk0 + c;
            ra[e] = (gi < M && gk < ke) ? A(gi, gk) : 0.0;
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int idx = tid + e * kGThreads;
            int r, c;
            if (!BLoad::kContigCol) { c = idx / kGBK; r = idx % kGBK; } else { c = idx % kGBN; r = idx / kGBN; }
            const long long gk = kk0 + r, gj = j0 + c;
            rb[e] = (gk < ke && gj < N) ? B(gk, gj) : 0.0;
        }
    };
    auto store_tiles = [&]() {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int idx = tid + e * kGThreads;
            int r, c;
            if (ALoad::kContigCol) { r = idx / kGBK; c = idx % kGBK; } else { r = idx % kGBM; c = idx / kGBM; }
            As[c][r] = ra[e];
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int idx = tid + e * kGThreads;
            int r, c;
            if (!BLoad::kContigCol) { c = idx / kGBK; r = idx % kGBK; } else { c = idx % kGBN; r = idx / kGBN; }
            Bs[r][c] = rb[e];
        }
    };

    if (kb < ke) {
        load_tiles(kb);
        for (long long kk0 = kb; kk0 < ke; kk0 += kGBK) {
            __syncthreads();
            store_tiles();
            __syncthreads();
            if (kk0 + kGBK < ke) load_tiles(kk0 + kGBK);
#pragma unroll
            for (int k = 0; k < kGBK; ++k) {
                double av[8], bv[4];
#pragma unroll
                for (int a = 0; a < 8; ++a) av[a] = As[k][ty * 8 + a];
#pragma unroll
                for (int b = 0; b < 4; ++b) bv[b] = Bs[k][tx * 4 + b];
#pragma unroll
                for (int a = 0; a < 8; ++a)
#pragma unroll
                    for (int b = 0; b < 4; ++b) acc[a][b] = fma(av[a], bv[b], acc[a][b]);
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        const long long gi = i0 + ty * 8 + a;
        if (gi >= M) continue;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const long long gj = j0 + tx * 4 + b;
            if (gj < N) epi(gi, gj, acc[a][b], split);
        }
    }
}

template <class ALoad, class BLoad, class Epi>
static int launch_gemm_f64(long long M, long long N, long long K, int splits, const ALoad& A, const BLoad& B,
                           const Epi& epi, cudaStream_t st) {
    if (M < 1 || N < 1 || K < 0 || splits < 1) return VB_ERR_ARG;
    long long kps = (K + splits - 1) / splits;
    kps = (kps + kGBK - 1) / kGBK * kGBK;
    if (kps < kGBK) kps = kGBK;
    dim3 grid((unsigned)((N + kGBN - 1) / kGBN), (unsigned)((M + kGBM - 1) / kGBM), (unsigned)splits);
    if (grid.y > 65535u) return VB_ERR_UNSUPPORTED;
    gemm_f64_kernel<ALoad, BLoad, Epi><<<grid, kGThreads, 0, st>>>(M, N, K, kps, A, B, epi);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? VB_OK : (int)e;
}

// ---- stock functors --------------------------------------------------------------------
template <typename T>
struct RowMajor {                       // element (r, c) = p[r*ld + c]
    static constexpr bool kContigCol = true;
    const T* p; long long ld;
    __device__ __forceinline__ double operator()(long long r, long long c) const { return (double)p[r * ld + c]; }
};
template <typename T>
struct ColMajor {                       // element (r, c) = p[c*ld + r]  (a row-major matrix read transposed)
    static constexpr bool kContigCol = false;
    const T* p; long long ld;
    __device__ __forceinline__ double operator()(long long r, long long c) const { return (double)p[c * ld + r]; }
};

}  // namespace vb
