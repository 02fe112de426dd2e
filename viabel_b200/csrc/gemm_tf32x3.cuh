// 3xTF32 GEMM on tcgen05 (sm_100a): C = A B with fp32 operands carried as tf32 (hi, lo) pairs,
//   A B ~= A_hi B_hi + A_hi B_lo + A_lo B_hi        (lo*lo dropped: 2^-22 relative)
// accumulated in fp32 in tensor memory.  Operands are "split" matrices (see tc05.cuh):
// producers write the hi/lo planes once, the GEMM streams them with TMA (SWIZZLE_128B) into a
// 3-stage shared-memory ring, a single thread issues tcgen05.mma (kind::tf32, 128x128x8), and four
// epilogue warps drain double-buffered TMEM accumulators through a fused epilogue functor.
//
// Warp roles (192 threads): warp 0 TMA producer, warp 1 TMEM owner + MMA issuer, warps 2-5 epilogue.
// Persistent CTAs walk (m tile, n tile, k split) work items round-robin.
//
// Layouts: A is either K-major (rows = M, cols = K) or MN-major (rows = K, cols = M); same for B
// with N.  K-major tiles are one (32 x 128 x 2) box; MN-major tiles are four (32 x 32 x 2) boxes.
#pragma once
#include "tc05.cuh"

namespace vb {

constexpr int kTBM = 128, kTBN = 128, kTBK = 32, kTStages = 3;
constexpr int kTThreads = 192;
constexpr uint32_t kTPlaneBytes = 128 * 32 * 4;             // one 128 x 32 fp32 tile
constexpr uint32_t kTOperandBytes = 2 * kTPlaneBytes;       // hi + lo
constexpr uint32_t kTStageBytes = 2 * kTOperandBytes;       // A + B
constexpr size_t kTSmemBytes = (size_t)kTStages * kTStageBytes + 1024 /*align*/ + 256 /*barriers*/;

struct SplitMat {
    const float* base;          // plane 0 (hi); plane 1 (lo) at base + plane_stride
    long long rows, cols, ld, plane_stride;
};

struct GemmShape {
    long long M, N, K;
    int splits;                 // k splits (>= 1), every split non-empty
    int k_tiles_per_split;
};

// Smem offsets of the hi / lo planes and descriptor strides for an operand tile
template <bool MN_MAJOR>
struct OperandLayout {
    // K-major : [plane][128 rows][32 k]            hi +0, lo +16384, SBO 1024 (8 rows), k-step 32 B,
    //           SWIZZLE_128B
    // MN-major: [block(4)][plane][32 k][32 mn]     hi +0, lo +4096,  LBO 8192 (next 32 mn),
    //           SBO 512 (next 4 k rows), k-step 1024 B, SWIZZLE_128B_BASE32B (tf32 MN-major rule)
    static constexpr uint32_t lo_offset = MN_MAJOR ? 4096u : kTPlaneBytes;
    static constexpr uint32_t lbo = MN_MAJOR ? 8192u : 16u;
    static constexpr uint32_t sbo = MN_MAJOR ? 512u : 1024u;
    static constexpr uint32_t layout_type = MN_MAJOR ? 1u : 2u;
    static constexpr uint32_t k_step = MN_MAJOR ? 1024u : 32u;     // per UMMA_K = 8 fp32
};

// EW = number of epilogue warps (4, 8 or 16): warp (q, h) drains TMEM lane quarter q = warp % 4 and the
// column blocks h, h + EW/4, ... of the 128-column accumulator.  Heavy epilogues (transcendentals per
// element) need 16 warps to hide their latency; a plain store is fine with 4.
template <bool A_MN, bool B_MN, class Epi, int EW = 4>
__global__ void __launch_bounds__(64 + 32 * EW, 1)
gemm_tf32x3_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                   const GemmShape shape, const Epi epi) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) &
                                                           ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)kTStages * kTStageBytes);
    uint64_t* full_bar = bars;                      // [stages]
    uint64_t* empty_bar = bars + kTStages;          // [stages]
    uint64_t* tmem_full = bars + 2 * kTStages;      // [2]
    uint64_t* tmem_empty = bars + 2 * kTStages + 2; // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * kTStages + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long m_tiles = (shape.M + kTBM - 1) / kTBM, n_tiles = (shape.N + kTBN - 1) / kTBN;
    const long long total = m_tiles * n_tiles * shape.splits;
    const long long k_tiles_total = (shape.K + kTBK - 1) / kTBK;

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tmap_a);
        tc::prefetch_tmap(&tmap_b);
#pragma unroll
        for (int s = 0; s < kTStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&tmem_full[0], 1); mbar_init(&tmem_full[1], 1);
        mbar_init(&tmem_empty[0], 32 * EW); mbar_init(&tmem_empty[1], 32 * EW);
        mbar_fence_init();
    }
    if (warp == 1) tc::tmem_alloc<256>(tmem_ptr);      // 2 accumulator stages x 128 columns
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ===================================================== TMA producer (whole warp loops, lane 0 issues)
        int stage = 0;
        uint32_t phase = 0;
        for (long long w = blockIdx.x; w < total; w += gridDim.x) {
            const long long tile = w % (m_tiles * n_tiles);
            const int split = (int)(w / (m_tiles * n_tiles));
            const int n0 = (int)(tile % n_tiles) * kTBN, m0 = (int)(tile / n_tiles) * kTBM;   // n fastest: CTAs that run together share the A tile in L2
            const long long kt0 = (long long)split * shape.k_tiles_per_split;
            long long kt1 = kt0 + shape.k_tiles_per_split;
            if (kt1 > k_tiles_total) kt1 = k_tiles_total;
            for (long long kt = kt0; kt < kt1; ++kt) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                if (lane == 0) {
                    unsigned char* sa = smem + (size_t)stage * kTStageBytes;
                    unsigned char* sb = sa + kTOperandBytes;
                    mbar_expect_tx(&full_bar[stage], kTStageBytes);
                    const int k0 = (int)kt * kTBK;
                    if (A_MN) {
#pragma unroll
                        for (int b = 0; b < 4; ++b) tc::tma_load_3d(sa + b * 8192, &tmap_a, &full_bar[stage], m0 + 32 * b, k0, 0);
                    } else {
                        tc::tma_load_3d(sa, &tmap_a, &full_bar[stage], k0, m0, 0);
                    }
                    if (B_MN) {
#pragma unroll
                        for (int b = 0; b < 4; ++b) tc::tma_load_3d(sb + b * 8192, &tmap_b, &full_bar[stage], n0 + 32 * b, k0, 0);
                    } else {
                        tc::tma_load_3d(sb, &tmap_b, &full_bar[stage], k0, n0, 0);
                    }
                }
                __syncwarp();
                if (++stage == kTStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (whole warp loops, lane 0 issues)
        constexpr uint32_t idesc = tc::idesc_tf32(kTBM, kTBN, A_MN, B_MN);
        using LA = OperandLayout<A_MN>;
        using LB = OperandLayout<B_MN>;
        int stage = 0;
        uint32_t phase = 0;
        int it = 0;
        for (long long w = blockIdx.x; w < total; w += gridDim.x, ++it) {
            const int split = (int)(w / (m_tiles * n_tiles));
            const long long kt0 = (long long)split * shape.k_tiles_per_split;
            long long kt1 = kt0 + shape.k_tiles_per_split;
            if (kt1 > k_tiles_total) kt1 = k_tiles_total;
            const int as = it & 1;
            mbar_wait(&tmem_empty[as], (uint32_t)(((it >> 1) & 1) ^ 1));
            tc::fence_after_sync();
            const uint32_t d_addr = tmem_base + (uint32_t)as * kTBN;
            bool first = true;
            for (long long kt = kt0; kt < kt1; ++kt) {
                mbar_wait(&full_bar[stage], phase);
                tc::fence_after_sync();
                if (lane == 0) {
                    const uint32_t a_addr = smem_u32(smem + (size_t)stage * kTStageBytes);
                    const uint32_t b_addr = a_addr + kTOperandBytes;
#pragma unroll
                    for (int k = 0; k < kTBK / 8; ++k) {
                        const uint64_t a_hi = tc::smem_desc_sw128(a_addr + k * LA::k_step, LA::lbo, LA::sbo, LA::layout_type);
                        const uint64_t a_lo = tc::smem_desc_sw128(a_addr + LA::lo_offset + k * LA::k_step, LA::lbo, LA::sbo, LA::layout_type);
                        const uint64_t b_hi = tc::smem_desc_sw128(b_addr + k * LB::k_step, LB::lbo, LB::sbo, LB::layout_type);
                        const uint64_t b_lo = tc::smem_desc_sw128(b_addr + LB::lo_offset + k * LB::k_step, LB::lbo, LB::sbo, LB::layout_type);
                        tc::mma_tf32(d_addr, a_lo, b_hi, idesc, !(first && k == 0));   // small terms first
                        tc::mma_tf32(d_addr, a_hi, b_lo, idesc, true);
                        tc::mma_tf32(d_addr, a_hi, b_hi, idesc, true);
                    }
                    tc::mma_commit(&empty_bar[stage]);        // frees the smem slot when these MMAs retire
                }
                __syncwarp();
                first = false;
                if (++stage == kTStages) { stage = 0; phase ^= 1; }
            }
            if (lane == 0) tc::mma_commit(&tmem_full[as]);    // accumulator ready for the epilogue
            __syncwarp();
        }
    } else {
        // ===================================================== epilogue (warps 2..5)
        const int q = warp & 3;                      // TMEM lane quarter this warp may access
        const int h = (warp - 2) >> 2;               // which column blocks of the quarter this warp takes
        int it = 0;
        for (long long w = blockIdx.x; w < total; w += gridDim.x, ++it) {
            const long long tile = w % (m_tiles * n_tiles);
            const int split = (int)(w / (m_tiles * n_tiles));
            const long long n0 = (tile % n_tiles) * kTBN, m0 = (tile / n_tiles) * kTBM;
            const int as = it & 1;
            mbar_wait(&tmem_full[as], (uint32_t)((it >> 1) & 1));
            tc::fence_after_sync();
            const long long row = m0 + q * 32 + lane;
#pragma unroll 1
            for (int c = h; c < kTBN / 32; c += EW / 4) {
                const long long col0 = n0 + c * 32;
                long long ncols = shape.N - col0;
                if (ncols > 32) ncols = 32;
                const bool live = row < shape.M && ncols > 0;
                // operands the epilogue needs from global memory are requested before the TMEM read
                typename Epi::Pre pre;
                if (live) pre = epi.prefetch(row, col0, (int)ncols);
                float v[32];
                tc::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * kTBN + c * 32), v);
                if (live) epi(row, col0, v, (int)ncols, split, pre);
            }
            tc::fence_before_sync();
            mbar_arrive(&tmem_empty[as]);
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc<256>(tmem_base);
}

// choose k splits so that the persistent grid has about one work item per SM
inline GemmShape plan_gemm(long long M, long long N, long long K, int want_splits) {
    GemmShape s;
    s.M = M; s.N = N; s.K = K;
    const long long kt = (K + kTBK - 1) / kTBK;
    long long sp = want_splits < 1 ? 1 : want_splits;
    if (sp > kt) sp = kt;
    const long long per = (kt + sp - 1) / sp;
    s.k_tiles_per_split = (int)per;
    s.splits = (int)((kt + per - 1) / per);
    return s;
}
inline int auto_splits(long long M, long long N, long long K) {
    const long long tiles = ((M + kTBM - 1) / kTBM) * ((N + kTBN - 1) / kTBN);
    if (tiles >= kNumSMs) return 1;
    long long sp = kNumSMs / tiles;
    const long long kt = (K + kTBK - 1) / kTBK;
    if (sp > kt / 8) sp = kt / 8;          // keep >= 8 k tiles per split
    return sp < 1 ? 1 : (int)sp;
}

int tf32x3_num_sms();

template <bool A_MN, bool B_MN, class Epi, int EW = 4>
static int launch_tf32x3(const SplitMat& A, const SplitMat& B, const GemmShape& shape, const Epi& epi,
                         cudaStream_t st) {
    CUtensorMap ta, tb;
    int rc = make_split_tmap(&ta, A.base, A.rows, A.cols, A.ld, A.plane_stride, A_MN ? kTBK : kTBM, A_MN);
    if (rc != VB_OK) return rc;
    rc = make_split_tmap(&tb, B.base, B.rows, B.cols, B.ld, B.plane_stride, B_MN ? kTBK : kTBN, B_MN);
    if (rc != VB_OK) return rc;
    auto kern = gemm_tf32x3_kernel<A_MN, B_MN, Epi, EW>;
    static bool configured = false;
    if (!configured) {
        VB_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTSmemBytes));
        configured = true;
    }
    const long long total = ((shape.M + kTBM - 1) / kTBM) * ((shape.N + kTBN - 1) / kTBN) * shape.splits;
    long long grid = tf32x3_num_sms();
    if (grid > total) grid = total;
    kern<<<(unsigned)grid, 64 + 32 * EW, kTSmemBytes, st>>>(ta, tb, shape, epi);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

}  // namespace vb
