// 3xTF32 GEMM on tcgen05 (sm_100a): C = A B with fp32 operands carried as tf32 (hi, lo) pairs,
//   A B ~= A_hi B_hi + A_hi B_lo + A_lo B_hi        (lo*lo dropped: 2^-22 relative)
// accumulated in fp32 in tensor memory.  Operands are "split" matrices (see tc05.cuh):
// producers write the hi/lo planes once, the GEMM streams them with TMA (SWIZZLE_128B) into a
// shared-memory ring, a single thread issues tcgen05.mma (kind::tf32, 128x128x8), and the
// epilogue warps drain double-buffered TMEM accumulators through a fused epilogue functor.
//
// Warp roles: warp 0 TMA producer, warp 1 TMEM owner + MMA issuer, warps 2 .. 2+EW-1 epilogue,
// (A_RAW only) four more warps that split a raw fp32 A tile into its hi / lo planes in shared memory.
// Persistent CTAs walk (m tile, n tile, k split) work items.
//
// Layouts: A is either K-major (rows = M, cols = K) or MN-major (rows = K, cols = M); same for B
// with N.  K-major tiles are one (32 x 128 x 2) box; MN-major tiles are four (32 x 32 x 2) boxes.
//
// Variants (template flags) -- both take the A operand of tcgen05.mma from TENSOR MEMORY: three tf32 MMAs per
// product re-read both operands from shared memory (96 KB per 32-wide k tile against 128 B/clk of shared-memory
// bandwidth: ~70 % of it for the MMAs alone), so anything else that touches shared memory stalls the tensor pipe.
//   A_RES  (K <= 128, K-major A, splits == 1): the whole K extent of the A tile (hi and lo: 256 TMEM columns) is
//          written to tensor memory once by four loader warps and stays there while the CTA walks a contiguous
//          run of n tiles; only B is streamed (L2 -> SM traffic per tile halves, too).
//   A_RAW  (K-major A): A is a plain fp32 matrix (one plane in HBM, half the traffic of a split operand); the raw
//          tile is loaded by TMA, split into hi / lo by four converter warps and written to a TMEM stage.
//   Epi::kStaged: the epilogue writes its 32x32 output blocks to a per-warp swizzled shared-memory buffer and
//          stores them with TMA (full lines, no LSU pressure) instead of 32 scattered sectors per warp store.
#pragma once
#include <cstdlib>
#include <type_traits>
#include "tc05.cuh"

namespace vb {

constexpr int kTBM = 128, kTBN = 128, kTBK = 32, kTStages = 3;
constexpr uint32_t kTPlaneBytes = 128 * 32 * 4;             // one 128 x 32 fp32 tile
constexpr uint32_t kTOperandBytes = 2 * kTPlaneBytes;       // hi + lo
constexpr uint32_t kTStageBytes = 2 * kTOperandBytes;       // A + B
constexpr int kTResKTiles = 4;
constexpr size_t kTResABytes = (size_t)kTResKTiles * kTOperandBytes;
constexpr uint32_t kTStagingBytes = 32 * 32 * 4;            // one 32 x 32 fp32 block per epilogue warp

struct SplitMat {
    const float* base;          // plane 0 (hi); plane 1 (lo) at base + plane_stride (ignored for a raw operand)
    long long rows, cols, ld, plane_stride;
};

struct GemmShape {
    long long M, N, K;
    int splits;                 // k splits (>= 1), every split non-empty
    int k_tiles_per_split;
    int debug = 0;              // experiments only (VB_GEMM_DEBUG): 1 = epilogue body skipped, 2 = MMAs skipped
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

template <class E, class = void> struct epi_colvec : std::false_type {};
template <class E> struct epi_colvec<E, std::void_t<decltype(E::kColVec)>> : std::bool_constant<E::kColVec> {};
template <class E, class = void> struct epi_staged : std::false_type {};
template <class E> struct epi_staged<E, std::void_t<decltype(E::kStaged)>> : std::bool_constant<E::kStaged> {};

// Shared-memory / tensor-memory plan of one instantiation
template <class Epi, int EW, bool A_RES, bool A_RAW, bool B_RAW>
struct GemmCfg {
    static constexpr bool kStaged = epi_staged<Epi>::value;
    static constexpr bool kColVec = epi_colvec<Epi>::value;     // the epilogue reads a per-column vector (staged per tile)
    static constexpr bool kATmem = A_RES || A_RAW;                 // A operand read from tensor memory
    static constexpr bool kConvert = A_RAW || B_RAW;               // auxiliary warps touch every k tile
    // ring slot: [A hi+lo | B hi+lo], or [B] (A_RES), or [A raw | B] (A_RAW); a raw B tile lands in the hi part
    // of its [hi | lo] region and is split in place
    static constexpr size_t kBOff = A_RES ? 0 : (A_RAW ? kTPlaneBytes : kTOperandBytes);
    static constexpr size_t kSlot = kBOff + kTOperandBytes;
    static constexpr uint32_t kSlotTx = (uint32_t)kBOff + (B_RAW ? kTPlaneBytes : kTOperandBytes);   // bytes TMA delivers per slot
    static constexpr size_t kColVecBytes = kColVec ? 2 * kTBN * 4 : 0;          // one 128-float segment per accumulator stage
    static constexpr size_t kStagingBytes = (kStaged ? (size_t)EW * kTStagingBytes : 0) + kColVecBytes;
    // as many ring slots as fit (a 128-wide K is only four k tiles: the ring must hold more than one output tile's
    // worth of operands or the loads of the next tile cannot start before the current one retires)
    static constexpr int kFit = (int)((232448 - 1344 - kStagingBytes) / kSlot);
    static constexpr int kStages = kATmem ? (kFit > 8 ? 8 : kFit) : kTStages;
    static constexpr size_t kStagingOff = (size_t)kStages * kSlot;
    static constexpr size_t kBarOff = kStagingOff + kStagingBytes;
    static constexpr size_t kSmemBytes = kBarOff + 320 /*barriers*/ + 1024 /*align*/;
    static constexpr int kThreads = 64 + 32 * EW + (kATmem ? 128 : 0);
    static constexpr uint32_t kTmemCols = kATmem ? 512 : 256;      // accumulators 0..255, A operand 256..511
    static constexpr uint32_t kTmemA = 256;
    static_assert(!A_RAW || kStages <= 4, "TMEM holds four staged A tiles");
    static_assert(kSmemBytes <= 232448, "exceeds the 227 KB of shared memory a CTA may use");
};

// The A operand as the loader warps of the A_RES variant read it (plain fp32 in global memory, split in registers)
struct AGlobal {
    const float* base; long long ld;
};

// EW = number of epilogue warps (4, 8 or 16): warp (q, h) drains TMEM lane quarter q = warp % 4 and the
// column blocks h, h + EW/4, ... of the 128-column accumulator.
template <bool A_MN, bool B_MN, class Epi, int EW = 4, bool A_RES = false, bool A_RAW = false, bool B_RAW = false>
__global__ void __launch_bounds__(GemmCfg<Epi, EW, A_RES, A_RAW, B_RAW>::kThreads, 1)
gemm_tf32x3_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                   const GemmShape shape, const AGlobal ag, const __grid_constant__ Epi epi) {
    using Cfg = GemmCfg<Epi, EW, A_RES, A_RAW, B_RAW>;
    static_assert(!(A_RES && A_MN) && !(A_RAW && A_MN) && !(A_RES && A_RAW) && !(B_RAW && !Cfg::kATmem),
                  "unsupported variant combination");
    constexpr int kStages = Cfg::kStages;
    constexpr size_t kSlot = Cfg::kSlot;
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) &
                                                           ~(uintptr_t)1023);
    unsigned char* ring = smem;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kBarOff);
    uint64_t* full_bar = bars;                      // [8] TMA bytes of the slot landed
    uint64_t* empty_bar = bars + 8;                 // [8] MMAs reading the slot (and its TMEM A stage) retired
    uint64_t* conv_bar = bars + 16;                 // [8] raw operands of the slot are split (A in its TMEM stage)
    uint64_t* tmem_full = bars + 24;                // [2]
    uint64_t* tmem_empty = bars + 26;               // [2]
    uint64_t* a_full = bars + 28;                   // A_RES: resident A tile written / all MMAs on it retired
    uint64_t* a_empty = bars + 29;
    uint64_t* vec_full = bars + 30;                 // [2] kColVec: the tile's column-vector segment landed
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 32);
    float* colvec_sm = reinterpret_cast<float*>(smem + Cfg::kBarOff - Cfg::kColVecBytes);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long m_tiles = (shape.M + kTBM - 1) / kTBM, n_tiles = (shape.N + kTBN - 1) / kTBN;
    const long long total = m_tiles * n_tiles * shape.splits;
    const long long k_tiles_total = (shape.K + kTBK - 1) / kTBK;
    // work items: round-robin over the grid, or (A_RES) one contiguous run of n tiles per CTA
    const long long w_begin = A_RES ? total * blockIdx.x / gridDim.x : blockIdx.x;
    const long long w_end = A_RES ? total * (blockIdx.x + 1) / gridDim.x : total;
    const long long w_step = A_RES ? 1 : gridDim.x;

    if (threadIdx.x == 0) {
        if (!Cfg::kATmem || A_RAW) tc::prefetch_tmap(&tmap_a);
        tc::prefetch_tmap(&tmap_b);
#pragma unroll
        for (int s = 0; s < 8; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); mbar_init(&conv_bar[s], 128); }
        mbar_init(&tmem_full[0], 1); mbar_init(&tmem_full[1], 1);
        mbar_init(&tmem_empty[0], 32 * EW); mbar_init(&tmem_empty[1], 32 * EW);
        mbar_init(a_full, 128); mbar_init(a_empty, 1);
        mbar_init(&vec_full[0], 1); mbar_init(&vec_full[1], 1);
        mbar_fence_init();
    }
    if (warp == 1) tc::tmem_alloc<Cfg::kTmemCols>(tmem_ptr);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ===================================================== TMA producer (whole warp loops, lane 0 issues)
        int stage = 0;
        uint32_t phase = 0;
        for (long long w = w_begin; w < w_end; w += w_step) {
            const long long tile = w % (m_tiles * n_tiles);
            const int split = (int)(w / (m_tiles * n_tiles));
            // n fastest: CTAs that run together share the A tile in L2
            const int n0 = (int)(tile % n_tiles) * kTBN, m0 = (int)(tile / n_tiles) * kTBM;
            const long long kt0 = (long long)split * shape.k_tiles_per_split;
            long long kt1 = kt0 + shape.k_tiles_per_split;
            if (kt1 > k_tiles_total) kt1 = k_tiles_total;
            for (long long kt = kt0; kt < kt1; ++kt) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                if (lane == 0) {
                    unsigned char* sa = ring + (size_t)stage * kSlot;
                    unsigned char* sb = sa + Cfg::kBOff;
                    mbar_expect_tx(&full_bar[stage], Cfg::kSlotTx);
                    const int k0 = (int)kt * kTBK;
                    if (A_RES) {
                    } else if (A_RAW) {
                        tc::tma_load_2d(sa, &tmap_a, &full_bar[stage], k0, m0);
                    } else if (A_MN) {
#pragma unroll
                        for (int b = 0; b < 4; ++b) tc::tma_load_3d(sa + b * 8192, &tmap_a, &full_bar[stage], m0 + 32 * b, k0, 0);
                    } else {
                        tc::tma_load_3d(sa, &tmap_a, &full_bar[stage], k0, m0, 0);
                    }
                    if (B_RAW && B_MN) {
#pragma unroll
                        for (int b = 0; b < 4; ++b) tc::tma_load_2d(sb + b * 8192, &tmap_b, &full_bar[stage], n0 + 32 * b, k0);
                    } else if (B_RAW) {
                        tc::tma_load_2d(sb, &tmap_b, &full_bar[stage], k0, n0);
                    } else if (B_MN) {
#pragma unroll
                        for (int b = 0; b < 4; ++b) tc::tma_load_3d(sb + b * 8192, &tmap_b, &full_bar[stage], n0 + 32 * b, k0, 0);
                    } else {
                        tc::tma_load_3d(sb, &tmap_b, &full_bar[stage], k0, n0, 0);
                    }
                }
                __syncwarp();
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (whole warp loops, lane 0 issues)
        constexpr uint32_t idesc = tc::idesc_tf32(kTBM, kTBN, A_MN, B_MN);
        using LA = OperandLayout<A_MN>;
        using LB = OperandLayout<B_MN>;
        int stage = 0;
        uint32_t phase = 0, a_phase = 0;
        int it = 0;
        long long cur_m = -1;
        const uint32_t tmem_u = __reduce_or_sync(0xffffffffu, tmem_base);      // same value, now provably uniform
        const uint32_t ring_u = smem_u32(ring);
        const uint32_t bars_u = smem_u32(bars);
        for (long long w = w_begin; w < w_end; w += w_step, ++it) {
            const int split = (int)(w / (m_tiles * n_tiles));
            const long long kt0 = (long long)split * shape.k_tiles_per_split;
            long long kt1 = kt0 + shape.k_tiles_per_split;
            if (kt1 > k_tiles_total) kt1 = k_tiles_total;
            if (A_RES && w / n_tiles != cur_m) {
                cur_m = w / n_tiles;
                mbar_wait(a_full, a_phase);
                a_phase ^= 1;
            }
            const int as = it & 1;
            mbar_wait(&tmem_empty[as], (uint32_t)(((it >> 1) & 1) ^ 1));
            tc::fence_after_sync();
            if constexpr (Cfg::kColVec) {
                // the epilogue of the tile that used this accumulator stage is done, so its vector buffer is free
                const long long n0v = ((w % (m_tiles * n_tiles)) % n_tiles) * kTBN;
                long long nv = shape.N - n0v;
                if (nv > kTBN) nv = kTBN;
                const uint32_t vbytes = (uint32_t)(nv / 4) * 16;
                if (lane == 0 && vbytes) {
                    mbar_expect_tx(&vec_full[as], vbytes);
                    bulk_g2s(colvec_sm + as * kTBN, epi.colvec() + n0v, vbytes, &vec_full[as]);
                }
                __syncwarp();
            }
            const uint32_t d_addr = tmem_u + (uint32_t)as * kTBN;
            bool first = true;
            for (long long kt = kt0; kt < kt1; ++kt) {
                mbar_wait(&full_bar[stage], phase);
                if (Cfg::kConvert) mbar_wait(&conv_bar[stage], phase);
                tc::fence_after_sync();
                // All 32 lanes run this code converged; one lane is elected inside each MMA / commit asm.  Issuing
                // under `if (lane == 0)` makes the compiler wrap every MMA in an elect / R2UR.BROADCAST / branch loop
                // (~100 clk per MMA: longer than the ~90 clk the 128x128x8 MMA itself takes).
                const uint32_t slot = ring_u + (uint32_t)stage * (uint32_t)kSlot;
                const uint32_t b_addr = slot + (uint32_t)Cfg::kBOff;
                // TMEM A: resident tile = [hi: 128 cols | lo: 128 cols]; staged tile = [hi: 32 | lo: 32] per stage
                const uint32_t ta_hi = tmem_u + Cfg::kTmemA + (A_RES ? (uint32_t)kt * kTBK : (uint32_t)stage * 2 * kTBK);
                const uint32_t ta_lo = ta_hi + (A_RES ? (uint32_t)kTResKTiles * kTBK : (uint32_t)kTBK);
                if (!(shape.debug & 2)) {
#pragma unroll
                    for (int k = 0; k < kTBK / 8; ++k) {
                        const uint64_t b_hi = tc::smem_desc_sw128(b_addr + k * LB::k_step, LB::lbo, LB::sbo, LB::layout_type);
                        const uint64_t b_lo = tc::smem_desc_sw128(b_addr + LB::lo_offset + k * LB::k_step, LB::lbo, LB::sbo, LB::layout_type);
                        const bool acc0 = !(first && k == 0);
                        if constexpr (Cfg::kATmem) {
                            tc::mma_tf32_ts_warp(d_addr, ta_lo + k * 8, b_hi, idesc, acc0);    // small terms first
                            tc::mma_tf32_ts_warp(d_addr, ta_hi + k * 8, b_lo, idesc, true);
                            tc::mma_tf32_ts_warp(d_addr, ta_hi + k * 8, b_hi, idesc, true);
                        } else {
                            const uint64_t a_hi = tc::smem_desc_sw128(slot + k * LA::k_step, LA::lbo, LA::sbo, LA::layout_type);
                            const uint64_t a_lo = tc::smem_desc_sw128(slot + LA::lo_offset + k * LA::k_step, LA::lbo, LA::sbo, LA::layout_type);
                            tc::mma_tf32_warp(d_addr, a_lo, b_hi, idesc, acc0);
                            tc::mma_tf32_warp(d_addr, a_hi, b_lo, idesc, true);
                            tc::mma_tf32_warp(d_addr, a_hi, b_hi, idesc, true);
                        }
                    }
                }
                tc::mma_commit_warp(bars_u + (uint32_t)(8 + stage) * 8);        // empty_bar[stage]: frees the slot when these MMAs retire
                first = false;
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
            tc::mma_commit_warp(bars_u + (uint32_t)(24 + as) * 8);            // tmem_full[as]: accumulator ready
            if (A_RES && w + 1 < w_end && (w + 1) / n_tiles != cur_m) tc::mma_commit_warp(bars_u + 29 * 8);   // a_empty: last use of this A tile
        }
    } else if (Cfg::kATmem && warp >= 2 + EW) {
        // ===================================================== auxiliary warps: operand preparation
        // Four warps, one per TMEM lane quarter (a warp may only touch lanes 32 (warp % 4) ..+31): thread = A row.
        //   A_RES: at every change of m tile, read this thread's raw fp32 A row, split it and write hi / lo to TMEM
        //   A_RAW: per k tile, the raw A tile TMA put into the slot (128-byte swizzle: 16-byte unit j of row r at
        //          r*128 + ((j ^ (r & 7)) << 4)) -> tf32 hi / lo -> the slot's TMEM stage
        //   B_RAW: per k tile, split the raw B tile in place (element-wise, so independent of the swizzle: every
        //          16-byte unit is rewritten as hi and mirrored as lo at the plane offset of the layout)
        const int qa = warp & 3;
        const uint32_t lane_base = (uint32_t)(qa * 32) << 16;
        const int tid_c = (int)threadIdx.x - (64 + 32 * EW);
        using LB = OperandLayout<B_MN>;
        int stage = 0;
        uint32_t phase = 0, a_phase = 0;
        long long cur_m = -1;
        for (long long w = w_begin; w < w_end; w += w_step) {
            if constexpr (A_RES) {
                const long long m = w / n_tiles;
                if (m != cur_m) {
                    if (cur_m >= 0) { mbar_wait(a_empty, a_phase); a_phase ^= 1; tc::fence_after_sync(); }
                    cur_m = m;
                    const long long row = m * kTBM + qa * 32 + lane;
                    const bool live = row < shape.M;
                    const float* src = ag.base + row * ag.ld;
#pragma unroll 1
                    for (int kt = 0; kt < kTResKTiles; ++kt) {
                        float hi[32], lo[32];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int c = kt * kTBK + 4 * j;
                            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                            if (live && c < shape.K) t = __ldg(reinterpret_cast<const float4*>(src + c));
                            tc::split_tf32_fast(t.x, hi[4 * j], lo[4 * j]);
                            tc::split_tf32_fast(t.y, hi[4 * j + 1], lo[4 * j + 1]);
                            tc::split_tf32_fast(t.z, hi[4 * j + 2], lo[4 * j + 2]);
                            tc::split_tf32_fast(t.w, hi[4 * j + 3], lo[4 * j + 3]);
                        }
                        const uint32_t ta = tmem_base + lane_base + Cfg::kTmemA + (uint32_t)(kt * kTBK);
                        tc::tmem_st_32x32(ta, hi);
                        tc::tmem_st_32x32(ta + kTResKTiles * kTBK, lo);
                    }
                    tc::tmem_st_wait();
                    tc::fence_before_sync();
                    mbar_arrive(a_full);
                }
            }
            if constexpr (Cfg::kConvert) {
                const int split = (int)(w / (m_tiles * n_tiles));
                const long long kt0 = (long long)split * shape.k_tiles_per_split;
                long long kt1 = kt0 + shape.k_tiles_per_split;
                if (kt1 > k_tiles_total) kt1 = k_tiles_total;
                for (long long kt = kt0; kt < kt1; ++kt) {
                    mbar_wait(&full_bar[stage], phase);
                    unsigned char* slot = ring + (size_t)stage * kSlot;
                    if constexpr (A_RAW) {
                        const unsigned char* my = slot + (qa * 32 + lane) * 128;
                        float hi[32], lo[32];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 x = *reinterpret_cast<const float4*>(my + ((j ^ (lane & 7)) << 4));
                            tc::split_tf32_fast(x.x, hi[4 * j], lo[4 * j]);
                            tc::split_tf32_fast(x.y, hi[4 * j + 1], lo[4 * j + 1]);
                            tc::split_tf32_fast(x.z, hi[4 * j + 2], lo[4 * j + 2]);
                            tc::split_tf32_fast(x.w, hi[4 * j + 3], lo[4 * j + 3]);
                        }
                        const uint32_t ta = tmem_base + lane_base + Cfg::kTmemA + (uint32_t)stage * 2 * kTBK;
                        tc::tmem_st_32x32(ta, hi);
                        tc::tmem_st_32x32(ta + kTBK, lo);
                    }
                    if (B_RAW && !(shape.debug & 8)) {
                        unsigned char* sb = slot + Cfg::kBOff;
#pragma unroll
                        for (int i = 0; i < (int)(kTPlaneBytes / 16 / 128); ++i) {
                            const int u = tid_c + i * 128;
                            // K-major: [hi 16 KB | lo 16 KB]; MN-major: four 8 KB blocks of [hi 4 KB | lo 4 KB]
                            const uint32_t off = B_MN ? (uint32_t)((u >> 8) * 8192 + (u & 255) * 16) : (uint32_t)u * 16;
                            float4* hp = reinterpret_cast<float4*>(sb + off);
                            const float4 x = *hp;
                            float4 hv, lv;
                            tc::split_tf32_fast(x.x, hv.x, lv.x); tc::split_tf32_fast(x.y, hv.y, lv.y);
                            tc::split_tf32_fast(x.z, hv.z, lv.z); tc::split_tf32_fast(x.w, hv.w, lv.w);
                            *hp = hv;
                            *reinterpret_cast<float4*>(sb + off + LB::lo_offset) = lv;
                        }
                        if (!(shape.debug & 4)) tc::fence_proxy_async_smem();
                    }
                    if constexpr (A_RAW) { tc::tmem_st_wait(); tc::fence_before_sync(); }
                    mbar_arrive(&conv_bar[stage]);
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else {
        // ===================================================== epilogue (warps 2 .. 2+EW-1)
        const int q = warp & 3;                      // TMEM lane quarter this warp may access
        const int h = (warp - 2) >> 2;               // which column blocks of the quarter this warp takes
        unsigned char* staging = smem + Cfg::kStagingOff + (size_t)(warp - 2) * kTStagingBytes;
        int it = 0;
        for (long long w = w_begin; w < w_end; w += w_step, ++it) {
            const long long tile = w % (m_tiles * n_tiles);
            const int split = (int)(w / (m_tiles * n_tiles));
            const long long n0 = (tile % n_tiles) * kTBN, m0 = (tile / n_tiles) * kTBM;
            const int as = it & 1;
            mbar_wait(&tmem_full[as], (uint32_t)((it >> 1) & 1));
            tc::fence_after_sync();
            if constexpr (Cfg::kColVec) {
                if (shape.N - n0 >= 4) mbar_wait(&vec_full[as], (uint32_t)((it >> 1) & 1));
            }
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
                if (shape.debug & 1) continue;
                if constexpr (Cfg::kStaged) {
                    if (lane == 0) tc::bulk_wait_read0();      // the previous store has finished reading the buffer
                    __syncwarp();
                    if (live) epi.staged(row, col0, v, (int)ncols, split, pre, staging, lane, colvec_sm + as * kTBN + c * 32);
                    tc::fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0 && ncols > 0 && m0 + q * 32 < shape.M) {   // rows / columns past the edge are clipped by the map
                        tc::tma_store_2d(&epi.tmap, staging, (int)col0, (int)(m0 + q * 32));
                        tc::bulk_commit();
                    }
                } else {
                    if (live) epi(row, col0, v, (int)ncols, split, pre);
                }
            }
            tc::fence_before_sync();
            mbar_arrive(&tmem_empty[as]);
        }
        if (Cfg::kStaged && lane == 0) tc::bulk_wait0();
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc<Cfg::kTmemCols>(tmem_base);
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
// k splits for a GEMM with fewer output tiles than SMs: minimise (waves of work items) x (k tiles per item) with
// at least 8 k tiles per split; among schedules within 3 % of the best, the fewest splits (fewer partial tiles).  max_chain bounds the k
// tiles accumulated in one TMEM chain (the tensor core's fp32 accumulate truncates: the relative error grows
// by about 7e-9 per unit of K), 0 = unbounded.
inline int auto_splits(long long M, long long N, long long K, long long max_chain = 0) {
    const long long tiles = ((M + kTBM - 1) / kTBM) * ((N + kTBN - 1) / kTBN);
    const long long kt = (K + kTBK - 1) / kTBK;
    const long long sp_min = (max_chain > 0) ? (kt + max_chain - 1) / max_chain : 1;
    if (tiles >= kNumSMs && sp_min <= 1) return 1;
    const long long sp0 = sp_min < 1 ? 1 : sp_min;
    auto cost_of = [&](long long sp) {
        const long long per = (kt + sp - 1) / sp;
        const long long items = tiles * ((kt + per - 1) / per);
        return ((items + kNumSMs - 1) / kNumSMs) * per;
    };
    long long best_cost = cost_of(sp0);
    for (long long sp = sp0 + 1; sp <= sp0 + 32 && sp * 8 <= kt; ++sp) best_cost = cost_of(sp) < best_cost ? cost_of(sp) : best_cost;
    long long best = sp0;                                   // fewest splits within 3 % of the best schedule
    while (cost_of(best) * 100 > best_cost * 103) ++best;
    return (int)best;
}

int tf32x3_num_sms();

template <bool A_MN, bool B_MN, class Epi, int EW = 4, bool A_RES = false, bool A_RAW = false, bool B_RAW = false>
static int launch_tf32x3(const SplitMat& A, const SplitMat& B, const GemmShape& shape, const Epi& epi,
                         cudaStream_t st) {
    using Cfg = GemmCfg<Epi, EW, A_RES, A_RAW, B_RAW>;
    if (A_RES && (shape.splits != 1 || shape.K > (long long)kTResKTiles * kTBK || (shape.K % 4) || (A.ld % 4) ||
                  (reinterpret_cast<uintptr_t>(A.base) & 15u)))
        return VB_ERR_ARG;
    CUtensorMap ta, tb;
    int rc = VB_OK;
    if (A_RES) memset(&ta, 0, sizeof(ta));               // A is read by the loader warps, not by TMA
    else if (A_RAW) rc = make_raw_tmap(&ta, A.base, A.rows, A.cols, A.ld, kTBM, false);
    else rc = make_split_tmap(&ta, A.base, A.rows, A.cols, A.ld, A.plane_stride, A_MN ? kTBK : kTBM, A_MN);
    if (rc != VB_OK) return rc;
    rc = B_RAW ? make_raw_tmap(&tb, B.base, B.rows, B.cols, B.ld, B_MN ? kTBK : kTBN, B_MN)
               : make_split_tmap(&tb, B.base, B.rows, B.cols, B.ld, B.plane_stride, B_MN ? kTBK : kTBN, B_MN);
    if (rc != VB_OK) return rc;
    auto kern = gemm_tf32x3_kernel<A_MN, B_MN, Epi, EW, A_RES, A_RAW, B_RAW>;
    static DeviceOnce configured;                          // per template instantiation, per device
    const int dev = current_device();
    if (configured.pending(dev)) {
        VB_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::kSmemBytes));
        configured.done(dev);
    }
    const long long total = ((shape.M + kTBM - 1) / kTBM) * ((shape.N + kTBN - 1) / kTBN) * shape.splits;
    long long grid = tf32x3_num_sms();
    if (grid > total) grid = total;
    GemmShape sh = shape;
    { const char* e = getenv("VB_GEMM_DEBUG"); if (e && Cfg::kStaged) sh.debug = atoi(e); }
    kern<<<(unsigned)grid, Cfg::kThreads, Cfg::kSmemBytes, st>>>(ta, tb, sh, AGlobal{A.base, A.ld}, epi);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

}  // namespace vb
