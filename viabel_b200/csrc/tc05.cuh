// Thin inline-PTX layer for the Blackwell tensor path: TMA tensor loads, mbarriers, TMEM
// allocation, tcgen05.mma (kind::tf32) with shared-memory descriptors, tcgen05.ld.
// Bit layouts follow the PTX ISA "tcgen05" chapter (shared-memory matrix descriptor,
// instruction descriptor) — the same fields CUTLASS names in cute/arch/mma_sm100_desc.hpp.
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace vb {
namespace tc {

// ---------------------------------------------------------------------------- mbarrier (on top of common.cuh)
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) { mbar_expect_tx(bar, bytes); }

// ---------------------------------------------------------------------------- TMA
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 3-D tiled load global -> shared, completion on `bar` (bytes counted by the TMA unit)
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                            int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        :: "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)),
           "r"(c0), "r"(c1), "r"(c2) : "memory");
}

// 2-D tiled load global -> shared (plain fp32 matrix)
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        :: "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
// 2-D tiled store shared -> global (bulk async group of the issuing thread); out-of-range box elements are clipped
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* smem_src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 :: "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all bulk groups of this thread have finished READING their shared-memory source
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the async proxy (TMA store, tcgen05.mma operand reads)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---------------------------------------------------------------------------- TMEM
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result) {      // one full warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 :: "r"(smem_u32(smem_result)), "n"(kCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {            // same warp that allocated
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(taddr), "n"(kCols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// 32 lanes x 32 columns of 32-bit: thread t of the warp gets lane (base_lane + t), columns c .. c+31
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// registers -> TMEM: thread t of the warp writes lane (base_lane + t), columns c .. c+31 (wait with tmem_st_wait)
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const float (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        :: "r"(taddr),
           "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]),
           "f"(v[8]), "f"(v[9]), "f"(v[10]), "f"(v[11]), "f"(v[12]), "f"(v[13]), "f"(v[14]), "f"(v[15]),
           "f"(v[16]), "f"(v[17]), "f"(v[18]), "f"(v[19]), "f"(v[20]), "f"(v[21]), "f"(v[22]), "f"(v[23]),
           "f"(v[24]), "f"(v[25]), "f"(v[26]), "f"(v[27]), "f"(v[28]), "f"(v[29]), "f"(v[30]), "f"(v[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---------------------------------------------------------------------------- descriptors
// Shared-memory matrix descriptor, 128-byte swizzle.
//   bits  0-13 start address >> 4 | 16-29 leading byte offset >> 4 | 32-45 stride byte offset >> 4
//   bits 46-47 version (1 on sm_100) | 61-63 layout type (2 = SWIZZLE_128B)
//   layout type 2 = SWIZZLE_128B (16-byte atoms; K-major operands), 1 = SWIZZLE_128B_BASE32B (32-byte
//   atoms: the only swizzled layout the hardware accepts for MN-major tf32 operands).
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                    uint32_t layout_type = 2) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)layout_type << 61;
    return d;
}
// Instruction descriptor for kind::tf32, fp32 accumulate.
//   bits 4-5 D format (1 = F32) | 7-9 A format (2 = TF32) | 10-12 B format (2 = TF32)
//   bit 15 A major (1 = MN) | bit 16 B major (1 = MN) | 17-22 N >> 3 | 24-28 M >> 4
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N, bool a_mn_major, bool b_mn_major) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         bool accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n"
        :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate) : "memory");
}
// Same with the A operand read from tensor memory (rows on lanes 0..127, one tf32 per 32-bit column, K along columns)
__device__ __forceinline__ void mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                            bool accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n"
        "}\n"
        :: "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate) : "memory");
}
// Warp-collective forms: called by ALL 32 converged lanes with warp-uniform operands; one elected lane issues.
// Keeping the election inside the asm leaves the C++ control flow convergent, so the compiler holds the operands
// in uniform registers instead of handing each one over from a vector register before every MMA.
__device__ __forceinline__ void mma_tf32_ts_warp(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                                 bool accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p, e;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "@e tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n"
        "}\n"
        :: "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate) : "memory");
}
__device__ __forceinline__ void mma_tf32_warp(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              bool accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p, e;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "@e tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n"
        :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate) : "memory");
}
__device__ __forceinline__ void mma_commit_warp(uint32_t bar_smem_addr) {
    asm volatile(
        "{\n"
        ".reg .pred e;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n"
        "}\n"
        :: "r"(bar_smem_addr) : "memory");
}
// mbarrier arrive once all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                 :: "r"(smem_u32(bar)) : "memory");
}

// ---------------------------------------------------------------------------- kind::f16 (fp16 operands, fp32 accumulate)
// Instruction descriptor: bits 4-5 D format (1 = F32) | 7-9 A format (0 = F16) | 10-12 B format (0 = F16)
//   bit 15 A major (1 = MN) | bit 16 B major (1 = MN) | 17-22 N >> 3 | 24-28 M >> 4.   UMMA_K = 16.
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N, bool a_mn_major, bool b_mn_major) {
    return (1u << 4) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// A from tensor memory (lane = row; a 32-bit column holds TWO consecutive K elements, low half first), B from
// shared memory; warp-collective form (all 32 lanes converged, one elected lane issues)
__device__ __forceinline__ void mma_f16_ts_warp(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                                bool accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p, e;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}\n"
        :: "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate) : "memory");
}
// 32 lanes x 16 columns of 32-bit, registers -> TMEM
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        :: "r"(taddr),
           "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
           "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
// TMEM -> registers without the wait (issue several loads, then tmem_ld_wait once)
__device__ __forceinline__ void tmem_ld_32x32_nowait(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16_nowait(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_st_32x8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
        :: "r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// two fp32 -> packed f16x2 (round to nearest even); `lo16` lands in the low half
__device__ __forceinline__ uint32_t pack_f16x2(float lo16, float hi16) {
    uint32_t r;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi16), "f"(lo16));
    return r;
}
// Cheap fp16 hi / lo split of a pair for streaming epilogues: hi = x TRUNCATED to 11 significant bits (one
// round-toward-zero pack), lo = fp16_rn(x - hi) with x - hi exact in fp32 (|lo| < 2^-10 |x|).  The truncation is
// not an error source: lo carries the remainder, so x = hi + lo up to the unbiased rounding of lo, 2^-21 |x|.
__device__ __forceinline__ void split_f16_pair_rz(float2 x, uint32_t& hi2, uint32_t& lo2) {
    asm("cvt.rz.f16x2.f32 %0, %1, %2;" : "=r"(hi2) : "f"(x.y), "f"(x.x));
    float2 h;                                               // hi as fp32: the low 13 mantissa bits masked off
    h.x = __uint_as_float(__float_as_uint(x.x) & 0xffffe000u);
    h.y = __uint_as_float(__float_as_uint(x.y) & 0xffffe000u);
    const float2 l = __ffma2_rn(h, make_float2(-1.f, -1.f), x);     // x - hi, exact
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(lo2) : "f"(l.y), "f"(l.x));
}
// 16-byte shared-memory load through a 32-bit shared address (a generic pointer makes the compiler emit LD.E)
__device__ __forceinline__ float4 lds128(uint32_t saddr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
    return v;
}
// fp16 hi / lo split of a pair: hi = x rounded to 11 significant bits (integer add + mask: exact in fp16 when
// normal), lo = x - hi (exact in fp32), both packed as f16x2.  22 significant bits in total, like tf32 hi / lo.
__device__ __forceinline__ void split_f16_pair(float a, float b, uint32_t& hi2, uint32_t& lo2) {
    const float ah = __uint_as_float((__float_as_uint(a) + 0x1000u) & 0xffffe000u);
    const float bh = __uint_as_float((__float_as_uint(b) + 0x1000u) & 0xffffe000u);
    hi2 = pack_f16x2(ah, bh);
    lo2 = pack_f16x2(a - ah, b - bh);
}

// round-to-nearest fp32 -> tf32 (low 13 mantissa bits cleared)
__device__ __forceinline__ float to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
// Cheap split for finite data: hi = round-half-away to tf32 by integer add + mask, lo = x - hi left unrounded
// (the tensor core reads only the upper 19 bits of a tf32 operand, i.e. truncates lo: error 2^-22 |x|)
__device__ __forceinline__ void split_tf32_fast(float x, float& hi, float& lo) {
    hi = __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
    lo = x - hi;
}
// x = hi + lo + O(2^-22 |x|), both exactly representable in tf32
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
    hi = to_tf32(x);
    lo = to_tf32(x - hi);
}

}  // namespace tc

// ---------------------------------------------------------------------------- host: tensor maps
// cuTensorMapEncodeTiled resolved at run time through the runtime's driver entry point query,
// so the library has no link-time dependency on libcuda.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode_tiled();

// A "split" fp32 matrix: plane 0 = tf32 hi part, plane 1 = tf32 lo part, each (rows, cols) row-major
// with leading dimension ld (elements).  The map is 3-D (cols, rows, 2) with a (32, box_rows, 2) box
// and 128-byte swizzle: one load brings the hi and the lo tile of a 128-byte-wide column block.
// mn_major operands use the 32-byte-atom flavour of the 128-byte swizzle (see smem_desc_sw128).
int make_split_tmap(CUtensorMap* out, const float* base, long long rows, long long cols, long long ld,
                    long long plane_stride_elems, int box_rows, bool mn_major);
// A plain fp32 (rows, cols) row-major matrix: 2-D map, (32, box_rows) box, 128-byte swizzle.  Loaded tiles have
// the layout of ONE plane of a K-major split tile (so a tile can be split into hi / lo in shared memory);
// the same map stores (32 x box_rows) tiles staged in shared memory.
int make_raw_tmap(CUtensorMap* out, const float* base, long long rows, long long cols, long long ld, int box_rows,
                  bool atom32 = false);
// fp16 hi / lo planes of a (rows, cols) matrix, leading dimension ld (elements, multiple of 8), plane 1 following
// plane 0 at plane_stride elements: 3-D map (cols, rows, 2), box (64, box_rows, 2), 128-byte swizzle.
int make_f16_planes_tmap(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld,
                         long long plane_stride_elems, int box_rows);

}  // namespace vb
