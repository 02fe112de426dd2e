// Fused Student-t regression plugin: log p and its gradient for S coefficient draws over N data
// rows in ONE persistent tcgen05 kernel -- the S x N residual / psi matrix never leaves the SM.
//   reference: notebooks/robust-regression.ipynb:73-86 evaluated through make_stan_log_density (viabel/vb.py:301-321)
//   log p(b) = -|b|^2 / (2 sd^2) - (df+1)/2 sum_r log1p(r_r^2 / df),   r = y - X b
//   grad     = -b / sd^2 + c X' psi,   psi_r / c = r_r / (1 + r_r^2 / df),   c = (df + 1) / df
//
// Shape (flash-attention like).  A CTA owns a 128-draw tile B_t (A operand, resident in tensor memory) and walks
// 64-row tiles X_j of its row range:
//   GEMM1   acc1 (128 x 64)  = B_t X_j'            tcgen05.mma kind::f16, A from TMEM, B = X_j K-major in smem
//   epilogue (8 warps)       r, sum log(u), psi    acc1 TMEM -> registers -> psi (fp16 hi / lo) -> TMEM
//   GEMM2   G (128 x 128)   += Psi_j X_j           A = Psi_j from TMEM, B = THE SAME smem bytes of X_j read MN-major
// G stays in tensor memory for a chain of row tiles and is flushed to a per-(row range, draw tile) fp32 partial.
//
// Arithmetic: "fp16 x 3".  Every fp32 operand x is carried as two fp16 numbers, hi = x rounded to 11 significant
// bits and lo = fp16(x - hi): 22 significant bits, exactly what the tf32 (hi, lo) pairs of the 3xTF32 scheme carry,
// and a b ~= a_lo b_hi + a_hi b_lo + a_hi b_hi accumulates in fp32 in tensor memory.  fp16 MMAs run at twice the
// tf32 rate and the operands take half the bytes.  fp16's narrow exponent is handled by power-of-two pre-scaling
// (exact): X by one factor for the whole matrix (max |x| -> [2^13, 2^14)), every draw by its own factor, psi by a
// constant; values more than 2^-27 below the row / matrix maximum lose relative (not absolute) precision.
#include <cstdlib>
#include <cuda_fp16.h>
#include "tc05.cuh"
#include "dense.cuh"
#include "models.cuh"

namespace vb {

constexpr int kFM = 128;                 // draws per tile (TMEM lanes)
constexpr int kFR = 64;                  // data rows per tile
constexpr int kFD = 128;                 // max dimension (K of GEMM1, N of GEMM2)
constexpr int kFStages = 5;              // X ring depth
constexpr uint32_t kFXBytes = 2u * 2u * kFR * 64 * 2;         // [col block 2][plane 2][64 rows][64 halves] = 32 KB
constexpr uint32_t kFStageBytes = kFXBytes + 256;             // + the tile's 64 y values
constexpr int kFEW = 16;                 // epilogue warps: warp (q, h) = TMEM lane quarter q, columns 16 h .. 16 h + 15
constexpr int kFThreads = 128 + 32 * kFEW;                    // warp 0 TMA, warp 1 MMA, warps 2-3 idle, then the epilogue
constexpr int kFEpiThreads = 32 * kFEW;
constexpr int kFG = kFEW / 4;            // column groups per lane quarter
constexpr int kFC = kFR / kFG;           // columns of a row tile per epilogue thread (16)
static_assert(kFC == 16 && kFD / kFG == 32, "epilogue register tiling assumes 16 epilogue warps");
// tensor-memory columns
constexpr uint32_t kTG = 0, kTAcc = 128, kTBhi = 256, kTBlo = 320, kTPsi = 384;
constexpr size_t kFSmemBytes = (size_t)kFStages * kFStageBytes + 512 + 1024;

struct F16Header { float scale, inv_scale; unsigned int absmax_bits; int ld; };

// ------------------------------------------------------------------------------------------ X -> scaled fp16 planes
__global__ void f16_absmax_kernel(const float* __restrict__ src, long long rows, long long cols, long long lds,
                                  F16Header* hdr) {
    unsigned int m = 0;
    const long long total = rows * cols;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const long long r = e / cols, c = e - r * cols;
        const unsigned int b = __float_as_uint(src[r * lds + c]) & 0x7fffffffu;
        if (b < 0x7f800000u && b > m) m = b;             // finite values only
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(&hdr->absmax_bits, m);
}
// scale = 2^(13 - floor(log2 max)); 1 when the matrix is zero / subnormal
__device__ __forceinline__ float pow2_scale_for(unsigned int absmax_bits) {
    const unsigned int eb = absmax_bits >> 23;
    if (eb < 14u) return 1.f;
    return __uint_as_float((267u - eb) << 23);
}
__global__ void f16_split_kernel(const float* __restrict__ src, long long rows, long long cols, long long lds,
                                 __half* __restrict__ dst, long long ldd, long long plane, F16Header* hdr) {
    const float sc = pow2_scale_for(hdr->absmax_bits);
    const long long total = rows * ldd;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const long long r = e / ldd, c = e - r * ldd;
        float hi = 0.f, lo = 0.f;
        if (c < cols) {
            const float x = sc * src[r * lds + c];
            hi = __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
            lo = x - hi;
        }
        dst[e] = __float2half_rn(hi);
        dst[e + plane] = __float2half_rn(lo);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { hdr->scale = sc; hdr->inv_scale = 1.f / sc; hdr->ld = (int)ldd; }
}

static inline long long pad8(long long x) { return (x + 7) / 8 * 8; }
static inline int grid_for_f(long long work_items, int threads, int max_blocks) {
    long long g = (work_items + threads - 1) / threads;
    if (g > max_blocks) g = max_blocks;
    if (g < 1) g = 1;
    return (int)g;
}
size_t split_f16_bytes(long long rows, long long cols) {
    return align_up((size_t)2 * rows * pad8(cols) * 2, 256) + 256;
}
static F16Header* f16_header(const void* planes, long long rows, long long cols) {
    return (F16Header*)((char*)planes + align_up((size_t)2 * rows * pad8(cols) * 2, 256));
}

// ------------------------------------------------------------------------------------------ epilogue arithmetic
// log(p) for p in [1, 3e38): p = 2^e m, m in [2/3, 4/3); log m = f - f^2/2 + f^3 Q(f), f = m - 1 (degree-6 minimax Q,
// max abs error 3.4e-8: scripts/fit_log_poly.py); packed fp32x2 arithmetic (FADD2 / FMUL2 / FFMA2)
__device__ __forceinline__ float2 log_ge1_x2f(float2 p) {
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
__device__ __noinline__ float log1p_sum4_slow_f(float r0, float r1, float r2, float r3, float inv_df) {
    return (log1pf(r0 * r0 * inv_df) + log1pf(r1 * r1 * inv_df)) + (log1pf(r2 * r2 * inv_df) + log1pf(r3 * r3 * inv_df));
}
// sixteen residuals, everything in register PAIRS (packed fp32x2 arithmetic needs aligned pairs; unpaired operands
// cost a move each): acc = tensor-memory accumulator pairs, ys = psi_scale * y pairs.
//   r' = psi_scale r = fma(acc, c1, ys),  c1 = -psi_scale / (draw scale * X scale)
//   u  = 1 + r^2/df = fma(r' r', idf, 1), idf = 1 / (df psi_scale^2)
//   sum log u: ONE packed log per sixteen (two products of eight; a product overflows only for |r| > ~1600 sqrt(df/40):
//   then the slow path);   psi' = psi_scale psi / c = r' / u -> fp16 (hi, lo)
__device__ __forceinline__ void psi_hex(const float2 (&acc)[8], const float2 (&ys)[8], float c1, float idf, float inv_df,
                                        float inv_ps, float2& lsum, uint32_t (&hi2)[8], uint32_t (&lo2)[8]) {
    const float2 c1v = make_float2(c1, c1), one = make_float2(1.f, 1.f), idfv = make_float2(idf, idf);
    float2 r[8], u[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        r[i] = __ffma2_rn(acc[i], c1v, ys[i]);
        u[i] = __ffma2_rn(__fmul2_rn(r[i], r[i]), idfv, one);
    }
    const float2 p = __fmul2_rn(__fmul2_rn(__fmul2_rn(u[0], u[1]), __fmul2_rn(u[2], u[3])),
                                __fmul2_rn(__fmul2_rn(u[4], u[5]), __fmul2_rn(u[6], u[7])));
    if (p.x < 3.0e38f && p.y < 3.0e38f) {
        lsum = __fadd2_rn(lsum, log_ge1_x2f(p));
    } else {                                                   // rare: a product of eight overflowed
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
            lsum.x += log1p_sum4_slow_f(r[i].x * inv_ps, r[i].y * inv_ps, r[i + 1].x * inv_ps, r[i + 1].y * inv_ps, inv_df);
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float2 iv;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iv.x) : "f"(u[i].x));
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iv.y) : "f"(u[i].y));
        tc::split_f16_pair_rz(__fmul2_rn(r[i], iv), hi2[i], lo2[i]);
    }
}

struct FusedArgs {
    const float* b;            // (S, D) fp32 draws
    const float* ys;           // (N padded to 64) psi_scale * y, zero past N (made per call by reg_fused_prep_kernel)
    const F16Header* xhdr;     // scale of the X planes
    float* gpart;              // [R][S_pad][128] fp32 partial gradients (scaled by psi_scale * x_scale)
    double* lpart;             // [R][kFG][S_pad] partial sums of log(1 + r^2/df)
    long long S, S_pad, N;
    int D, n_dt, DG, R, flush_tiles;
    float inv_df, psi_scale;
};

// ------------------------------------------------------------------------------------------ the kernel
// CTA c -> (row range rr = c / DG, draw group dg = c % DG).  It runs one PASS per draw tile dt = dg, dg + DG, ... :
// load the 128 draws into tensor memory, walk the row tiles [rt0, rt1) of its range.
__global__ void __launch_bounds__(kFThreads, 1)
reg_fused_kernel(const __grid_constant__ CUtensorMap tmap_x, const FusedArgs a) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)kFStages * kFStageBytes);
    uint64_t* full = bars;                    // [8]
    uint64_t* empty = bars + 8;               // [8]
    uint64_t* acc_full = bars + 16;           // [2]
    uint64_t* acc_empty = bars + 18;          // [2]
    uint64_t* psi_full = bars + 20;           // [2]
    uint64_t* psi_empty = bars + 22;          // [2]
    uint64_t* b_full = bars + 24;
    uint64_t* b_empty = bars + 25;
    uint64_t* g_full = bars + 26;
    uint64_t* g_empty = bars + 27;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 28);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rr = blockIdx.x / a.DG, dg = blockIdx.x % a.DG;
    const long long T = (a.N + kFR - 1) / kFR;
    const long long rt0 = T * rr / a.R, rt1 = T * (rr + 1) / a.R;
    const long long nt = rt1 - rt0;                               // row tiles per pass (>= 1: R <= T)
    const int n_pass = (a.n_dt - dg + a.DG - 1) / a.DG;

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tmap_x);
#pragma unroll
        for (int s = 0; s < 8; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], kFEpiThreads);
            mbar_init(&psi_full[s], kFEpiThreads); mbar_init(&psi_empty[s], 1);
        }
        mbar_init(b_full, kFEpiThreads); mbar_init(b_empty, 1);
        mbar_init(g_full, 1); mbar_init(g_empty, kFEpiThreads);
        mbar_fence_init();
    }
    if (warp == 1) tc::tmem_alloc<512>(tmem_ptr);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ===================================================== TMA producer
        int s = 0;
        uint32_t ph = 0;
        for (int p = 0; p < n_pass; ++p) {
            for (long long t = rt0; t < rt1; ++t) {
                mbar_wait(&empty[s], ph ^ 1);
                if (lane == 0) {
                    unsigned char* dst = smem + (size_t)s * kFStageBytes;
                    const long long r0 = t * kFR;
                    mbar_expect_tx(&full[s], kFXBytes + 256);
                    tc::tma_load_3d(dst, &tmap_x, &full[s], 0, (int)r0, 0);          // rows past N: zero-filled
                    tc::tma_load_3d(dst + 16384, &tmap_x, &full[s], 64, (int)r0, 0);
                    bulk_g2s(dst + kFXBytes, a.ys + r0, 256, &full[s]);
                }
                __syncwarp();
                if (++s == kFStages) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (all lanes converged; one elected per MMA)
        constexpr uint32_t idesc1 = tc::idesc_f16(kFM, kFR, false, false);      // B = X tile K-major, N = 64 rows
        constexpr uint32_t idesc2 = tc::idesc_f16(kFM, kFD, false, true);       // B = X tile MN-major, N = 128 dims
        const uint32_t tm = __reduce_or_sync(0xffffffffu, tmem_base);
        const uint32_t smem_u = smem_u32(smem);
        const uint32_t bars_u = smem_u32(bars);
        long long it = 0;
        int chain = 0;
        uint32_t flushes = 0;
        int s1 = 0, s2 = 0;                           // X ring slots of the next GEMM1 / GEMM2 tile
        uint32_t ph1 = 0;
        auto gemm2 = [&](long long j, bool last_of_pass) {
            const int s = s2, pb = (int)(j & 1);
            if (++s2 == kFStages) s2 = 0;
            mbar_wait(&psi_full[pb], (uint32_t)((j >> 1) & 1));
            if (chain == 0) mbar_wait(g_empty, (flushes & 1) ^ 1);
            tc::fence_after_sync();
            const uint32_t xs = smem_u + (uint32_t)s * kFStageBytes;
            const uint32_t psi = tm + kTPsi + (uint32_t)pb * 64;
#pragma unroll
            for (int ks = 0; ks < kFR / 16; ++ks) {
                // MN-major: 64 dims = one 128-byte line; next 64 dims at +16384 (LBO), next 8 rows at +1024 (SBO)
                const uint64_t x_hi = tc::smem_desc_sw128(xs + ks * 2048, 16384, 1024, 2);
                const uint64_t x_lo = tc::smem_desc_sw128(xs + 8192 + ks * 2048, 16384, 1024, 2);
                tc::mma_f16_ts_warp(tm + kTG, psi + 32 + ks * 8, x_hi, idesc2, !(chain == 0 && ks == 0));   // small terms first
                tc::mma_f16_ts_warp(tm + kTG, psi + ks * 8, x_lo, idesc2, true);
                tc::mma_f16_ts_warp(tm + kTG, psi + ks * 8, x_hi, idesc2, true);
            }
            tc::mma_commit_warp(bars_u + (uint32_t)(8 + s) * 8);              // empty[s]: X slot free
            tc::mma_commit_warp(bars_u + (uint32_t)(22 + pb) * 8);            // psi_empty[pb]
            ++chain;
            if (chain == a.flush_tiles || last_of_pass) {
                tc::mma_commit_warp(bars_u + 26 * 8);                         // g_full: chain complete
                chain = 0;
                ++flushes;
            }
        };
        // GEMM1 runs kLook tiles ahead of GEMM2: the epilogue of tile t (TMEM load, ~250 instructions per thread,
        // TMEM store) takes longer than one GEMM1, so with a look-ahead of one the tensor pipe idled on psi_full
        constexpr int kLook = 2;
        for (int p = 0; p < n_pass; ++p) {
            mbar_wait(b_full, (uint32_t)(p & 1));
            const long long it0 = it;
            for (long long t = 0; t < nt + kLook; ++t) {
                if (t < nt) {
                    const int s = s1, ab = (int)(it & 1);
                    mbar_wait(&full[s], ph1);
                    if (++s1 == kFStages) { s1 = 0; ph1 ^= 1; }
                    mbar_wait(&acc_empty[ab], (uint32_t)(((it >> 1) & 1) ^ 1));
                    tc::fence_after_sync();
                    const uint32_t xs = smem_u + (uint32_t)s * kFStageBytes;
                    const uint32_t d1 = tm + kTAcc + (uint32_t)ab * kFR;
#pragma unroll
                    for (int k = 0; k < kFD / 16; ++k) {
                        // K-major: col block (k / 4) at +16384, 16 dims = 32 bytes inside the 128-byte line; 8 rows = 1024 B
                        const uint32_t off = (uint32_t)(k >> 2) * 16384 + (uint32_t)(k & 3) * 32;
                        const uint64_t x_hi = tc::smem_desc_sw128(xs + off, 16, 1024, 2);
                        const uint64_t x_lo = tc::smem_desc_sw128(xs + 8192 + off, 16, 1024, 2);
                        tc::mma_f16_ts_warp(d1, tm + kTBlo + k * 8, x_hi, idesc1, k != 0);
                        tc::mma_f16_ts_warp(d1, tm + kTBhi + k * 8, x_lo, idesc1, true);
                        tc::mma_f16_ts_warp(d1, tm + kTBhi + k * 8, x_hi, idesc1, true);
                    }
                    tc::mma_commit_warp(bars_u + (uint32_t)(16 + ab) * 8);        // acc_full[ab]
                    if (t == nt - 1) tc::mma_commit_warp(bars_u + 25 * 8);        // b_empty: last GEMM1 on this draw tile
                    ++it;
                }
                if (t >= kLook) gemm2(it0 + t - kLook, t - kLook == nt - 1);
            }
        }
    } else if (warp >= 4) {
        // ===================================================== epilogue warps: thread = draw row (TMEM lane) x 16 columns
        const int q = warp & 3, h = (warp - 4) >> 2;                  // lane quarter, column group
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
        const float x_inv = a.xhdr->inv_scale;
        const float idf = a.inv_df / (a.psi_scale * a.psi_scale), inv_ps = 1.f / a.psi_scale;
        long long it = 0;
        uint32_t flushes = 0;
        int st = 0;                                  // X ring slot / phase of tile `it`, kept incrementally
        uint32_t st_ph = 0;
        const uint32_t smem_sa = smem_u32(smem);
        for (int p = 0; p < n_pass; ++p) {
            const int dt = dg + p * a.DG;
            const long long row = (long long)dt * kFM + q * 32 + lane;
            // ---- this pass's draws -> tensor memory (fp16 hi / lo, per-row power-of-two scale); this thread converts
            // dims 32 h .. 32 h + 31 of its row
            if (p > 0) { mbar_wait(b_empty, (uint32_t)((p - 1) & 1)); tc::fence_after_sync(); }
            float b_scale = 1.f;
            {
                const float* src = a.b + row * a.D;
                const bool live = row < a.S;
                unsigned int mx = 0;
                if (live)
                    for (int j = 0; j < a.D; ++j) {
                        const unsigned int bb = __float_as_uint(__ldg(src + j)) & 0x7fffffffu;
                        if (bb < 0x7f800000u && bb > mx) mx = bb;
                    }
                b_scale = pow2_scale_for(mx);
                uint32_t hi2[16], lo2[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int c = h * 32 + 2 * j;
                    const float v0 = (live && c < a.D) ? b_scale * __ldg(src + c) : 0.f;
                    const float v1 = (live && c + 1 < a.D) ? b_scale * __ldg(src + c + 1) : 0.f;
                    tc::split_f16_pair(v0, v1, hi2[j], lo2[j]);
                }
                tc::tmem_st_32x16(lane_addr + kTBhi + h * 16, hi2);
                tc::tmem_st_32x16(lane_addr + kTBlo + h * 16, lo2);
                tc::tmem_st_wait();
                tc::fence_before_sync();
                mbar_arrive(b_full);
            }
            const float c1 = -(a.psi_scale * x_inv) / b_scale;
            double lsum = 0.0;
            float lc = 0.f, lcomp = 0.f;            // per-chain log-sum in compensated fp32 (the fp64 pipe is too slow per tile)
            int chain = 0;
            bool first_flush = true, pending = false;
            for (long long t = 0; t < nt; ++t, ++it) {
                const int ab = (int)(it & 1);
                const uint32_t ph2 = (uint32_t)((it >> 1) & 1);
                mbar_wait(&acc_full[ab], ph2);
                mbar_wait(&full[st], st_ph);                                   // y segment of the slot (already complete)
                tc::fence_after_sync();
                uint32_t raw[16];
                tc::tmem_ld_32x16_nowait(lane_addr + kTAcc + ab * kFR + h * kFC, raw);
                const uint32_t ya = smem_sa + (uint32_t)st * kFStageBytes + kFXBytes + (uint32_t)h * (kFC * 4);
                const float4 y0 = tc::lds128(ya), y1 = tc::lds128(ya + 16), y2 = tc::lds128(ya + 32), y3 = tc::lds128(ya + 48);
                if (++st == kFStages) { st = 0; st_ph ^= 1; }
                tc::tmem_ld_wait();
                tc::fence_before_sync();
                mbar_arrive(&acc_empty[ab]);
                uint32_t hi2[8], lo2[8];
                float2 ls = make_float2(0.f, 0.f);
                {
                    float2 av[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) av[i] = make_float2(__uint_as_float(raw[2 * i]), __uint_as_float(raw[2 * i + 1]));
                    const float2 yv[8] = {make_float2(y0.x, y0.y), make_float2(y0.z, y0.w), make_float2(y1.x, y1.y),
                                          make_float2(y1.z, y1.w), make_float2(y2.x, y2.y), make_float2(y2.z, y2.w),
                                          make_float2(y3.x, y3.y), make_float2(y3.z, y3.w)};
                    psi_hex(av, yv, c1, idf, a.inv_df, inv_ps, ls, hi2, lo2);
                }
                {
                    const float v = (ls.x + ls.y) - lcomp;       // Kahan
                    const float tsum = lc + v;
                    lcomp = (tsum - lc) - v;
                    lc = tsum;
                }
                mbar_wait(&psi_empty[ab], ph2 ^ 1);
                tc::fence_after_sync();
                // packed fp16 pairs: tile rows 16 h + 2 j, 16 h + 2 j + 1 -> column 8 h + j of the hi / lo block
                tc::tmem_st_32x8(lane_addr + kTPsi + ab * 64 + h * 8, hi2);
                tc::tmem_st_32x8(lane_addr + kTPsi + ab * 64 + 32 + h * 8, lo2);
                tc::tmem_st_wait();
                tc::fence_before_sync();
                mbar_arrive(&psi_full[ab]);
                ++chain;
                // ---- chain complete: G (this thread: one row, 32 dims) -> fp32 partial of (row range, draw tile).  The
                // drain is deferred by one tile (the epilogue works on tile t + 1 while the chain's last GEMM2 retires)
                // except at the end of a pass.
                auto drain = [&]() {
                    mbar_wait(g_full, flushes & 1);
                    tc::fence_after_sync();
                    float* dst = a.gpart + ((size_t)rr * a.S_pad + row) * kFD + h * 32;
                    uint32_t g[32];
                    tc::tmem_ld_32x32_nowait(lane_addr + kTG + h * 32, g);
                    tc::tmem_ld_wait();
                    tc::fence_before_sync();
                    mbar_arrive(g_empty);                           // G is in registers: the next chain may start
                    float4* d4 = reinterpret_cast<float4*>(dst);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        float4 v = make_float4(__uint_as_float(g[4 * j]), __uint_as_float(g[4 * j + 1]),
                                               __uint_as_float(g[4 * j + 2]), __uint_as_float(g[4 * j + 3]));
                        if (!first_flush) { const float4 o = d4[j]; v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
                        d4[j] = v;
                    }
                    ++flushes;
                    first_flush = false;
                };
                if (pending) { drain(); pending = false; }
                if (chain == a.flush_tiles || t == nt - 1) {
                    chain = 0;
                    lsum += (double)lc - (double)lcomp;
                    lc = 0.f; lcomp = 0.f;
                    if (t == nt - 1) drain(); else pending = true;
                }
            }
            a.lpart[((size_t)rr * kFG + h) * a.S_pad + row] = lsum;
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc<512>(tmem_base);
}

// logp[s] = -|b_s|^2 / (2 sd^2) - (df+1)/2 sum_{rr,h} lpart;  grad[s, j] = -b_sj / sd^2 + c / (psi_scale x_scale) sum_rr gpart
__global__ void reg_fused_final_kernel(const FusedArgs a, double inv_var, double half_dfp1, double c,
                                       double* __restrict__ logp, float* __restrict__ grad) {
    const int lane = threadIdx.x & 31;
    const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const double gs = c * (double)a.xhdr->inv_scale / (double)a.psi_scale;
    for (long long s = warp; s < a.S; s += nwarps) {
        double bb = 0.0;
        for (int j = lane; j < a.D; j += 32) {
            const double v = (double)a.b[s * a.D + j];
            bb = fma(v, v, bb);
            if (grad) {
                double g = 0.0;
                for (int r = 0; r < a.R; ++r) g += (double)a.gpart[((size_t)r * a.S_pad + s) * kFD + j];      // fixed order
                grad[s * a.D + j] = (float)fma(gs, g, -v * inv_var);
            }
        }
        bb = warp_sum(bb);
        if (lane == 0) {
            double l = 0.0;
            for (int r = 0; r < kFG * a.R; ++r) l += a.lpart[(size_t)r * a.S_pad + s];
            logp[s] = -0.5 * bb * inv_var - half_dfp1 * l;
        }
    }
}

// ys[r] = psi_scale * y[r] for r < N, 0 up to the next multiple of 64 (+ one spare tile)
__global__ void reg_fused_prep_kernel(const float* __restrict__ y, long long N, long long n_pad, float psi_scale,
                                      float* __restrict__ ys) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_pad; i += (long long)gridDim.x * blockDim.x)
        ys[i] = i < N ? psi_scale * y[i] : 0.f;
}

struct FusedPlan { int n_dt, DG, R; long long S_pad, n_pad; size_t off_g, off_l, off_y, total; };
static FusedPlan fused_plan(long long S, long long N) {
    FusedPlan p;
    p.n_dt = (int)((S + kFM - 1) / kFM);
    p.S_pad = (long long)p.n_dt * kFM;
    const long long T = (N + kFR - 1) / kFR;
    const int sms = device_sm_count();
    // draw groups x row ranges: minimise passes x row tiles per pass; within 3 % of the best, the MOST draw groups
    // (every pass re-streams the X planes of the CTA's row range: fewer passes = less DRAM traffic)
    int forced = 0;
    { const char* e = getenv("VB_REG_DG"); if (e) forced = atoi(e); }
    auto cost = [&](int dg) {
        long long R = sms / dg;
        if (R > T) R = T;
        if (R < 1) R = 1;
        return (long long)((p.n_dt + dg - 1) / dg) * ((T + R - 1) / R);
    };
    const int dg_max = p.n_dt < sms ? p.n_dt : sms;
    int best = 1;
    long long best_cost = cost(1);
    for (int dg = 2; dg <= dg_max; ++dg) if (cost(dg) < best_cost) { best_cost = cost(dg); best = dg; }
    for (int dg = dg_max; dg > best; --dg) if (cost(dg) * 100 <= best_cost * 103) { best = dg; break; }
    if (forced >= 1 && forced <= dg_max) best = forced;
    p.DG = best;
    long long R = sms / best;
    if (R > T) R = T;
    if (R < 1) R = 1;
    p.R = (int)R;
    p.off_g = 0;
    p.off_l = align_up((size_t)p.R * p.S_pad * kFD * 4, 256);
    p.off_y = p.off_l + align_up((size_t)p.R * kFG * p.S_pad * 8, 256);
    p.n_pad = T * kFR;
    p.total = p.off_y + align_up((size_t)p.n_pad * 4, 256);
    return p;
}

bool regression_fused_ok(const ModelHost* m) {
    const char* e = getenv("VB_REG_PATH");                   // "1" = the two-GEMM 3xTF32 pipeline (A/B, wide models)
    const bool off = e != nullptr && atoi(e) == 1;
    return !off && m->d_X_planes16 != nullptr && m->dim <= kFD;
}
size_t regression_fused_workspace_bytes(long long n, long long n_rows) { return fused_plan(n, n_rows).total + 512; }

int regression_logp_grad_fused(const ModelHost* m, const float* b, long long S, double* out_logp,
                               float* out_grad_or_null, void* ws, size_t ws_bytes, cudaStream_t st) {
    const int D = m->dim;
    const long long N = m->n_rows;
    const FusedPlan p = fused_plan(S, N);
    if (!ws || ws_bytes < p.total + 512) return VB_ERR_WORKSPACE;
    char* w = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
    const long long ldp = pad8(D);
    CUtensorMap tx;
    int rc = make_f16_planes_tmap(&tx, m->d_X_planes16, N, D, ldp, N * ldp, kFR);
    if (rc != VB_OK) return rc;
    FusedArgs a;
    a.b = b; a.ys = (const float*)(w + p.off_y); a.xhdr = f16_header(m->d_X_planes16, N, D);
    a.gpart = (float*)(w + p.off_g); a.lpart = (double*)(w + p.off_l);
    a.S = S; a.S_pad = p.S_pad; a.N = N; a.D = D; a.n_dt = p.n_dt; a.DG = p.DG; a.R = p.R;
    a.flush_tiles = 128;                         // 8192 rows per TMEM accumulation chain (DESIGN: truncating accumulate)
    { const char* e = getenv("VB_REG_FLUSH"); if (e && atoi(e) > 0) a.flush_tiles = atoi(e); }
    a.inv_df = (float)(1.0 / m->df);
    // |psi / c| <= sqrt(df) / 2: largest power of two that keeps psi_scale * that below 2^15
    double ps = 1.0;
    while (ps * 2.0 * 0.5 * sqrt(m->df) < 32768.0 && ps < 65536.0) ps *= 2.0;
    a.psi_scale = (float)ps;
    static DeviceOnce configured;
    const int dev = current_device();
    if (configured.pending(dev)) {
        VB_CUDA_TRY(cudaFuncSetAttribute(reg_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFSmemBytes));
        configured.done(dev);
    }
    reg_fused_prep_kernel<<<grid_for_f(p.n_pad, 256, kNumSMs * 8), 256, 0, st>>>((const float*)m->d_y, N, p.n_pad, a.psi_scale,
                                                                                 (float*)(w + p.off_y));
    VB_LAUNCH_CHECK();
    reg_fused_kernel<<<(unsigned)(p.R * p.DG), kFThreads, kFSmemBytes, st>>>(tx, a);
    VB_LAUNCH_CHECK();
    const double inv_var = 1.0 / (m->prior_sd * m->prior_sd);
    reg_fused_final_kernel<<<grid_for_f(S * 32, 256, kNumSMs * 8), 256, 0, st>>>(
        a, inv_var, 0.5 * (m->df + 1.0), (m->df + 1.0) / m->df, out_logp, out_grad_or_null);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

}  // namespace vb

// ------------------------------------------------------------------------------------------ C ABI
extern "C" size_t vb_split_f16_bytes(long long rows, long long cols) {
    if (rows < 1 || cols < 1) return 0;
    return vb::split_f16_bytes(rows, cols);
}
extern "C" int vb_split_f16(const float* d_src, long long rows, long long cols, long long ld, void* d_dst_planes,
                            void* stream) {
    using namespace vb;
    if (!d_src || !d_dst_planes || rows < 1 || cols < 1 || ld < cols) return VB_ERR_ARG;
    if (reinterpret_cast<uintptr_t>(d_dst_planes) & 255u) return VB_ERR_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    F16Header* hdr = f16_header(d_dst_planes, rows, cols);
    VB_CUDA_TRY(cudaMemsetAsync(hdr, 0, sizeof(F16Header), st));
    f16_absmax_kernel<<<grid_for_f(rows * cols, 256, kNumSMs * 8), 256, 0, st>>>(d_src, rows, cols, ld, hdr);
    VB_LAUNCH_CHECK();
    const long long ldd = pad8(cols);
    f16_split_kernel<<<grid_for_f(rows * ldd, 256, kNumSMs * 8), 256, 0, st>>>(d_src, rows, cols, ld, (__half*)d_dst_planes,
                                                                             ldd, rows * ldd, hdr);
    VB_LAUNCH_CHECK();
    return VB_OK;
}
