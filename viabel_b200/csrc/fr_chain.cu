// The O(d^3) parameter chain of the full-rank Student-t family (kernel K7) as TWO persistent
// cooperative kernels, with no eigendecomposition and no library call:
//   forward   phi -> L (paragami free PSD pattern, viabel/vb.py:185-189) -> S = sqrtm(L L')      (vb.py:207)
//   backward  G_S -> Y solving S Y + Y S = sym(G_S) (autograd's sqrtm VJP; only Y + Y' is used) -> G_L = 2 Y L ->
//             d/dphi, and the assembly of the KLVI / CHIVI (value, gradient) vector                (vb.py:236-266)
// Everything is fp64 GEMMs on the fp64 tensor cores (mma.sync m8n8k4) inside one launch, separated by grid barriers:
//   * S is the symmetric polar factor of L (L = S Q, Q orthogonal): Q by a polynomial iteration on
//     X0 = L / (|L|_F |X0'X0|_F^1/2): six minimax quintics a X + X (b A + c A^2), A = X'X (kPolarSched: built for
//     sigma_min / sigma_max >= 1e-3, re-entered in cycles below that), then the cubic 1.5 X - 0.5 X A, which converges
//     quadratically; S = sym(L Q').
//   * Sylvester by squared-Smith doubling: with M = S + mu I, A0 = (S - mu I) M^-1, Y0 = 2 mu M^-1 G M^-1,
//     Y = sum_j A0^j Y0 A0^j:  Y <- Y + A Y A, A <- A^2;  M^-1 by Newton-Schulz (X <- 2 X - X M X) from
//     2 / (ub + 2 mu) I, ub >= lambda_max(S) left by the forward.
//   * Products whose result is symmetric (52 of 73 at C3) compute only the lower-triangular 32 x 32 blocks.
// Every iteration's stopping test is a Frobenius norm summed in fixed order from per-CTA partials, so all CTAs take
// the same decision and repeated runs are bit-identical.  Accuracy vs the eigh formulation: S 5e-15, Y 4e-13
// (config C3, cond(L) = 216); 73 GEMMs of 512^3 (9 + 7 + 10 iterations) instead of cuSOLVER syevd + 5 GEMMs.
#include <cstdlib>
#include "common.cuh"

namespace vb {

constexpr int kCConsumers = 256;           // MMA warps: two groups of 4 warps, each takes half of every k slab
constexpr int kCGroups = kCConsumers / 128;
constexpr int kCProducers = 128;           // copy warps: issue the cp.async of the operand slabs, nothing else
constexpr int kCT = kCConsumers + kCProducers;
constexpr int kCM = 64, kCN = 32, kCK = 32;    // CTA tile and k slab
constexpr int kCMaxGrid = 148;

#ifdef VB_CHAIN_TRACE
// debug build only: CTA 0 / thread 0 records (SM clock, tag) pairs; dumped by vb_fr_chain_trace_dump
__device__ unsigned long long g_chain_trace[8192];
__device__ unsigned int g_chain_trace_n;
#define CHAIN_TRACE(tag)                                                                           \
    do {                                                                                           \
        if (blockIdx.x == 0 && threadIdx.x == 0) {                                                 \
            const unsigned int n_ = g_chain_trace_n++;                                             \
            if (n_ < 8192) g_chain_trace[n_] = ((unsigned long long)clock64() << 8) | (tag);       \
        }                                                                                          \
    } while (0)
#else
#define CHAIN_TRACE(tag) do { } while (0)
#endif

// all CTAs of the (cooperative: co-resident) grid; `phase` counts the barriers this thread block has passed
__device__ __forceinline__ void grid_barrier(unsigned int* counter, unsigned int& phase) {
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int target = (phase + 1) * gridDim.x;
        // release: the CTA's writes (ordered before this thread by the bar.sync above) are visible before the count
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
        unsigned int v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
        } while (v < target);
    }
    ++phase;
    __syncthreads();
    CHAIN_TRACE(6);
}

__device__ __forceinline__ void dmma(double (&c)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

// C (d x d, ld = d) = alpha op(A) op(B) + beta E + gamma I, every CTA of the grid takes tiles t = blockIdx.x, + gridDim.x ..
// norm_mode: 0 none, 1 partial of |C - I|_F^2, 2 partial of |C|_F^2  -> returned (this CTA's sum, valid in every thread)
//
// Operand slabs (k = 32) travel global (L2) -> shared memory by cp.async into a ring of kCStages buffers, in their
// GLOBAL layout (no transposing store: the m8n8k4 fragments are 8-byte scalar loads, so either orientation is read in
// place; row pitches = 4 mod 16 doubles keep both conflict-free).  The copies are issued by four PRODUCER warps and
// complete on a per-stage mbarrier (cp.async.mbarrier.arrive); the eight MMA warps only wait on it, load fragments and
// issue DMMAs, and hand the stage back through a second mbarrier: no __syncthreads in the k loop.  The operand
// orientations are template parameters, so that every fragment offset is an immediate of its shared-memory load (as
// run-time strides they cost an integer multiply-add per load and 20 % of the kernel).  `ring` counts the slabs a CTA
// has consumed since the kernel began: stage and mbarrier parity follow from it, across tiles and across GEMMs.
constexpr int kCStages = 5;
constexpr int kCPitchK = kCK + 4;          // rows of k-values (A not transposed, B transposed)
constexpr int kCPitchM = kCM + 4;          // rows of m-values (A transposed)
constexpr int kCPitchN = kCN + 4;          // rows of n-values (B not transposed)
constexpr int kCABytes = (kCM * kCPitchK > kCK * kCPitchM ? kCM * kCPitchK : kCK * kCPitchM) * 8;
constexpr int kCBBytes = (kCN * kCPitchK > kCK * kCPitchN ? kCN * kCPitchK : kCK * kCPitchN) * 8;
struct GemmSmem {
    unsigned char stage[kCStages][kCABytes + kCBBytes];
    double xch[(kCGroups - 1) * 16 * 128];             // k-split exchange
    double red[kCT / 32];
    unsigned long long full[kCStages], empty[kCStages];
};                                                     // 151 KB: dynamic

__device__ __forceinline__ void chain_smem_init(GemmSmem& sm) {
    if (threadIdx.x == 0) {
        for (int s = 0; s < kCStages; ++s) {
            mbar_init(reinterpret_cast<uint64_t*>(&sm.full[s]), kCProducers);
            mbar_init(reinterpret_cast<uint64_t*>(&sm.empty[s]), kCConsumers / 32);
        }
        mbar_fence_init();
    }
    __syncthreads();
}
template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem_dst, const void* gsrc, bool valid) {
    const unsigned int dst = (unsigned int)__cvta_generic_to_shared(smem_dst);
    const int sz = valid ? BYTES : 0;                  // src-size 0: the destination is zero-filled
    if (BYTES == 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(gsrc), "r"(sz) : "memory");
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst), "l"(gsrc), "r"(sz) : "memory");
}
// the executing thread's earlier cp.async copies arrive on `bar` when they have landed (counted in the expected arrivals)
__device__ __forceinline__ void cp_async_arrive(unsigned long long* bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kCConsumers) : "memory"); }

// VEC = doubles per copy: 2 when d is even (every row start is 16-byte aligned), else 1
template <int VEC, bool ta, bool tb>
__device__ __noinline__ double chain_gemm_impl(GemmSmem& sm, unsigned int& ring, int d, const double* __restrict__ A,
                                               const double* __restrict__ B, double alpha, double beta,
                                               const double* E, double gamma, double* __restrict__ C, int norm_mode) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tiles_m = (d + kCM - 1) / kCM, tiles_n = (d + kCN - 1) / kCN;
    const int nslab = (d + kCK - 1) / kCK;
    double nrm = 0.0;
    if (tid >= kCConsumers) {
        // ---------------------------------------------------------------- producer warps
        const int ptid = tid - kCConsumers;
        for (int t = blockIdx.x; t < tiles_m * tiles_n; t += gridDim.x) {
            const int i0 = (t / tiles_n) * kCM, j0 = (t % tiles_n) * kCN;
            for (int slab = 0; slab < nslab; ++slab, ++ring) {
                const int st = ring % kCStages;
                mbar_wait(reinterpret_cast<uint64_t*>(&sm.empty[st]), ((ring / kCStages) & 1) ^ 1);
                const int k0 = slab * kCK;
                double* as = reinterpret_cast<double*>(sm.stage[st]);
                double* bs = reinterpret_cast<double*>(sm.stage[st] + kCABytes);
                constexpr int kChunksA = kCM * kCK / VEC, kChunksB = kCN * kCK / VEC;
#pragma unroll
                for (int c = ptid; c < kChunksA; c += kCProducers) {
                    // rows of the global layout: (m, k..) when not transposed, (k, m..) when transposed
                    int row, col;
                    if (!ta) { row = c / (kCK / VEC); col = (c % (kCK / VEC)) * VEC; }
                    else { row = c / (kCM / VEC); col = (c % (kCM / VEC)) * VEC; }
                    const int gr = (ta ? k0 : i0) + row, gc = (ta ? i0 : k0) + col;
                    const bool ok = gr < d && gc < d;
                    cp_async<VEC * 8>(as + row * (ta ? kCPitchM : kCPitchK) + col, ok ? A + (size_t)gr * d + gc : A, ok);
                }
#pragma unroll
                for (int c = ptid; c < kChunksB; c += kCProducers) {
                    int row, col;                        // (k, n..) or (n, k..)
                    if (!tb) { row = c / (kCN / VEC); col = (c % (kCN / VEC)) * VEC; }
                    else { row = c / (kCK / VEC); col = (c % (kCK / VEC)) * VEC; }
                    const int gr = (tb ? j0 : k0) + row, gc = (tb ? k0 : j0) + col;
                    const bool ok = gr < d && gc < d;
                    cp_async<VEC * 8>(bs + row * (tb ? kCPitchK : kCPitchN) + col, ok ? B + (size_t)gr * d + gc : B, ok);
                }
                cp_async_arrive(&sm.full[st]);
            }
        }
    } else {
        // ---------------------------------------------------------------- MMA warps
        const int grp = warp >> 2;                                      // k-split group of this warp
        const int wm = (warp & 1) * 32, wn = ((warp >> 1) & 1) * 16;
        // fragment addressing: A(m, k) at m * sa_m + k * sa_k, B(k, n) at k * sb_k + n * sb_n (doubles within a stage)
        constexpr int sa_m = ta ? 1 : kCPitchK, sa_k = ta ? kCPitchM : 1;
        constexpr int sb_k = tb ? 1 : kCPitchN, sb_n = tb ? kCPitchK : 1;
        const int a_off = (wm + (lane >> 2)) * sa_m + (lane & 3) * sa_k;
        const int b_off = (lane & 3) * sb_k + (wn + (lane >> 2)) * sb_n;
        for (int t = blockIdx.x; t < tiles_m * tiles_n; t += gridDim.x) {
            const int i0 = (t / tiles_n) * kCM, j0 = (t % tiles_n) * kCN;
            double acc[4][2][2];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 2; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
            CHAIN_TRACE(1);
            for (int slab = 0; slab < nslab; ++slab, ++ring) {
                const int st = ring % kCStages;
                mbar_wait(reinterpret_cast<uint64_t*>(&sm.full[st]), (ring / kCStages) & 1);
                if (slab == 0) CHAIN_TRACE(2);
                const double* as = reinterpret_cast<const double*>(sm.stage[st]) + a_off;
                const double* bs = reinterpret_cast<const double*>(sm.stage[st] + kCABytes) + b_off;
#pragma unroll
                for (int kk = grp * 4; kk < kCK; kk += 4 * kCGroups) {
                    double af[4], bf[2];
#pragma unroll
                    for (int a = 0; a < 4; ++a) af[a] = as[kk * sa_k + a * 8 * sa_m];
#pragma unroll
                    for (int b = 0; b < 2; ++b) bf[b] = bs[kk * sb_k + b * 8 * sb_n];
#pragma unroll
                    for (int a = 0; a < 4; ++a)
#pragma unroll
                        for (int b = 0; b < 2; ++b) dmma(acc[a][b], af[a], bf[b]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(reinterpret_cast<uint64_t*>(&sm.empty[st]));
            }
            CHAIN_TRACE(3);
            // the k-split groups meet in shared memory; group 0 adds in fixed order
            {
                const int t128 = tid & 127;
                if (grp != 0) {
#pragma unroll
                    for (int a = 0; a < 4; ++a)
#pragma unroll
                        for (int b = 0; b < 2; ++b)
#pragma unroll
                            for (int c = 0; c < 2; ++c)
                                sm.xch[((grp - 1) * 16 + (a * 2 + b) * 2 + c) * 128 + t128] = acc[a][b][c];
                }
                consumer_sync();
                if (grp == 0) {
#pragma unroll
                    for (int g = 0; g < kCGroups - 1; ++g)
#pragma unroll
                        for (int a = 0; a < 4; ++a)
#pragma unroll
                            for (int b = 0; b < 2; ++b)
#pragma unroll
                                for (int c = 0; c < 2; ++c) acc[a][b][c] += sm.xch[(g * 16 + (a * 2 + b) * 2 + c) * 128 + t128];
                }
                consumer_sync();
            }
            CHAIN_TRACE(4);
            // epilogue (group 0): thread holds C(row = lane / 4, cols = 2 (lane % 4), + 1) of each 8 x 8 block.  All E
            // values are requested before the first store (E may alias nothing the compiler can prove: interleaved,
            // the 16 load -> store pairs ran as a chain of L2 round trips, 3.7 K cycles per GEMM)
            if (grp == 0) {
                double ev[4][2][2];
                const bool use_e = beta != 0.0;
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 2; ++b) {
                        const int gi = i0 + wm + a * 8 + (lane >> 2), gj = j0 + wn + b * 8 + 2 * (lane & 3);
                        ev[a][b][0] = ev[a][b][1] = 0.0;
                        if (use_e && gi < d) {
                            if (VEC == 2) {
                                if (gj < d) {
                                    const double2 q = *reinterpret_cast<const double2*>(E + (size_t)gi * d + gj);
                                    ev[a][b][0] = q.x; ev[a][b][1] = q.y;
                                }
                            } else {
                                if (gj < d) ev[a][b][0] = E[(size_t)gi * d + gj];
                                if (gj + 1 < d) ev[a][b][1] = E[(size_t)gi * d + gj + 1];
                            }
                        }
                    }
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 2; ++b) {
                        const int gi = i0 + wm + a * 8 + (lane >> 2), gj = j0 + wn + b * 8 + 2 * (lane & 3);
                        double v[2];
#pragma unroll
                        for (int c = 0; c < 2; ++c) {
                            v[c] = fma(beta, ev[a][b][c], alpha * acc[a][b][c]);
                            if (gi == gj + c) v[c] += gamma;
                            if (norm_mode && gi < d && gj + c < d) {
                                const double r = (norm_mode == 1 && gi == gj + c) ? v[c] - 1.0 : v[c];
                                nrm = fma(r, r, nrm);
                            }
                        }
                        if (gi < d) {
                            if (VEC == 2) {
                                if (gj < d) *reinterpret_cast<double2*>(C + (size_t)gi * d + gj) = make_double2(v[0], v[1]);
                            } else {
                                if (gj < d) C[(size_t)gi * d + gj] = v[0];
                                if (gj + 1 < d) C[(size_t)gi * d + gj + 1] = v[1];
                            }
                        }
                    }
            }
        }
    }
    CHAIN_TRACE(5);
    if (!norm_mode) return 0.0;
    nrm = warp_sum(nrm);
    __syncthreads();
    if (lane == 0) sm.red[warp] = nrm;
    __syncthreads();
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kCT / 32; ++w) s += sm.red[w];
    return s;
}

// The same product when the RESULT is symmetric (X'X, A A, products of commuting symmetric matrices, A Y A): only the
// 32 x 32 blocks on and below the diagonal are computed (nb (nb + 1) / 2 tiles: 136 at d = 512, one wave) and each
// off-diagonal block is also stored transposed.  E, if given, must be symmetric too (it is read at the lower position).
// Where symmetry holds only mathematically (M X with X a polynomial in M, A Y A) the mirror replaces an O(eps) rounding
// asymmetry.  Same ring, barriers and fragment scheme as above with 16 x 16 warp tiles.
constexpr int kCSym = 32;                  // block size of the symmetric variant (kCSym == kCK: all slab pitches are 36)
static_assert(kCSym == kCK && kCSym == kCN, "symmetric tiles reuse the B-slab geometry");
template <int VEC, bool ta, bool tb>
__device__ __noinline__ double chain_gemm_sym_impl(GemmSmem& sm, unsigned int& ring, int d, const double* __restrict__ A,
                                                   const double* __restrict__ B, double alpha, double beta,
                                                   const double* E, double gamma, double* __restrict__ C, int norm_mode) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nb = (d + kCSym - 1) / kCSym;
    const int ntiles = nb * (nb + 1) / 2;
    const int nslab = (d + kCK - 1) / kCK;
    constexpr int P = kCSym + 4;
    auto tile_of = [&](int t, int& bi, int& bj) {        // row-major lower triangle: t -> (bi, bj <= bi)
        bi = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
        while (bi * (bi + 1) / 2 > t) --bi;
        while ((bi + 1) * (bi + 2) / 2 <= t) ++bi;
        bj = t - bi * (bi + 1) / 2;
    };
    double nrm = 0.0;
    if (tid >= kCConsumers) {
        const int ptid = tid - kCConsumers;
        for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
            int bi, bj;
            tile_of(t, bi, bj);
            const int i0 = bi * kCSym, j0 = bj * kCSym;
            for (int slab = 0; slab < nslab; ++slab, ++ring) {
                const int st = ring % kCStages;
                mbar_wait(reinterpret_cast<uint64_t*>(&sm.empty[st]), ((ring / kCStages) & 1) ^ 1);
                const int k0 = slab * kCK;
                double* as = reinterpret_cast<double*>(sm.stage[st]);
                double* bs = reinterpret_cast<double*>(sm.stage[st] + kCABytes);
                constexpr int kChunks = kCSym * kCK / VEC;
#pragma unroll
                for (int c = ptid; c < kChunks; c += kCProducers) {
                    const int row = c / (kCSym / VEC), col = (c % (kCSym / VEC)) * VEC;
                    {   // A: rows (m, k..) or (k, m..)
                        const int gr = (ta ? k0 : i0) + row, gc = (ta ? i0 : k0) + col;
                        const bool ok = gr < d && gc < d;
                        cp_async<VEC * 8>(as + row * P + col, ok ? A + (size_t)gr * d + gc : A, ok);
                    }
                    {   // B: rows (k, n..) or (n, k..)
                        const int gr = (tb ? j0 : k0) + row, gc = (tb ? k0 : j0) + col;
                        const bool ok = gr < d && gc < d;
                        cp_async<VEC * 8>(bs + row * P + col, ok ? B + (size_t)gr * d + gc : B, ok);
                    }
                }
                cp_async_arrive(&sm.full[st]);
            }
        }
    } else {
        const int grp = warp >> 2;
        const int wm = (warp & 1) * 16, wn = ((warp >> 1) & 1) * 16;
        constexpr int sa_m = ta ? 1 : P, sa_k = ta ? P : 1;
        constexpr int sb_k = tb ? 1 : P, sb_n = tb ? P : 1;
        const int a_off = (wm + (lane >> 2)) * sa_m + (lane & 3) * sa_k;
        const int b_off = (lane & 3) * sb_k + (wn + (lane >> 2)) * sb_n;
        for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
            int bi, bj;
            tile_of(t, bi, bj);
            const int i0 = bi * kCSym, j0 = bj * kCSym;
            double acc[2][2][2];
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int b = 0; b < 2; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
            for (int slab = 0; slab < nslab; ++slab, ++ring) {
                const int st = ring % kCStages;
                mbar_wait(reinterpret_cast<uint64_t*>(&sm.full[st]), (ring / kCStages) & 1);
                const double* as = reinterpret_cast<const double*>(sm.stage[st]) + a_off;
                const double* bs = reinterpret_cast<const double*>(sm.stage[st] + kCABytes) + b_off;
#pragma unroll
                for (int kk = grp * 4; kk < kCK; kk += 4 * kCGroups) {
                    double af[2], bf[2];
#pragma unroll
                    for (int a = 0; a < 2; ++a) af[a] = as[kk * sa_k + a * 8 * sa_m];
#pragma unroll
                    for (int b = 0; b < 2; ++b) bf[b] = bs[kk * sb_k + b * 8 * sb_n];
#pragma unroll
                    for (int a = 0; a < 2; ++a)
#pragma unroll
                        for (int b = 0; b < 2; ++b) dmma(acc[a][b], af[a], bf[b]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(reinterpret_cast<uint64_t*>(&sm.empty[st]));
            }
            {
                const int t128 = tid & 127;
                if (grp != 0) {
#pragma unroll
                    for (int a = 0; a < 2; ++a)
#pragma unroll
                        for (int b = 0; b < 2; ++b)
#pragma unroll
                            for (int c = 0; c < 2; ++c)
                                sm.xch[((grp - 1) * 8 + (a * 2 + b) * 2 + c) * 128 + t128] = acc[a][b][c];
                }
                consumer_sync();
                if (grp == 0) {
#pragma unroll
                    for (int g = 0; g < kCGroups - 1; ++g)
#pragma unroll
                        for (int a = 0; a < 2; ++a)
#pragma unroll
                            for (int b = 0; b < 2; ++b)
#pragma unroll
                                for (int c = 0; c < 2; ++c) acc[a][b][c] += sm.xch[(g * 8 + (a * 2 + b) * 2 + c) * 128 + t128];
                }
                consumer_sync();
            }
            if (grp == 0) {
                const bool offdiag = bi != bj;
                double ev[2][2][2];
                const bool use_e = beta != 0.0;
#pragma unroll
                for (int a = 0; a < 2; ++a)
#pragma unroll
                    for (int b = 0; b < 2; ++b) {
                        const int gi = i0 + wm + a * 8 + (lane >> 2), gj = j0 + wn + b * 8 + 2 * (lane & 3);
                        ev[a][b][0] = ev[a][b][1] = 0.0;
                        if (use_e && gi < d) {
                            if (VEC == 2) {
                                if (gj < d) {
                                    const double2 q = *reinterpret_cast<const double2*>(E + (size_t)gi * d + gj);
                                    ev[a][b][0] = q.x; ev[a][b][1] = q.y;
                                }
                            } else {
                                if (gj < d) ev[a][b][0] = E[(size_t)gi * d + gj];
                                if (gj + 1 < d) ev[a][b][1] = E[(size_t)gi * d + gj + 1];
                            }
                        }
                    }
#pragma unroll
                for (int a = 0; a < 2; ++a)
#pragma unroll
                    for (int b = 0; b < 2; ++b) {
                        const int gi = i0 + wm + a * 8 + (lane >> 2), gj = j0 + wn + b * 8 + 2 * (lane & 3);
                        double v[2];
#pragma unroll
                        for (int c = 0; c < 2; ++c) {
                            v[c] = fma(beta, ev[a][b][c], alpha * acc[a][b][c]);
                            if (gi == gj + c) v[c] += gamma;
                            if (norm_mode && gi < d && gj + c < d) {
                                const double r = (norm_mode == 1 && gi == gj + c) ? v[c] - 1.0 : v[c];
                                nrm = fma(offdiag ? 2.0 * r : r, r, nrm);
                            }
                        }
                        if (gi < d) {
                            if (VEC == 2) {
                                if (gj < d) *reinterpret_cast<double2*>(C + (size_t)gi * d + gj) = make_double2(v[0], v[1]);
                            } else {
                                if (gj < d) C[(size_t)gi * d + gj] = v[0];
                                if (gj + 1 < d) C[(size_t)gi * d + gj + 1] = v[1];
                            }
                            if (offdiag) {
                                if (gj < d) C[(size_t)gj * d + gi] = v[0];
                                if (gj + 1 < d) C[(size_t)(gj + 1) * d + gi] = v[1];
                            }
                        }
                    }
            }
        }
    }
    if (!norm_mode) return 0.0;
    nrm = warp_sum(nrm);
    __syncthreads();
    if (lane == 0) sm.red[warp] = nrm;
    __syncthreads();
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kCT / 32; ++w) s += sm.red[w];
    return s;
}

template <int VEC>
__device__ __forceinline__ double chain_gemm_v(GemmSmem& sm, unsigned int& ring, int d, bool ta, const double* __restrict__ A,
                                               bool tb, const double* __restrict__ B, double alpha, double beta,
                                               const double* E, double gamma, double* __restrict__ C, int norm_mode, bool sym) {
    if (sym) {
        if (!ta && !tb) return chain_gemm_sym_impl<VEC, false, false>(sm, ring, d, A, B, alpha, beta, E, gamma, C, norm_mode);
        if (ta && !tb) return chain_gemm_sym_impl<VEC, true, false>(sm, ring, d, A, B, alpha, beta, E, gamma, C, norm_mode);
    }
    if (!ta && !tb) return chain_gemm_impl<VEC, false, false>(sm, ring, d, A, B, alpha, beta, E, gamma, C, norm_mode);
    if (ta && !tb) return chain_gemm_impl<VEC, true, false>(sm, ring, d, A, B, alpha, beta, E, gamma, C, norm_mode);
    if (!ta && tb) return chain_gemm_impl<VEC, false, true>(sm, ring, d, A, B, alpha, beta, E, gamma, C, norm_mode);
    return chain_gemm_impl<VEC, true, true>(sm, ring, d, A, B, alpha, beta, E, gamma, C, norm_mode);
}
// sym: the caller asserts that the result is symmetric (see chain_gemm_sym_impl)
__device__ __forceinline__ double chain_gemm(GemmSmem& sm, unsigned int& ring, int d, bool ta, const double* __restrict__ A,
                                             bool tb, const double* __restrict__ B, double alpha, double beta,
                                             const double* E, double gamma, double* __restrict__ C, int norm_mode,
                                             bool sym = false) {
    if ((d & 1) == 0) return chain_gemm_v<2>(sm, ring, d, ta, A, tb, B, alpha, beta, E, gamma, C, norm_mode, sym);
    return chain_gemm_v<1>(sm, ring, d, ta, A, tb, B, alpha, beta, E, gamma, C, norm_mode, sym);
}

// sum of the per-CTA partials in fixed order (every thread of every CTA gets the same value)
__device__ __forceinline__ double sum_partials(const volatile double* parts) {
    // lane l adds parts[l], parts[l + 32], ... then a fixed butterfly: the same tree in every warp of every CTA
    double s = 0.0;
    for (unsigned int b = threadIdx.x & 31; b < gridDim.x; b += 32) s += parts[b];
    return warp_sum(s);
}

// Minimax odd quintics a x + b x^3 + c x^5 for the polar iteration: polynomial k is the best approximation of 1 on the
// image [1 - t, 1 + t] of its predecessor, starting from [1e-3, 1] (scripts/polar_schedule.py; errors t = 0.9915,
// 0.9646, 0.8598, 0.5459, 0.1134, 9.2e-4).
__constant__ double kPolarSched[6][3] = {
    {8.470329, -25.108079, 18.629279}, {4.182834, -3.108701, 0.580607}, {3.961856, -2.954063, 0.562976},
    {3.286584, -2.464719, 0.507358},   {2.273748, -1.644659, 0.416191}, {1.888716, -1.265157, 0.376519}};

struct ChainWs {
    double *X0, *X1, *T0, *T1, *T2, *T3;        // six d x d scratch matrices
    double* parts;                              // [3][grid] norm partials (rotating: a slot is rewritten two barriers later)
    unsigned int* barrier;
    int* info;                                  // [0] polar iterations [1] inverse iterations [2] smith iterations [3] status
                                                // [4] GEMMs executed by the forward [5] by the backward [6], [7] symmetric ones among them
};

// ------------------------------------------------------------------------------------------ forward
struct FwdArgs {
    const double* vp;          // (d + d (d + 1) / 2) variational parameters
    int d, mode;               // mode 0 = sqrtm (reference), 1 = chol (x = mu + L ztilde)
    double* L;                 // out (d, d)
    double* S;                 // out (d, d) fp64: row-form scale matrix (x = mu + ztilde S): sqrtm(Sigma) or L'
    float* S32;                // out (d, d) fp32 copy or null
    double* scal;              // out [0] entropy = sum phi_ii  [1] |L|_F^2  [2] upper bound of sigma_max(L) (sqrtm mode)
    ChainWs ws;
};

__global__ void __launch_bounds__(kCT)
fr_chain_forward_kernel(const FwdArgs a) {
    extern __shared__ __align__(16) unsigned char chain_smem[];
    GemmSmem& sm = *reinterpret_cast<GemmSmem*>(chain_smem);
    const int d = a.d;
    const long long dd = (long long)d * d;
    const long long gtid = blockIdx.x * (long long)blockDim.x + threadIdx.x, gth = (long long)gridDim.x * blockDim.x;
    unsigned int phase = 0, ring = 0;
    chain_smem_init(sm);
    double* parts = a.ws.parts;
    int slot = 0;
    auto next_slot = [&]() { double* p = parts + (size_t)slot * gridDim.x; slot = (slot + 1) % 3; return p; };
    // ---- fold: L = unvec_tril_rowmajor(phi) with exp on the diagonal; partial sums of |L|_F^2 and of the log diagonal
    double f2 = 0.0, ent = 0.0;
    for (long long e = gtid; e < dd; e += gth) {
        const int r = (int)(e / d), c = (int)(e - (long long)r * d);
        double v = 0.0;
        if (c <= r) {
            v = a.vp[d + (long long)r * (r + 1) / 2 + c];
            if (c == r) { ent += v; v = exp(v); }
        }
        a.L[e] = v;
        f2 = fma(v, v, f2);
    }
    {
        f2 = warp_sum(f2); ent = warp_sum(ent);
        __syncthreads();
        double* red2 = sm.xch;
        if ((threadIdx.x & 31) == 0) { sm.red[threadIdx.x >> 5] = f2; red2[threadIdx.x >> 5] = ent; }
        __syncthreads();
        double s = 0.0, t = 0.0;
        for (int w = 0; w < kCT / 32; ++w) { s += sm.red[w]; t += red2[w]; }
        __syncthreads();
        double* p = next_slot();
        double* p2 = next_slot();
        if (threadIdx.x == 0) { p[blockIdx.x] = s; p2[blockIdx.x] = t; }
        grid_barrier(a.ws.barrier, phase);
        f2 = sum_partials(p);
        ent = sum_partials(p2);
        if (gtid == 0) { a.scal[0] = ent; a.scal[1] = f2; }
    }
    if (a.mode == 1) {                                   // chol: S = L' (row form), no iteration
        for (long long e = gtid; e < dd; e += gth) {
            const int r = (int)(e / d), c = (int)(e - (long long)r * d);
            const double v = a.L[(size_t)c * d + r];
            a.S[e] = v;
            if (a.S32) a.S32[e] = (float)v;
        }
        if (gtid == 0) { a.ws.info[0] = 0; a.ws.info[3] = 0; a.ws.info[4] = 0; a.ws.info[6] = 0; a.scal[2] = 0.0; }
        return;
    }
    // ---- polar factor of L: X0 = L / |L|_F, then X0 / s with s^2 = |X0'X0|_F >= sigma_max(X0)^2 (Frobenius alone
    // leaves the spectrum a factor ~sqrt(d) below 1; this bound, free with the first product, leaves d^(1/4))
    const double inv_f = 1.0 / sqrt(f2);
    double* X = a.ws.X0;
    double* Xn = a.ws.X1;
    for (long long e = gtid; e < dd; e += gth) X[e] = a.L[e] * inv_f;
    grid_barrier(a.ws.barrier, phase);
    const double tol2 = 16.0 * (double)d * 1e-28;        // |X'X - I|_F < 4 sqrt(d) 1e-14
    int its = 0, status = 1, ng = 0, nsym = 0;
    double ps = 1.0;                                     // pending scale: the iterate is X / ps, its Gram matrix T0 / ps^2
    double err2 = INFINITY;
    bool finishing = false;
    for (int it = 0; it < 80; ++it) {
        double* p = next_slot();
        const double part = chain_gemm(sm, ring, d, true, X, false, X, 1.0, 0.0, nullptr, 0.0, a.ws.T0, it == 0 ? 2 : 1, true);   // A = X'X
        ++ng; ++nsym;
        if (threadIdx.x == 0) p[blockIdx.x] = part;
        grid_barrier(a.ws.barrier, phase);
        if (it == 0) {
            ps = sqrt(sqrt(sum_partials(p)) * (1.0 + 1e-12));
            if (gtid == 0) a.scal[2] = ps * sqrt(f2);    // >= sigma_max(L) = lambda_max(S): the backward's inverse starts from it
        } else {
            err2 = sum_partials(p);
            if (err2 < tol2) { status = 0; break; }
        }
        // Schedule: six minimax quintics (kPolarSched: [l, 1] -> 1 +- 9e-4 for l = 1e-3), then the cubic Newton-Schulz
        // step; a spectrum reaching below 1e-3 re-enters the schedule at its second polynomial (valid on [0, 1.99]) in
        // cycles of five (x234 per cycle) until the error seen at a cycle boundary is small.
        if (it >= 6 && (it - 6) % 5 == 0 && err2 <= 0.25) finishing = true;
        if (err2 <= 0.0025) finishing = true;
        if (!finishing) {
            const double* q = kPolarSched[it < 6 ? it : 1 + (it - 6) % 5];
            const double i2 = 1.0 / (ps * ps);
            // B = c A^2 + b A, X <- a X + X B with the pending scale folded into the coefficients
            chain_gemm(sm, ring, d, false, a.ws.T0, false, a.ws.T0, q[2] * i2 * i2, q[1] * i2, a.ws.T0, 0.0, a.ws.T1, 0, true);
            grid_barrier(a.ws.barrier, phase);
            chain_gemm(sm, ring, d, false, X, false, a.ws.T1, 1.0 / ps, q[0] / ps, X, 0.0, Xn, 0);
            ng += 2; ++nsym;
        } else {
            chain_gemm(sm, ring, d, false, X, false, a.ws.T0, -0.5, 1.5, X, 0.0, Xn, 0);                          // 1.5 X - 0.5 X A
            ++ng;
        }
        ps = 1.0;
        grid_barrier(a.ws.barrier, phase);
        double* t = X; X = Xn; Xn = t;
        ++its;
        if (finishing && err2 < 1e-14) { status = 0; break; }   // quadratic: the next error is ~err^2 < 1e-14, no product spent on seeing it
    }
    // ---- S = sym(L Q'), Q = X
    chain_gemm(sm, ring, d, false, a.L, true, X, 1.0, 0.0, nullptr, 0.0, a.ws.T0, 0);
    ++ng;
    grid_barrier(a.ws.barrier, phase);
    for (long long e = gtid; e < dd; e += gth) {
        const int r = (int)(e / d), c = (int)(e - (long long)r * d);
        const double v = 0.5 * (a.ws.T0[e] + a.ws.T0[(size_t)c * d + r]);
        a.S[e] = v;
        if (a.S32) a.S32[e] = (float)v;
    }
    if (gtid == 0) { a.ws.info[0] = its; a.ws.info[3] = status; a.ws.info[4] = ng; a.ws.info[6] = nsym; }
}

// ------------------------------------------------------------------------------------------ backward
struct BwdArgs {
    int d, mode, objective;    // objective: 0 KLVI, 1 CHIVI, 2 KLVI path-derivative value
    double alpha, n_total;
    const double* L;           // (d, d)
    const double* S;           // (d, d)
    const double* GS;          // (d, d) G_S = Ztilde' diag(w) G
    const double* gmu;         // (d)    sum_i w_i g_i
    const double* rows;        // [max lw, sum w, sum log p, sum lw] (+ [4] global sample count, read when n_total == 0)
    const double* scal;        // forward scalars: [0] entropy, [1] |L|_F^2
    double* out;               // (1 + P): value, gradient
    ChainWs ws;
};

__global__ void __launch_bounds__(kCT)
fr_chain_backward_kernel(const BwdArgs a) {
    extern __shared__ __align__(16) unsigned char chain_smem[];
    GemmSmem& sm = *reinterpret_cast<GemmSmem*>(chain_smem);
    const int d = a.d;
    const long long dd = (long long)d * d;
    const long long gtid = blockIdx.x * (long long)blockDim.x + threadIdx.x, gth = (long long)gridDim.x * blockDim.x;
    unsigned int phase = 0, ring = 0;
    chain_smem_init(sm);
    double* parts = a.ws.parts;
    int slot = 0;
    auto next_slot = [&]() { double* p = parts + (size_t)slot * gridDim.x; slot = (slot + 1) % 3; return p; };
    const double* GL = nullptr;          // d f / d L as (d, d) row-major
    bool gl_transposed = false;
    int inv_its = 0, smith_its = 0, status = 0, ng = 0, nsym = 0;
    if (a.mode == 1) {
        GL = a.GS;                       // chol: G_L = G_S' (x = mu + L ztilde)
        gl_transposed = true;
    } else {
        // ---- M^-1, M = S + mu I, mu = rms eigenvalue of S = sqrt(|L|_F^2 / d).  S is positive semi-definite and the
        // forward left ub >= lambda_max(S) in scal[2], so spec(M) lies in [mu, ub + mu] and X0 = 2 / (ub + 2 mu) I
        // starts Newton-Schulz at |I - M X0|_2 <= ub / (ub + 2 mu) < 1 (the classic I / |M|_1 start cost two more steps)
        const double mu = sqrt(a.scal[1] / (double)d);
        double* Mx = a.ws.T3;            // M
        const double ub = a.scal[2];
        double* X = a.ws.X0;
        double* Xn = a.ws.X1;
        const double x0 = 2.0 / (ub + 2.0 * mu);
        for (long long e = gtid; e < dd; e += gth) {
            const int r = (int)(e / d), c = (int)(e - (long long)r * d);
            Mx[e] = a.S[e] + (r == c ? mu : 0.0);
            X[e] = (r == c) ? x0 : 0.0;
            // only Y + Y' is used below and the equation is linear: solve for the symmetric part of G, then every
            // product of the doubling except A Y is symmetric
            a.ws.T2[e] = 0.5 * (a.GS[e] + a.GS[(size_t)c * d + r]);
        }
        grid_barrier(a.ws.barrier, phase);
        const double tol2 = 16.0 * (double)d * 1e-28;
        status = 2;
        for (int it = 0; it < 60; ++it) {
            double* p = next_slot();
            const double part = chain_gemm(sm, ring, d, false, Mx, false, X, 1.0, 0.0, nullptr, 0.0, a.ws.T0, 1, true);   // R = M X
            ++ng; ++nsym;
            if (threadIdx.x == 0) p[blockIdx.x] = part;
            grid_barrier(a.ws.barrier, phase);
            const double err2 = sum_partials(p);
            if (err2 < tol2) { status = 0; break; }
            chain_gemm(sm, ring, d, false, X, false, a.ws.T0, -1.0, 2.0, X, 0.0, Xn, 0, true);                     // 2 X - X R
            ++ng; ++nsym;
            grid_barrier(a.ws.barrier, phase);
            double* t = X; X = Xn; Xn = t;
            ++inv_its;
            if (err2 < 1e-14) { status = 0; break; }     // the residual squares: now < 1e-14, no product spent on seeing it
        }
        // ---- A0 = S M^-1 - mu M^-1,  Y0 = 2 mu M^-1 G M^-1
        double* Aq = a.ws.T1;
        double* Aq2 = a.ws.T3;           // M is no longer needed after A0
        double* Y = a.ws.T2;
        double* Yn = Xn;                 // the spare inverse buffer
        chain_gemm(sm, ring, d, false, X, false, a.ws.T2, 1.0, 0.0, nullptr, 0.0, a.ws.T0, 0);                     // T = M^-1 sym(G)
        grid_barrier(a.ws.barrier, phase);
        chain_gemm(sm, ring, d, false, a.ws.T0, false, X, 2.0 * mu, 0.0, nullptr, 0.0, Y, 0, true);                // Y0
        ng += 3; nsym += 2;
        {
            double* p = next_slot();
            const double part = chain_gemm(sm, ring, d, false, a.S, false, X, 1.0, -mu, X, 0.0, Aq, 2, true);      // A0, |A0|_F^2
            if (threadIdx.x == 0) p[blockIdx.x] = part;
            grid_barrier(a.ws.barrier, phase);
            double na2 = sum_partials(p);
            // ---- doubling: Y <- Y + A Y A, A <- A^2.  What is still missing after a step is A_new Y A_new, relatively
            // |A_new|_2^2 <= |A_new|_F^2: stop once that is below 1e-17
            for (int it = 0; it < 60 && na2 > 1e-17; ++it) {
                chain_gemm(sm, ring, d, false, Aq, false, Y, 1.0, 0.0, nullptr, 0.0, a.ws.T0, 0);                  // T = A Y
                double* p2 = next_slot();
                const double part2 = chain_gemm(sm, ring, d, false, Aq, false, Aq, 1.0, 0.0, nullptr, 0.0, Aq2, 2, true);   // A^2 (independent)
                if (threadIdx.x == 0) p2[blockIdx.x] = part2;
                grid_barrier(a.ws.barrier, phase);
                chain_gemm(sm, ring, d, false, a.ws.T0, false, Aq, 1.0, 1.0, Y, 0.0, Yn, 0, true);                 // Y + T A
                grid_barrier(a.ws.barrier, phase);
                na2 = sum_partials(p2);
                double* t = Aq; Aq = Aq2; Aq2 = t;
                t = Y; Y = Yn; Yn = t;
                ++smith_its;
                ng += 3; nsym += 2;
            }
            if (na2 > 1e-17 && status == 0) status = 3;
        }
        // ---- G_L = (Y + Y') L = 2 Y L
        chain_gemm(sm, ring, d, false, Y, false, a.L, 2.0, 0.0, nullptr, 0.0, a.ws.T0, 0);      // Y is symmetric: Y + Y' = 2 Y
        ++ng;
        grid_barrier(a.ws.barrier, phase);
        GL = a.ws.T0;
    }
    // ---- assemble (value, gradient): vb.py:240-243 (KLVI), :261-264 (CHIVI)
    const double nt = a.n_total > 0.0 ? a.n_total : a.rows[4];      // 0: the global count travelled with the sums (rows[4])
    const double entropy = a.scal[0];
    const double sw = a.rows[1];
    const bool chivi = a.objective == 1;
    const double gscale = chivi ? a.alpha / nt : -1.0 / nt;
    if (gtid == 0) {
        double v;
        if (chivi) v = log(sw / nt) / a.alpha + a.rows[0];
        else if (a.objective == 0) v = -(entropy + a.rows[2] / nt);
        else v = -(a.rows[3] / nt);
        a.out[0] = v;
        a.ws.info[1] = inv_its; a.ws.info[2] = smith_its; a.ws.info[5] = ng; a.ws.info[7] = nsym;
        if (status) a.ws.info[3] = status;
    }
    for (long long j = gtid; j < d; j += gth) a.out[1 + j] = gscale * a.gmu[j];
    const long long ntri = (long long)d * (d + 1) / 2;
    for (long long e = gtid; e < ntri; e += gth) {
        // row-major lower triangle: e -> (r, c)
        long long r = (long long)((sqrt(8.0 * (double)e + 1.0) - 1.0) * 0.5);
        while (r * (r + 1) / 2 > e) --r;
        while ((r + 1) * (r + 2) / 2 <= e) ++r;
        const long long c = e - r * (r + 1) / 2;
        double g = gl_transposed ? GL[(size_t)c * d + r] : GL[(size_t)r * d + c];
        if (r == c) {
            g *= a.L[(size_t)r * d + r];                     // chain through exp on the diagonal
            g += chivi ? sw : (double)nt;                    // -log q term: + sum w (CHIVI), -1 after the KLVI scaling
        }
        a.out[1 + d + e] = gscale * g;
    }
}

static cudaError_t chain_configure() {
    static DeviceOnce once;
    const int dev = current_device();
    if (!once.pending(dev)) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(fr_chain_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(GemmSmem));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(fr_chain_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(GemmSmem));
    if (e == cudaSuccess) once.done(dev);
    return e;
}
static int chain_grid(int d) {
    int tiles = ((d + kCM - 1) / kCM) * ((d + kCN - 1) / kCN);
    const int nb = (d + kCSym - 1) / kCSym;
    if (nb * (nb + 1) / 2 > tiles) tiles = nb * (nb + 1) / 2;     // the symmetric products' tile count
    int g = device_sm_count();
    if (g > kCMaxGrid) g = kCMaxGrid;
    if (tiles < g) g = tiles < 8 ? 8 : tiles;              // a few CTAs even for tiny d (elementwise phases)
    if (g > device_sm_count()) g = device_sm_count();
    return g;
}
static size_t chain_ws_bytes(int d) {
    return align_up((size_t)6 * d * d * 8, 256) + align_up((size_t)3 * kCMaxGrid * 8, 256) + 256 + 256;
}
static ChainWs chain_ws(void* ws, int d) {
    char* p = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
    const size_t m = (size_t)d * d * 8;
    ChainWs w;
    w.X0 = (double*)p; w.X1 = (double*)(p + m); w.T0 = (double*)(p + 2 * m); w.T1 = (double*)(p + 3 * m);
    w.T2 = (double*)(p + 4 * m); w.T3 = (double*)(p + 5 * m);
    p += align_up(6 * m, 256);
    w.parts = (double*)p; p += align_up((size_t)3 * kCMaxGrid * 8, 256);
    w.barrier = (unsigned int*)p; p += 256;
    w.info = (int*)p;
    return w;
}

}  // namespace vb

using namespace vb;

#ifdef VB_CHAIN_TRACE
extern "C" int vb_fr_chain_trace_dump(unsigned long long* h_out, int cap, int reset) {
    unsigned int n = 0;
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(&n, g_chain_trace_n, sizeof(n));
    if ((int)n > cap) n = cap;
    if (n > 8192) n = 8192;
    cudaMemcpyFromSymbol(h_out, g_chain_trace, (size_t)n * 8);
    if (reset) { const unsigned int z = 0; cudaMemcpyToSymbol(g_chain_trace_n, &z, sizeof(z)); }
    return (int)n;
}
#endif

extern "C" size_t vb_fr_chain_workspace_bytes(int dim) { return dim < 1 ? 0 : chain_ws_bytes(dim) + 256; }

extern "C" int vb_fr_chain_forward(const double* d_var_param, int dim, int scale_mode, double* d_out_L, double* d_out_S,
                                   float* d_out_S32_or_null, double* d_out_scal, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_var_param || !d_out_L || !d_out_S || !d_out_scal || !d_ws || dim < 1 || (scale_mode != 0 && scale_mode != 1))
        return VB_ERR_ARG;
    if (ws_bytes < vb_fr_chain_workspace_bytes(dim)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    FwdArgs a;
    a.vp = d_var_param; a.d = dim; a.mode = scale_mode; a.L = d_out_L; a.S = d_out_S; a.S32 = d_out_S32_or_null;
    a.scal = d_out_scal; a.ws = chain_ws(d_ws, dim);
    VB_CUDA_TRY(cudaMemsetAsync(a.ws.barrier, 0, 256 + 256, st));
    void* args[] = {(void*)&a};
    VB_CUDA_TRY(chain_configure());
    VB_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)fr_chain_forward_kernel, dim3(chain_grid(dim)), dim3(kCT), args,
                                            sizeof(GemmSmem), st));
    return VB_OK;
}

extern "C" int vb_fr_chain_backward(int dim, int scale_mode, int objective, double alpha, double n_total,
                                    const double* d_L, const double* d_S, const double* d_GS, const double* d_gmu,
                                    const double* d_rows, const double* d_scal, double* d_out, void* d_ws, size_t ws_bytes,
                                    void* stream) {
    if (!d_L || !d_S || !d_GS || !d_gmu || !d_rows || !d_scal || !d_out || !d_ws || dim < 1) return VB_ERR_ARG;
    if (scale_mode != 0 && scale_mode != 1) return VB_ERR_ARG;
    if (objective < 0 || objective > 2 || !(n_total >= 0.0)) return VB_ERR_ARG;
    if (ws_bytes < vb_fr_chain_workspace_bytes(dim)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    BwdArgs a;
    a.d = dim; a.mode = scale_mode; a.objective = objective; a.alpha = alpha; a.n_total = n_total;
    a.L = d_L; a.S = d_S; a.GS = d_GS; a.gmu = d_gmu; a.rows = d_rows; a.scal = d_scal; a.out = d_out;
    a.ws = chain_ws(d_ws, dim);
    VB_CUDA_TRY(cudaMemsetAsync(a.ws.barrier, 0, 256, st));
    void* args[] = {(void*)&a};
    VB_CUDA_TRY(chain_configure());
    VB_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)fr_chain_backward_kernel, dim3(chain_grid(dim)), dim3(kCT), args,
                                            sizeof(GemmSmem), st));
    return VB_OK;
}

// iteration counts and status of the last forward / backward on this workspace:
// [polar, inverse, smith, status, GEMMs of the forward, GEMMs of the backward, symmetric (half-work) ones among them: forward, backward]
extern "C" int vb_fr_chain_info(const void* d_ws, int dim, int* h_out8, void* stream) {
    if (!d_ws || !h_out8 || dim < 1) return VB_ERR_ARG;
    const ChainWs w = chain_ws(const_cast<void*>(d_ws), dim);
    VB_CUDA_TRY(cudaMemcpyAsync(h_out8, w.info, 8 * sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    VB_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    return VB_OK;
}
