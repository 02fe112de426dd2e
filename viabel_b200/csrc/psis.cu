// Pareto smoothed importance sampling on the device (notebooks/psis.py:112-395).
//
// The reference does a full argsort of n log weights only to read ONE order statistic and
// then works on the <= ceil(min(0.2 n, 3 sqrt(n))) largest values.  Here:
//   P1  max                                              (8 B/weight read)
//   P2  prefilter: u = max - x; elements with u <= pivot (a ~1% sample quantile) are
//       appended as (u, index) candidates, all others contribute exp(-u) to the body sum
//                                                        (8 B/weight read)
//       optional refinement of the candidates with a 64 Ki-bin histogram of u
//   S   bitonic sort of the (few 1e4..1e5) candidates by (u descending, index ascending)
//       == ascending shifted log weight with ties broken by ascending index
//   T   exact cutoff = (K+1)-th largest, strict tail, GPD fit, inverse-CDF smoothing
//   P3  out = (x - max) - logsumexp, tail scattered      (16 B/weight read+write)
// Selection, tail membership and tail order are exact (bit-level comparisons on the same
// fp64 values the reference compares); k-hat and the smoothed values differ from numpy only
// by summation order.
#include <cstdlib>
#include "common.cuh"
#include <cfloat>

namespace vb {

static inline int grid_for(long long work_items, int threads, int max_blocks) {
    long long g = (work_items + threads - 1) / threads;
    if (g > max_blocks) g = max_blocks;
    if (g < 1) g = 1;
    return (int)g;
}

struct __align__(16) Cand { unsigned long long key; unsigned long long idx; };
// key = ~bits(u), u >= 0: ascending key == descending u == ascending shifted log weight.
__device__ __forceinline__ unsigned long long u_to_key(double u) {
    return ~(unsigned long long)__double_as_longlong(u);
}
__device__ __forceinline__ double key_to_u(unsigned long long k) {
    return __longlong_as_double((long long)~k);
}
__device__ __forceinline__ bool cand_less(const Cand& a, const Cand& b) {
    return a.key < b.key || (a.key == b.key && a.idx < b.idx);
}

constexpr int kSortChunk = 4096;          // elements sorted per CTA in shared memory (64 KiB)
constexpr int kSortThreads = 1024;
constexpr int kHistBins = 8192;          // 32 KiB of counters: privatised per CTA in shared memory
constexpr int kSampleSize = 4096;
constexpr long long kDirectSortMax = 1ll << 17;
constexpr double kNegLogTiny = 708.3964185322641;      // -log(DBL_MIN), psis.py:158

// ------------------------------------------------------------------------------------------ bitonic sort
__device__ __forceinline__ void cmp_swap(Cand& a, Cand& b, bool ascending) {
    if (cand_less(b, a) == ascending) { const Cand t = a; a = b; b = t; }
}

// Sort each chunk of kSortChunk elements completely (stages k = 2 .. kSortChunk).
__global__ void __launch_bounds__(kSortThreads)
bitonic_chunk_sort_kernel(Cand* __restrict__ data) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Cand* s = reinterpret_cast<Cand*>(smem_raw);
    const long long base = (long long)blockIdx.x * kSortChunk;
    for (int i = threadIdx.x; i < kSortChunk; i += kSortThreads) s[i] = data[base + i];
    __syncthreads();
    for (int k = 2; k <= kSortChunk; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < kSortChunk / 2; t += kSortThreads) {
                const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const int p = i | j;
                const bool asc = (((base + i) & k) == 0);
                cmp_swap(s[i], s[p], asc);
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < kSortChunk; i += kSortThreads) data[base + i] = s[i];
}

// One global compare-exchange step (j >= kSortChunk) of stage k.
__global__ void bitonic_global_step_kernel(Cand* __restrict__ data, long long n_pad, long long k, long long j) {
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n_pad / 2;
         t += (long long)gridDim.x * blockDim.x) {
        const long long i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const long long p = i | j;
        const bool asc = ((i & k) == 0);
        Cand a = data[i], b = data[p];
        if (cand_less(b, a) == asc) { data[i] = b; data[p] = a; }
    }
}

// Remaining steps j = kSortChunk/2 .. 1 of stage k, inside shared memory.
__global__ void __launch_bounds__(kSortThreads)
bitonic_chunk_merge_kernel(Cand* __restrict__ data, long long k) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Cand* s = reinterpret_cast<Cand*>(smem_raw);
    const long long base = (long long)blockIdx.x * kSortChunk;
    for (int i = threadIdx.x; i < kSortChunk; i += kSortThreads) s[i] = data[base + i];
    __syncthreads();
    const bool asc = ((base & k) == 0);     // constant inside a chunk because k > kSortChunk
    for (int j = kSortChunk >> 1; j > 0; j >>= 1) {
        for (int t = threadIdx.x; t < kSortChunk / 2; t += kSortThreads) {
            const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
            cmp_swap(s[i], s[i | j], asc);
        }
        __syncthreads();
    }
    for (int i = threadIdx.x; i < kSortChunk; i += kSortThreads) data[base + i] = s[i];
}

__global__ void pad_cands_kernel(Cand* __restrict__ data, long long count, long long n_pad) {
    for (long long i = count + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_pad;
         i += (long long)gridDim.x * blockDim.x) {
        data[i].key = ~0ull;
        data[i].idx = ~0ull;
    }
}

static long long pad_pow2(long long count) {
    long long p = kSortChunk;
    while (p < count) p <<= 1;
    return p;
}

static DeviceOnce g_sort_attr_set;
// ascending sort of data[0..n_pad) (n_pad a power of two >= kSortChunk; caller padded the end)
static int bitonic_sort(Cand* data, long long n_pad, cudaStream_t st) {
    const size_t smem = (size_t)kSortChunk * sizeof(Cand);
    const int dev = current_device();
    if (g_sort_attr_set.pending(dev)) {
        VB_CUDA_TRY(cudaFuncSetAttribute(bitonic_chunk_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        VB_CUDA_TRY(cudaFuncSetAttribute(bitonic_chunk_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        g_sort_attr_set.done(dev);
    }
    const int chunks = (int)(n_pad / kSortChunk);
    bitonic_chunk_sort_kernel<<<chunks, kSortThreads, smem, st>>>(data);
    VB_LAUNCH_CHECK();
    for (long long k = 2ll * kSortChunk; k <= n_pad; k <<= 1) {
        for (long long j = k >> 1; j >= kSortChunk; j >>= 1) {
            bitonic_global_step_kernel<<<grid_for(n_pad / 2, 256, kNumSMs * 8), 256, 0, st>>>(data, n_pad, k, j);
            VB_LAUNCH_CHECK();
        }
        bitonic_chunk_merge_kernel<<<chunks, kSortThreads, smem, st>>>(data, k);
        VB_LAUNCH_CHECK();
    }
    return VB_OK;
}

// ------------------------------------------------------------------------------------------ pivot from a strided sample
// The pivot lives in x-space (raw log weights), so neither the sample nor the filter pass needs the
// maximum: sample[i] = (ordered key of x[floor(i n / S)], index); sorted ascending.
__global__ void sample_kernel(const double* __restrict__ lw, long long n, long long index0, Cand* __restrict__ sample) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < kSampleSize; i += gridDim.x * blockDim.x) {
        const long long e = ((long long)i * n) / kSampleSize;   // i < 2^12, n < 2^50
        Cand c;
        c.key = f64_to_key(lw[e]);
        c.idx = (unsigned long long)(index0 + e);
        sample[i] = c;
    }
}
// pivot_x = the (rank+1)-th largest sampled value, or -inf when rank >= count (everything is a candidate)
__global__ void pivot_kernel(const Cand* __restrict__ sorted_sample, long long count, long long rank,
                             double* __restrict__ out_pivot_x) {
    if (threadIdx.x == 0 && blockIdx.x == 0)
        *out_pivot_x = (rank >= count) ? -INFINITY : key_to_f64(sorted_sample[count - 1 - rank].key);
}
// The same selection without sorting: one CTA radix-selects the (rank+1)-th largest 64-bit key,
// eight 8-bit digits from the top (ties in the VALUE do not matter for a pivot).
__global__ void __launch_bounds__(1024)
radix_select_kernel(const Cand* __restrict__ sample, int count, int rank, double* __restrict__ out_pivot_x) {
    __shared__ unsigned int hist[256];
    __shared__ unsigned int wtot[8];
    __shared__ unsigned long long s_prefix;
    __shared__ int s_rank;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { s_prefix = 0ull; s_rank = rank; }
    unsigned long long mask = 0ull;
    for (int shift = 56; shift >= 0; shift -= 8) {
        if (tid < 256) hist[tid] = 0;
        __syncthreads();
        const unsigned long long prefix = s_prefix;
        const int r = s_rank;
        for (int i = tid; i < count; i += blockDim.x) {
            const unsigned long long k = sample[i].key;
            if ((k & mask) == prefix) atomicAdd(&hist[(unsigned)(k >> shift) & 255u], 1u);
        }
        __syncthreads();
        // above[d] = number of keys with a larger digit: suffix sums by warp shuffles (a serial walk over up to 255
        // counters per thread made this kernel 18 us)
        unsigned int h = 0, v = 0;
        if (tid < 256) {
            h = hist[tid];
            v = h;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned int t = __shfl_down_sync(0xffffffffu, v, o);
                if (lane + o < 32) v += t;
            }
            if (lane == 0) wtot[warp] = v;
        }
        __syncthreads();
        if (tid < 256) {
            unsigned int above = v - h;
            for (int w = warp + 1; w < 8; ++w) above += wtot[w];
            if ((int)above <= r && r < (int)(above + h)) {               // exactly one digit owns rank r
                s_prefix = prefix | ((unsigned long long)tid << shift);
                s_rank = r - (int)above;
            }
        }
        mask |= 255ull << shift;
        __syncthreads();
    }
    if (tid == 0) *out_pivot_x = key_to_f64(s_prefix);
}
__global__ void set_double_kernel(double* p, double v) { if (threadIdx.x == 0 && blockIdx.x == 0) *p = v; }
__global__ void set_counts_kernel(long long* p, long long a, long long b, long long c, long long d) {
    if (threadIdx.x == 0 && blockIdx.x == 0) { p[0] = a; p[1] = b; p[2] = c; p[3] = d; }
}

// ------------------------------------------------------------------------------------------ P1: fused filter pass
// ONE read of the weights: elements with x >= pivot_x are appended as (raw bits of x, global index)
// candidates, every other element adds exp(x - pivot_x) to the body sum (terms <= 1: no running
// max needed), and the maximum is taken over the candidates only (it is one of them).
// Candidates are staged per WARP in shared memory (no block barrier in the loop) and flushed with
// one global atomic per flush: a per-warp atomic per round on a single counter serialises in L2
// (measured 1.16 ms at n = 1.25e8).
struct FilterOut {
    Cand* cand;
    long long cap;
    unsigned long long* counter;     // number of candidates found (may exceed cap)
    double* block_body;              // [grid]
    double* block_max;               // [grid]
};

constexpr int kFThreads = 256;
constexpr int kFWarps = kFThreads / 32;
constexpr int kFPerThread = 8;                          // 4 x 16-byte loads in flight
constexpr int kFChunk = kFThreads * kFPerThread;        // elements per CTA per round
constexpr int kFWarpBuf = 128;                          // staged candidates per warp (2 KiB); overflow goes straight to global
constexpr size_t kFSmem = (size_t)kFWarps * kFWarpBuf * sizeof(Cand);

__device__ __forceinline__ void filter_flush(Cand* mybuf, unsigned int* wcount, unsigned int cnt, int lane,
                                             const FilterOut& o) {
    unsigned long long gbase = 0;
    if (lane == 0) gbase = atomicAdd(o.counter, (unsigned long long)cnt);
    gbase = __shfl_sync(0xffffffffu, gbase, 0);
    for (unsigned int k = lane; k < cnt; k += 32)
        if ((long long)(gbase + k) < o.cap) o.cand[gbase + k] = mybuf[k];
    __syncwarp();
}

template <bool ALIGNED, int MIXED, int MINB>
__global__ void __launch_bounds__(kFThreads, MINB)
filter_kernel(const double* __restrict__ lw, long long n, long long index0, const double* __restrict__ pivot_ptr,
              const FilterOut o) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ unsigned int wcounts[kFWarps];
    __shared__ double scratch[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    Cand* mybuf = reinterpret_cast<Cand*>(smem_raw) + warp * kFWarpBuf;
    unsigned int* wcount = &wcounts[warp];
    if (lane == 0) *wcount = 0;
    __syncwarp();
    __shared__ double s_table[64];
    if (tid < 64) s_table[tid] = kExp2Table[tid];
    __syncthreads();
    const double pivot = *pivot_ptr;
    double body = 0.0, cmax = -INFINITY;
    const long long n_chunks = (n + kFChunk - 1) / kFChunk;
    const long long full_chunks = n / kFChunk;

    // Candidates are compacted per warp without atomics: one ballot per element slot, ranks from the lane mask, the
    // fill count in a (warp-uniform) register.  With ~1 % candidates nearly every 256-element warp round holds one, so
    // the per-lane `if (candidate)` path with a shared-memory atomic that this replaces ran (divergent) in most rounds.
    unsigned int cnt = 0;                                // staged candidates of this warp (uniform across lanes)
    const unsigned int lt_mask = (1u << lane) - 1u;
    auto stage = [&](bool c, double x, long long i) {    // executed by the whole warp (converged)
        const unsigned int mask = __ballot_sync(0xffffffffu, c);
        if (mask) {
            const unsigned int add = __popc(mask);
            if (cnt + add > (unsigned)kFWarpBuf) {       // no room: flush what is staged first
                __syncwarp();
                filter_flush(mybuf, wcount, cnt, lane, o);
                cnt = 0;
            }
            if (c) {
                Cand cd;
                cd.key = (unsigned long long)__double_as_longlong(x);
                cd.idx = (unsigned long long)(index0 + i);
                mybuf[cnt + __popc(mask & lt_mask)] = cd;
                cmax = fmax(cmax, x);
            }
            cnt += add;
        }
    };
    double body2 = 0.0;                                  // second accumulator: two independent add chains

    // software pipeline: the next chunk's four 16-byte loads are in flight while this one is evaluated
    double2 v[kFPerThread / 2];
    {
        const long long c0 = blockIdx.x;
        if (ALIGNED && c0 < full_chunks) {
            const double2* p = reinterpret_cast<const double2*>(lw + c0 * kFChunk) + tid;
#pragma unroll
            for (int q = 0; q < kFPerThread / 2; ++q) v[q] = ld_stream2(p + q * kFThreads);
        }
    }
    for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const long long base = chunk * kFChunk;
        if (ALIGNED && chunk < full_chunks) {
            double2 nv[kFPerThread / 2];
            const long long next = chunk + gridDim.x;
            if (next < full_chunks) {
                const double2* p = reinterpret_cast<const double2*>(lw + next * kFChunk) + tid;
#pragma unroll
                for (int q = 0; q < kFPerThread / 2; ++q) nv[q] = ld_stream2(p + q * kFThreads);
            }
#pragma unroll
            for (int q = 0; q < kFPerThread / 2; ++q) {
                // every element is exponentiated (the evaluation is exact for the small positive arguments of the
                // candidates too), the add is predicated: no select in front of the exp.  NaN: not a candidate (the
                // compare is false), poisons the sum
                const bool c0 = v[q].x >= pivot, c1 = v[q].y >= pivot;
                const double e0 = exp_nonpos_flush_v<MIXED>(v[q].x - pivot, s_table);
                const double e1 = exp_nonpos_flush_v<MIXED>(v[q].y - pivot, s_table);
                if (!c0) body += e0;
                if (!c1) body2 += e1;
                if (__any_sync(0xffffffffu, c0 || c1)) {   // 0.25 % candidates: one vote per pair of slots skips most stagings
                    const long long i = base + 2 * (tid + q * kFThreads);
                    stage(c0, v[q].x, i);
                    stage(c1, v[q].y, i + 1);
                }
            }
#pragma unroll
            for (int q = 0; q < kFPerThread / 2; ++q) v[q] = nv[q];
        } else {
#pragma unroll 1
            for (int q = 0; q < kFPerThread; ++q) {
                const long long i = base + tid + q * kFThreads;
                const bool in = i < n;
                const double x = in ? lw[i] : 0.0;
                const bool c = in && x >= pivot;
                if (in && !c) body += exp_nonpos_flush(x - pivot, s_table);
                stage(c, x, i);
            }
        }
        if (cnt > (unsigned)(kFWarpBuf / 2)) {
            __syncwarp();
            filter_flush(mybuf, wcount, cnt, lane, o);
            cnt = 0;
        }
    }
    __syncwarp();
    if (cnt > 0) filter_flush(mybuf, wcount, cnt, lane, o);
    body = block_sum(body + body2, scratch);
    cmax = block_max(cmax, scratch);
    if (tid == 0) { o.block_body[blockIdx.x] = body; o.block_max[blockIdx.x] = cmax; }
}
// out_body = sum of the block bodies (fixed order), out_max = max of the block maxima
__global__ void __launch_bounds__(256)
filter_finish_kernel(const double* __restrict__ block_body, const double* __restrict__ block_max_in, int G,
                     double* __restrict__ out_body, double* __restrict__ out_max) {
    __shared__ double scratch[32];
    double s = 0.0, m = -INFINITY;
    for (int g = threadIdx.x; g < G; g += blockDim.x) { s += block_body[g]; m = fmax(m, block_max_in[g]); }
    s = block_sum(s, scratch);
    m = block_max(m, scratch);
    if (threadIdx.x == 0) { *out_body = s; *out_max = m; }
}
// Device-driven counts: a kernel given `cnt_ptr` takes the element count from device memory (clamped to the buffer
// capacity passed as `count`), so that no host read sits between the stages; null = the host value.
__device__ __forceinline__ long long dev_count(long long count, const unsigned long long* cnt_ptr) {
    if (!cnt_ptr) return count;
    const unsigned long long c = *cnt_ptr;
    return c < (unsigned long long)count ? (long long)c : count;
}

// Once the (global) maximum M is known: candidate keys become ~bits(M - x) (the reference sorts the
// SHIFTED weights, psis.py:165-167), pivot_u = M - pivot_x, and the body sum is rescaled to M.
__global__ void rekey_kernel(Cand* __restrict__ cand, long long count, const double* __restrict__ M_ptr,
                             const unsigned long long* __restrict__ cnt_ptr = nullptr) {
    count = dev_count(count, cnt_ptr);
    const double M = *M_ptr;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count;
         i += (long long)gridDim.x * blockDim.x)
        cand[i].key = u_to_key(M - __longlong_as_double((long long)cand[i].key));
}
// terms[0] = body_pre * exp(pivot_x - M); pivot_u = M - pivot_x
__global__ void derive_kernel(const double* __restrict__ M_ptr, const double* __restrict__ pivotx_ptr,
                              const double* __restrict__ body_pre, double* __restrict__ pivot_u,
                              double* __restrict__ term0) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const double M = *M_ptr, px = *pivotx_ptr, b = *body_pre;
    if (pivot_u) *pivot_u = M - px;                       // +inf when everything was a candidate
    *term0 = (b == 0.0) ? 0.0 : b * exp(px - M);
}

// ------------------------------------------------------------------------------------------ refinement histogram
__global__ void __launch_bounds__(512)
hist_kernel(const Cand* __restrict__ cand, long long count, const double* __restrict__ pivot_ptr,
            unsigned int* __restrict__ hist, const unsigned long long* __restrict__ cnt_ptr = nullptr) {
    count = dev_count(count, cnt_ptr);
    __shared__ unsigned int sh[kHistBins];
    for (int b = threadIdx.x; b < kHistBins; b += blockDim.x) sh[b] = 0;
    __syncthreads();
    const double pivot = *pivot_ptr;
    const double scale = (pivot > 0.0 && isfinite(pivot)) ? (double)kHistBins / pivot : 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count;
         i += (long long)gridDim.x * blockDim.x) {
        const double u = key_to_u(cand[i].key);
        long long b = (long long)(u * scale);
        if (b > kHistBins - 1) b = kHistBins - 1;
        atomicAdd(&sh[b], 1u);
    }
    __syncthreads();
    for (int b = threadIdx.x; b < kHistBins; b += blockDim.x)
        if (sh[b]) atomicAdd(&hist[b], sh[b]);
}
// smallest bin b* with cumulative count >= target; out[0] = b*, out[1] = cumulative count through b*
__global__ void __launch_bounds__(1024)
hist_scan_kernel(const unsigned int* __restrict__ hist, long long target, long long* __restrict__ out) {
    __shared__ long long warp_tot[32];
    constexpr int per = kHistBins / 1024;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    long long mine = 0;
    for (int b = 0; b < per; ++b) mine += hist[tid * per + b];
    long long incl = mine;                                   // inclusive scan inside the warp
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const long long v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        long long w = warp_tot[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const long long v = __shfl_up_sync(0xffffffffu, w, d);
            if (lane >= d) w += v;
        }
        warp_tot[lane] = w;                                  // inclusive totals of warps 0..lane
    }
    __syncthreads();
    const long long before_warp = (warp == 0) ? 0 : warp_tot[warp - 1];
    incl += before_warp;
    const long long excl = incl - mine;
    if (excl < target && target <= incl) {                   // exactly one thread owns the crossing
        long long cum = excl;
        int b = tid * per;
        for (;; ++b) {
            cum += hist[b];
            if (cum >= target) break;
        }
        out[0] = b;
        out[1] = cum;
    }
    if (tid == 1023 && incl < target) { out[0] = kHistBins - 1; out[1] = incl; }
}
__global__ void __launch_bounds__(256)
compact2_kernel(const Cand* __restrict__ cand, long long count, const double* __restrict__ pivot_ptr,
                const long long* __restrict__ bstar_ptr, Cand* __restrict__ out, unsigned long long* __restrict__ counter,
                double* __restrict__ block_body, const unsigned long long* __restrict__ cnt_ptr = nullptr,
                long long out_cap = 0x7fffffffffffffffll) {
    __shared__ double scratch[32];
    count = dev_count(count, cnt_ptr);
    const double pivot = *pivot_ptr;
    const double scale = (pivot > 0.0 && isfinite(pivot)) ? (double)kHistBins / pivot : 0.0;
    const long long bstar = *bstar_ptr;
    double body = 0.0;
    const long long gtid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long gthreads = (long long)gridDim.x * blockDim.x;
    const long long iters = (count + gthreads - 1) / gthreads;
    for (long long it = 0; it < iters; ++it) {
        const long long i = it * gthreads + gtid;
        bool keep = false;
        Cand c;
        c.key = 0; c.idx = 0;
        if (i < count) {
            c = cand[i];
            const double u = key_to_u(c.key);
            long long b = (long long)(u * scale);
            if (b > kHistBins - 1) b = kHistBins - 1;
            keep = (b <= bstar);
            if (!keep) body += exp_nonpos(-u);
        }
        const unsigned mask = __ballot_sync(0xffffffffu, keep);
        if (mask) {
            const int lane = threadIdx.x & 31;
            const int leader = __ffs(mask) - 1;
            unsigned long long basepos = 0;
            if (lane == leader) basepos = atomicAdd(counter, (unsigned long long)__popc(mask));
            basepos = __shfl_sync(0xffffffffu, basepos, leader);
            const unsigned long long pos = basepos + __popc(mask & ((1u << lane) - 1));
            if (keep && (long long)pos < out_cap) out[pos] = c;      // the counter still counts what did not fit
        }
    }
    body = block_sum(body, scratch);
    if (threadIdx.x == 0) block_body[blockIdx.x] = body;
}

// ------------------------------------------------------------------------------------------ counting sort by bin
// The refinement bins are monotone in u, so the kept candidates can be ordered bin by bin: an exclusive scan of
// the counters gives every bin its segment (largest u first == ascending key), a scatter fills the segments, and
// one warp per bin ranks its few elements by (key, index).  Replaces the 17-stage bitonic network over the padded
// candidate array (150 us at 65 536 entries, barrier-latency bound) when no bin is large; heavy ties or a
// degenerate pivot (everything in one bin) fall back to the bitonic sort.
constexpr int kBinSortMax = 512;               // largest bin the per-warp ranking takes
// offsets[b] = number of kept candidates in bins < b (b = 0 .. kHistBins); meta = [largest bin, total]
__global__ void __launch_bounds__(1024)
bin_scan_kernel(const unsigned int* __restrict__ hist, const long long* __restrict__ bstar_ptr,
                unsigned int* __restrict__ offsets, long long* __restrict__ meta) {
    __shared__ unsigned int warp_tot[32];
    __shared__ unsigned int warp_max[32];
    constexpr int per = kHistBins / 1024;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long bstar = bstar_ptr ? *bstar_ptr : (long long)kHistBins - 1;
    unsigned int c[per];
    unsigned int mine = 0, mx = 0;
#pragma unroll
    for (int b = 0; b < per; ++b) {
        const int bin = tid * per + b;
        c[b] = (bin <= bstar) ? hist[bin] : 0u;
        mine += c[b];
        mx = max(mx, c[b]);
    }
    unsigned int incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned int v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, d));
    if (lane == 31) warp_tot[warp] = incl;
    if (lane == 0) warp_max[warp] = mx;
    __syncthreads();
    if (warp == 0) {
        unsigned int w = warp_tot[lane], m = warp_max[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned int v = __shfl_up_sync(0xffffffffu, w, d);
            if (lane >= d) w += v;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, d));
        warp_tot[lane] = w;
        if (lane == 0) { meta[0] = (long long)m; }
        if (lane == 31) { meta[1] = (long long)w; offsets[kHistBins] = w; }
    }
    __syncthreads();
    unsigned int run = incl - mine + ((warp == 0) ? 0u : warp_tot[warp - 1]);
#pragma unroll
    for (int b = 0; b < per; ++b) {
        offsets[tid * per + b] = run;
        run += c[b];
    }
}
// Preconditions of the device-driven tail (the host-driven path decides these from three host reads):
//   K1 + 1 <= C1 <= cap1 (the filter pass found enough candidates and they fit), K1 + 1 <= C2 (kept by the
//   refinement), every kept bin small enough for the per-warp ranking, histogram total == kept count.
__global__ void psis_check_kernel(const unsigned long long* __restrict__ c1_ptr, long long cap1,
                                  const unsigned long long* __restrict__ c2_ptr, const long long* __restrict__ meta,
                                  long long K1, int max_bin, int* __restrict__ status,
                                  const int* __restrict__ status_in = nullptr) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const unsigned long long c1 = *c1_ptr, c2 = *c2_ptr;
    int st = 0;
    if (status_in && *status_in) st = *status_in;              // raised by an earlier stage (any rank)
    else if (c1 < (unsigned long long)(K1 + 1)) st = 1;
    else if (c1 > (unsigned long long)cap1) st = 2;
    else if (c2 < (unsigned long long)(K1 + 1) || c2 > c1) st = 3;
    else if (meta[0] > max_bin) st = 4;
    else if ((unsigned long long)meta[1] != c2) st = 5;
    *status = st;
}
__device__ __forceinline__ int cand_bin(const Cand& c, double scale) {
    long long b = (long long)(key_to_u(c.key) * scale);         // same expression as hist_kernel
    return (int)(b > kHistBins - 1 ? kHistBins - 1 : b);
}
// dst[total - offsets[b + 1] + slot]: bins in descending order, arbitrary order inside a bin
__global__ void __launch_bounds__(256)
bin_scatter_kernel(const Cand* __restrict__ src, const long long* __restrict__ meta, const double* __restrict__ pivot_ptr,
                   const unsigned int* __restrict__ offsets, unsigned int* __restrict__ cursor, Cand* __restrict__ dst,
                   const int* __restrict__ status = nullptr) {
    if (status && *status) return;
    const double pivot = *pivot_ptr;
    const double scale = (pivot > 0.0 && isfinite(pivot)) ? (double)kHistBins / pivot : 0.0;
    const long long total = meta[1];
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const Cand c = src[i];
        const int b = cand_bin(c, scale);
        const unsigned int slot = atomicAdd(&cursor[b], 1u);
        dst[total - (long long)offsets[b + 1] + slot] = c;
    }
}
// one warp per bin: rank of every element among the bin's elements by (key, index)
__global__ void __launch_bounds__(256)
bin_rank_kernel(const Cand* __restrict__ src, const long long* __restrict__ meta, const long long* __restrict__ bstar_ptr,
                const unsigned int* __restrict__ offsets, Cand* __restrict__ dst, const int* __restrict__ status = nullptr) {
    if (status && *status) return;
    const long long total = meta[1];
    const int n_bins = (int)((bstar_ptr ? *bstar_ptr : (long long)kHistBins - 1) + 1);
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < n_bins; b += warps) {
        const unsigned int lo = offsets[b], hi = offsets[b + 1];
        const int c = (int)(hi - lo);
        if (c == 0) continue;
        const long long base = total - (long long)hi;
        const Cand* seg = src + base;
        for (int i = lane; i < c; i += 32) {
            const Cand e = seg[i];
            int rank = 0;
            for (int j = 0; j < c; ++j) rank += cand_less(seg[j], e) ? 1 : 0;
            dst[base + rank] = e;
        }
    }
}

// ------------------------------------------------------------------------------------------ tail bookkeeping
// state[] (doubles): [0] xcutoff  [1] exp(xcutoff)  [2] k  [3] sigma  [4] smoothed  [5] lse  [6] b (GPD)  [7] spare
// counts[] (long long): [0] n2  [1] first tail position p in the sorted candidates  [2] m (quadrature points)
//                       [3] 1 when the GPD fit runs (psislw: n2 > 4, psis.py:177)
__global__ void tail_bounds_kernel(const Cand* __restrict__ sorted, long long count, long long K1,
                                   double* __restrict__ state, long long* __restrict__ counts,
                                   const unsigned long long* __restrict__ cnt_ptr = nullptr,
                                   const int* __restrict__ status = nullptr) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (status && *status) {                       // device-driven path found its preconditions violated: inert state
        state[0] = 0.0; state[1] = 1.0; counts[0] = 0; counts[1] = 0; counts[2] = 30; counts[3] = 0;
        return;
    }
    count = dev_count(count, cnt_ptr);
    // (K1+1)-th largest shifted weight == (K1+1)-th smallest u == sorted[count-1-K1]   (psis.py:157,169)
    const double u_cut = key_to_u(sorted[count - 1 - K1].key);
    const double xcut = fmax(-u_cut, -kNegLogTiny);                  // psis.py:169-172
    const double u_eff = -xcut;
    // sorted is descending in u: tail = suffix with u < u_eff  (x > xcutoff, strict; psis.py:174)
    long long lo = 0, hi = count;                                    // first position with u < u_eff
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        if (key_to_u(sorted[mid].key) < u_eff) hi = mid; else lo = mid + 1;
    }
    const long long n2 = count - lo;
    state[0] = xcut;
    state[1] = exp(xcut);
    counts[0] = n2;
    counts[1] = lo;
    counts[2] = 30 + (long long)sqrt((double)n2);                    // psis.py:268
    counts[3] = (n2 > 4) ? 1 : 0;                                    // psis.py:177
}

// x2[r] = exp(x_r) - exp(xcutoff) for the tail in ascending order (psis.py:184-185)
__global__ void tail_excess_kernel(const Cand* __restrict__ sorted, const double* __restrict__ state,
                                   const long long* __restrict__ counts, double* __restrict__ x2) {
    const long long n2 = counts[0], p = counts[1];
    const double ec = state[1];
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < n2;
         r += (long long)gridDim.x * blockDim.x)
        x2[r] = exp(-key_to_u(sorted[p + r].key)) - ec;
}

// ------------------------------------------------------------------------------------------ GPD fit (psis.py:211-331)
__device__ __forceinline__ double gpd_b(long long j /*0-based*/, long long m, double x_q, double x_max) {
    const double jj = (double)(j + 1) - 0.5;
    double b = 1.0 - sqrt((double)m / jj);          // psis.py:270-274
    b /= 3.0 * x_q;                                 // PRIOR = 3, psis.py:276/279
    b += 1.0 / x_max;                               // psis.py:277/280
    return b;
}
// kGpdSplit CTAs per quadrature point: parts[j][split] = partial sum_r log1p(-b_j x_r)
// (a single CTA per point leaves the GPU at ~11 warps/SM: latency-bound, 75 us at n2 = 33 541)
constexpr int kGpdSplit = 8;
__global__ void __launch_bounds__(256)
gpd_profile_kernel(const double* __restrict__ x, const long long* __restrict__ counts, double* __restrict__ parts) {
    __shared__ double scratch[32];
    const long long n = counts[0], m = counts[2];
    const long long j = blockIdx.x;
    if (counts[3] == 0 || j >= m) return;
    const double x_q = x[(long long)((double)n / 4.0 + 0.5) - 1];
    const double x_max = x[n - 1];
    const double b = gpd_b(j, m, x_q, x_max);
    double s = 0.0;
    // log(1 - b x) through the table-driven log (1 ulp of max(|log|, 1e-16)): the argument is rounded to 1 - b x, an
    // absolute 1e-16 per term of a sum of n2 O(1) terms -- 1e-14 in L_j = n2 (...), far inside the 1e-9 of the tests;
    // half the instructions of the library log1p (this kernel evaluates n2 x m = 7e6 logarithms at C5)
    for (long long r = (long long)blockIdx.y * blockDim.x + threadIdx.x; r < n; r += (long long)kGpdSplit * blockDim.x)
        s += fast_log(fma(-b, x[r], 1.0));
    s = block_sum(s, scratch);
    if (threadIdx.x == 0) parts[j * kGpdSplit + blockIdx.y] = s;
}
// profile log-likelihood L_j = n (log(-(b_j/ks_j)) - ks_j - 1) (psis.py:287-289)
__global__ void __launch_bounds__(256)
gpd_loglik_kernel(const double* __restrict__ x, const long long* __restrict__ counts, const double* __restrict__ parts,
                  double* __restrict__ bs, double* __restrict__ L) {
    const long long n = counts[0], m = counts[2];
    if (counts[3] == 0) return;
    const long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (j >= m) return;
    const double x_q = x[(long long)((double)n / 4.0 + 0.5) - 1];
    const double x_max = x[n - 1];
    double s = 0.0;
    for (int k = 0; k < kGpdSplit; ++k) s += parts[j * kGpdSplit + k];          // fixed order
    const double ks = s / (double)n;
    const double b = gpd_b(j, m, x_q, x_max);
    bs[j] = b;
    L[j] = (double)n * (log(-(b / ks)) - ks - 1.0);
}
// quadrature weights w_j = 1 / sum_i exp(L_i - L_j), one WARP per j over a grid of CTAs (as one CTA this was a chain of
// ~50 dependent fp64 exponentials per lane, 26 us at m = 213); the CTA that finishes last (ticket) takes the posterior
// mean of b in fixed order (psis.py:293-308)
constexpr int kGpdWWarps = 8;
__global__ void __launch_bounds__(kGpdWWarps * 32)
gpd_weights_kernel(const long long* __restrict__ counts, const double* __restrict__ bs, const double* __restrict__ L,
                   double* __restrict__ w_out, double* __restrict__ state, unsigned int* __restrict__ ticket) {
    __shared__ double scratch[32];
    __shared__ bool s_last;
    const long long m = counts[2];
    if (counts[3] == 0) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long j = blockIdx.x * (long long)kGpdWWarps + warp;
    if (j < m) {
        double s = 0.0;
        const double Lj = L[j];
        for (long long i = lane; i < m; i += 32) s += exp(L[i] - Lj);
        s = warp_sum(s);
        if (lane == 0) w_out[j] = 1.0 / s;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double sw = 0.0, sbw = 0.0;
    for (long long i = threadIdx.x; i < m; i += blockDim.x) {
        const double w = __ldcg(w_out + i);
        if (w >= 10.0 * DBL_EPSILON) { sw += w; sbw += bs[i] * w; }      // psis.py:300-305
    }
    sw = block_sum(sw, scratch);
    sbw = block_sum(sbw, scratch);
    if (threadIdx.x == 0) state[6] = sbw / sw;                            // b, psis.py:308
}
// partial sums of log1p(-b x_r)
__global__ void __launch_bounds__(256)
gpd_k_blocks_kernel(const double* __restrict__ x, const long long* __restrict__ counts, const double* __restrict__ state,
                    double* __restrict__ block_out) {
    __shared__ double scratch[32];
    const long long n = counts[0];
    const double b = state[6];
    double s = 0.0;
    if (counts[3] != 0)
        for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < n;
             r += (long long)gridDim.x * blockDim.x)
            s += log1p(-b * x[r]);
    s = block_sum(s, scratch);
    if (threadIdx.x == 0) block_out[blockIdx.x] = s;
}
__global__ void __launch_bounds__(256)
gpd_k_finish_kernel(const long long* __restrict__ counts, const double* __restrict__ blocks, int G,
                    double* __restrict__ state) {
    __shared__ double scratch[32];
    double s = 0.0;
    for (int g = threadIdx.x; g < G; g += blockDim.x) s += blocks[g];
    s = block_sum(s, scratch);
    if (threadIdx.x != 0) return;
    const long long n = counts[0];
    if (counts[3] == 0) {                                                 // psis.py:177-179
        state[2] = INFINITY; state[3] = NAN; state[4] = 0.0;
        return;
    }
    double k = s / (double)n;                                             // psis.py:311-313
    const double b = state[6];
    const double sigma = -k / b * (double)n / (double)(n - 0);            // psis.py:320
    const double a = 10.0;
    k = k * (double)n / ((double)n + a) + a * 0.5 / ((double)n + a);      // psis.py:322-323
    state[2] = k;
    state[3] = sigma;
    state[4] = (k >= 1.0 / 3.0 && !isinf(k)) ? 1.0 : 0.0;                 // psis.py:187 (NaN compares false)
}

// gpinv for p strictly inside (0,1) (psis.py:342-351)
__device__ __forceinline__ double gpinv_inside(double p, double k, double sigma) {
    if (!(sigma > 0.0)) return NAN;                                       // psis.py:338-339
    double t = log1p(-p);
    if (fabs(k) < DBL_EPSILON) t = -t;
    else t = expm1(-k * t) / k;
    return t * sigma;
}

// new tail values (psis.py:190-198) or the raw ones, and partial sums of exp(value)
__global__ void __launch_bounds__(256)
tail_smooth_kernel(const Cand* __restrict__ sorted, const double* __restrict__ state,
                   const long long* __restrict__ counts, double* __restrict__ tail_new,
                   long long* __restrict__ tail_idx_or_null, double* __restrict__ block_out) {
    __shared__ double scratch[32];
    const long long n2 = counts[0], p = counts[1];
    const bool smoothed = state[4] != 0.0;
    const double k = state[2], sigma = state[3], ec = state[1];
    double s = 0.0;
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < n2;
         r += (long long)gridDim.x * blockDim.x) {
        const Cand c = sorted[p + r];
        double v;
        if (smoothed) {
            const double pr = ((double)r + 0.5) / (double)n2;
            v = log(gpinv_inside(pr, k, sigma) + ec);
            if (v > 0.0) v = 0.0;                                         // psis.py:198
        } else {
            v = -key_to_u(c.key);
        }
        tail_new[r] = v;
        if (tail_idx_or_null) tail_idx_or_null[r] = (long long)c.idx;
        s += exp(v);
    }
    s = block_sum(s, scratch);
    if (threadIdx.x == 0) block_out[blockIdx.x] = s;
}
// body candidates (sorted positions [0, p)) contribute exp(-u)
__global__ void __launch_bounds__(256)
cand_body_kernel(const Cand* __restrict__ sorted, const long long* __restrict__ counts, long long skip_idx_lo,
                 long long skip_idx_hi, double* __restrict__ block_out) {
    __shared__ double scratch[32];
    const long long p = counts[1];
    double s = 0.0;
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < p;
         r += (long long)gridDim.x * blockDim.x) {
        const Cand c = sorted[r];
        const long long gi = (long long)c.idx;
        if (gi >= skip_idx_lo && gi < skip_idx_hi) s += exp(-key_to_u(c.key));
    }
    s = block_sum(s, scratch);
    if (threadIdx.x == 0) block_out[blockIdx.x] = s;
}

struct SumList { const double* p[4]; int n[4]; };
// lse = log(sum of all partial lists) (fixed order); fills the diag vector
__global__ void __launch_bounds__(256)
psis_finish_kernel(SumList lists, const double* __restrict__ M_ptr, const long long* __restrict__ counts,
                                   long long n_cand, double* __restrict__ state, double* __restrict__ diag,
                                   double* __restrict__ lse_partial_or_null,
                                   const unsigned long long* __restrict__ cnt_ptr = nullptr,
                                   const int* __restrict__ status = nullptr) {
    __shared__ double scratch[32];
    n_cand = dev_count(n_cand, cnt_ptr);
    double total = 0.0;
    for (int l = 0; l < 4; ++l) {                  // fixed tree per list, lists added in order
        double part = 0.0;
        for (int i = threadIdx.x; i < lists.n[l]; i += blockDim.x) part += lists.p[l][i];
        total += block_sum(part, scratch);
    }
    if (threadIdx.x != 0) return;
    if (lse_partial_or_null) { lse_partial_or_null[0] = 0.0; lse_partial_or_null[1] = total; }
    const double lse = log(total);
    state[5] = lse;
    if (diag) {
        diag[0] = state[2]; diag[1] = state[3]; diag[2] = state[0]; diag[3] = (double)counts[0];
        diag[4] = state[4]; diag[5] = *M_ptr; diag[6] = lse; diag[7] = (double)n_cand;
        if (status && *status) diag[7] = -(double)*status;       // device-driven path: the caller must take the host-driven one
    }
}

// P3: out = (x - max) - lse   (psis.py:165, 200)
__global__ void __launch_bounds__(256)
normalise_kernel(const double* __restrict__ lw, double* __restrict__ out, long long n, const double* __restrict__ M_ptr,
                 const double* __restrict__ lse_ptr) {
    const double M = *M_ptr, lse = *lse_ptr;
    const long long gtid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long gthreads = (long long)gridDim.x * blockDim.x;
    if (((reinterpret_cast<uintptr_t>(lw) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0) {
        const long long n2 = n >> 1;
        const double2* p = reinterpret_cast<const double2*>(lw);
        double2* q = reinterpret_cast<double2*>(out);
        long long i = gtid;
        for (; i + 3 * gthreads < n2; i += 4 * gthreads) {
            double2 v0 = ld_stream2(p + i), v1 = ld_stream2(p + i + gthreads);
            double2 v2 = ld_stream2(p + i + 2 * gthreads), v3 = ld_stream2(p + i + 3 * gthreads);
            v0.x = (v0.x - M) - lse; v0.y = (v0.y - M) - lse; v1.x = (v1.x - M) - lse; v1.y = (v1.y - M) - lse;
            v2.x = (v2.x - M) - lse; v2.y = (v2.y - M) - lse; v3.x = (v3.x - M) - lse; v3.y = (v3.y - M) - lse;
            q[i] = v0; q[i + gthreads] = v1; q[i + 2 * gthreads] = v2; q[i + 3 * gthreads] = v3;
        }
        for (; i < n2; i += gthreads) {
            double2 v = ld_stream2(p + i);
            v.x = (v.x - M) - lse; v.y = (v.y - M) - lse;
            q[i] = v;
        }
        if ((n & 1) && gtid == 0) out[n - 1] = (lw[n - 1] - M) - lse;
    } else {
        for (long long i = gtid; i < n; i += gthreads) out[i] = (lw[i] - M) - lse;
    }
}
// P3 + divergence statistics: the normalise pass is HBM-bound with idle issue slots and knows the exact maximum M,
// so the one-pass statistics of divergence_bound(alpha = 2) (bounds.py:169-176) ride along instead of costing a
// third read of the weights: per element e = exp(x - M) <= 1, sums of e^2 (= w), e^4 (= w^2), (x - K), (x - K)^2
// with K = lw[0] as the shift of the variance sums.  Block partials, merged in fixed order by the finish kernel.
constexpr int kNDThreads = 256;
__global__ void __launch_bounds__(kNDThreads, 3)
normalise_div_kernel(const double* __restrict__ lw, double* __restrict__ out, long long n, const double* __restrict__ M_ptr,
                     const double* __restrict__ lse_ptr, double* __restrict__ block_out /* [grid][4] */) {
    __shared__ double s_table[64];
    __shared__ double scratch[32];
    if (threadIdx.x < 64) s_table[threadIdx.x] = kExp2Table[threadIdx.x];
    __syncthreads();
    const double M = *M_ptr, lse = *lse_ptr, K = lw[0];
    const long long gtid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long gthreads = (long long)gridDim.x * blockDim.x;
    double a2 = 0.0, a4 = 0.0, b2 = 0.0, b4 = 0.0, sx = 0.0, sxx = 0.0, tx = 0.0, txx = 0.0;
    auto one = [&](double x, double& s2, double& s4, double& s1, double& s11) -> double {
        const double d = x - M;
        const double e = exp_nonpos_flush(d, s_table);
        const double w = e * e;
        s2 += w;
        s4 = fma(w, w, s4);
        const double c = x - K;
        s1 += c;
        s11 = fma(c, c, s11);
        return d - lse;
    };
    if (((reinterpret_cast<uintptr_t>(lw) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0) {
        const long long n2 = n >> 1;
        const double2* p = reinterpret_cast<const double2*>(lw);
        double2* q = reinterpret_cast<double2*>(out);
        long long i = gtid;
        // software pipeline: the next round's four 16-byte loads are in flight while this round's eight exponentials
        // are evaluated (without it the pass alternates between a memory phase and a compute phase)
        double2 v0, v1, v2, v3;
        bool have = i + 3 * gthreads < n2;
        if (have) {
            v0 = ld_stream2(p + i); v1 = ld_stream2(p + i + gthreads);
            v2 = ld_stream2(p + i + 2 * gthreads); v3 = ld_stream2(p + i + 3 * gthreads);
        }
        while (have) {
            const long long nx = i + 4 * gthreads;
            const bool more = nx + 3 * gthreads < n2;
            double2 w0, w1, w2, w3;
            if (more) {
                w0 = ld_stream2(p + nx); w1 = ld_stream2(p + nx + gthreads);
                w2 = ld_stream2(p + nx + 2 * gthreads); w3 = ld_stream2(p + nx + 3 * gthreads);
            }
            v0.x = one(v0.x, a2, a4, sx, sxx); v0.y = one(v0.y, b2, b4, tx, txx);
            v1.x = one(v1.x, a2, a4, sx, sxx); v1.y = one(v1.y, b2, b4, tx, txx);
            v2.x = one(v2.x, a2, a4, sx, sxx); v2.y = one(v2.y, b2, b4, tx, txx);
            v3.x = one(v3.x, a2, a4, sx, sxx); v3.y = one(v3.y, b2, b4, tx, txx);
            q[i] = v0; q[i + gthreads] = v1; q[i + 2 * gthreads] = v2; q[i + 3 * gthreads] = v3;
            i = nx;
            have = more;
            if (more) { v0 = w0; v1 = w1; v2 = w2; v3 = w3; }
        }
        for (; i < n2; i += gthreads) {
            double2 v = ld_stream2(p + i);
            v.x = one(v.x, a2, a4, sx, sxx); v.y = one(v.y, b2, b4, tx, txx);
            q[i] = v;
        }
        if ((n & 1) && gtid == 0) out[n - 1] = one(lw[n - 1], a2, a4, sx, sxx);
    } else {
        for (long long i = gtid; i < n; i += gthreads) out[i] = one(lw[i], a2, a4, sx, sxx);
    }
    double r[4] = {a2 + b2, a4 + b4, sx + tx, sxx + txx};
#pragma unroll
    for (int k = 0; k < 4; ++k) r[k] = block_sum(r[k], scratch);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < 4; ++k) block_out[(size_t)blockIdx.x * 4 + k] = r[k];
    }
}
// partials[6] in the layout of vb_bounds_divergence_partials: max, sum w, sum w^2, count, mean lw, M2 lw
__global__ void __launch_bounds__(256)
normalise_div_finish_kernel(const double* __restrict__ lw, const double* __restrict__ blocks, int G, long long n,
                            const double* __restrict__ M_ptr, double* __restrict__ out) {
    __shared__ double scratch[32];
    double acc[4] = {0, 0, 0, 0};
    for (int b = threadIdx.x; b < G; b += blockDim.x)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[k] += blocks[(size_t)b * 4 + k];
#pragma unroll
    for (int k = 0; k < 4; ++k) acc[k] = block_sum(acc[k], scratch);
    if (threadIdx.x != 0) return;
    const double cnt = (double)n;
    out[0] = *M_ptr; out[1] = acc[0]; out[2] = acc[1]; out[3] = cnt;
    out[4] = lw[0] + acc[2] / cnt;
    out[5] = acc[3] - acc[2] * acc[2] / cnt;
}
__global__ void tail_scatter_kernel(const Cand* __restrict__ sorted, const long long* __restrict__ counts,
                                    const double* __restrict__ tail_new, const double* __restrict__ state,
                                    double* __restrict__ out, long long index0, long long n_local) {
    if (state[4] == 0.0) return;                 // not smoothed: P3 already wrote x - max - lse
    const long long n2 = counts[0], p = counts[1];
    const double lse = state[5];
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < n2;
         r += (long long)gridDim.x * blockDim.x) {
        const long long li = (long long)sorted[p + r].idx - index0;
        if (li >= 0 && li < n_local) out[li] = tail_new[r] - lse;
    }
}

// ------------------------------------------------------------------------------------------ host orchestration
static long long psis_K1(long long n, double reff) {
    return (long long)ceil(fmin(0.2 * (double)n, 3.0 * sqrt((double)n / reff)));     // psis.py:157
}

// GPD fit on an ascending device array; n2 and m are read from counts[] on the device.
static int run_gpd_fit(const double* x2, const long long* counts, long long n2_upper, double* bs, double* Lq,
                       double* wq, double* blk, int g_small, double* state, cudaStream_t st) {
    const long long m_max = 30 + (long long)sqrt((double)n2_upper) + 1;
    double* parts = Lq + m_max;                     // the L region holds L[m_max] then parts[m_max][kGpdSplit]
    gpd_profile_kernel<<<dim3((unsigned)m_max, kGpdSplit), 256, 0, st>>>(x2, counts, parts);
    VB_LAUNCH_CHECK();
    gpd_loglik_kernel<<<(unsigned)((m_max + 255) / 256), 256, 0, st>>>(x2, counts, parts, bs, Lq);
    VB_LAUNCH_CHECK();
    unsigned int* ticket = (unsigned int*)blk;      // blk is written by gpd_k_blocks_kernel only after this
    VB_CUDA_TRY(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st));
    gpd_weights_kernel<<<(unsigned)((m_max + kGpdWWarps - 1) / kGpdWWarps), kGpdWWarps * 32, 0, st>>>(counts, bs, Lq, wq, state,
                                                                                                    ticket);
    VB_LAUNCH_CHECK();
    gpd_k_blocks_kernel<<<g_small, 256, 0, st>>>(x2, counts, state, blk);
    VB_LAUNCH_CHECK();
    gpd_k_finish_kernel<<<1, 256, 0, st>>>(counts, blk, g_small, state);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

constexpr int kSmallGrid = kNumSMs * 2;
static int filter_grid(long long n) {
    static const int per_sm = (getenv("VB_FILTER_OCC4") != nullptr && atoi(getenv("VB_FILTER_OCC4")) != 0) ? 4 : 3;
    return grid_for((n + kFPerThread - 1) / kFPerThread, kFThreads, kNumSMs * per_sm);
}

__global__ void sum_blocks_kernel(const double* __restrict__ blocks, int G, double* __restrict__ out) {
    __shared__ double scratch[32];
    double s = 0.0;
    for (int g = threadIdx.x; g < G; g += blockDim.x) s += blocks[g];
    s = block_sum(s, scratch);
    if (threadIdx.x == 0) *out = s;
}
__global__ void set_u64_kernel(unsigned long long* p, unsigned long long v) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *p = v;
}
__global__ void pad_low_kernel(Cand* __restrict__ data, long long count, long long cap) {
    for (long long i = count + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < cap;
         i += (long long)gridDim.x * blockDim.x) { data[i].key = 0ull; data[i].idx = 0ull; }
}

}  // namespace vb

using namespace vb;

extern "C" long long vb_psis_tail_capacity(long long n, double reff) {
    if (n < 2 || !(reff > 0.0)) return 0;
    return psis_K1(n, reff);
}

// ---- S1: strided sample of the raw weights (VB_PSIS_SAMPLE entries of 16 bytes)
extern "C" int vb_psis_sample(const double* d_lw, long long n_local, long long index0, void* d_out_sample, void* stream) {
    if (!d_lw || !d_out_sample || n_local < 1) return VB_ERR_ARG;
    sample_kernel<<<16, 256, 0, (cudaStream_t)stream>>>(d_lw, n_local, index0, (Cand*)d_out_sample);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ---- S2: sort the (gathered) sample and pick pivot_x with ceil(frac * count) sampled values above it
extern "C" int vb_psis_pivot(void* d_sample, long long count, long long capacity, double frac,
                             double* d_out_pivot_x, void* stream) {
    if (!d_sample || !d_out_pivot_x || count < 1 || capacity < count || capacity < kSortChunk ||
        (capacity & (capacity - 1)))
        return VB_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    const long long rank = (long long)ceil(frac * (double)count);
    if (rank >= count) {
        set_double_kernel<<<1, 1, 0, st>>>(d_out_pivot_x, -INFINITY);
        VB_LAUNCH_CHECK();
        return VB_OK;
    }
    if (count <= 16 * kSampleSize) {                // 8 counting passes in one CTA beat a full sort here
        radix_select_kernel<<<1, 1024, 0, st>>>((const Cand*)d_sample, (int)count, (int)rank, d_out_pivot_x);
        VB_LAUNCH_CHECK();
        return VB_OK;
    }
    if (capacity > count) {
        pad_low_kernel<<<grid_for(capacity - count, 256, kNumSMs), 256, 0, st>>>((Cand*)d_sample, count, capacity);
        VB_LAUNCH_CHECK();
    }
    int rc = bitonic_sort((Cand*)d_sample, capacity, st);
    if (rc != VB_OK) return rc;
    // pads (key 0) sort first, so ranks counted from the top are unaffected
    pivot_kernel<<<1, 1, 0, st>>>((const Cand*)d_sample, capacity, rank, d_out_pivot_x);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ---- S3: the fused filter pass
extern "C" size_t vb_psis_filter_workspace_bytes(long long n_local) {
    if (n_local < 1) return 0;
    return 2 * align_up((size_t)filter_grid(n_local) * 8, 256) + 256;
}
extern "C" int vb_psis_filter(const double* d_lw, long long n_local, long long index0, const double* d_pivot_x,
                              void* d_out_cand, long long cap, unsigned long long* d_out_count,
                              double* d_out_body_max, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_lw || !d_pivot_x || !d_out_cand || !d_out_count || !d_out_body_max || !d_ws || n_local < 1 || cap < 1)
        return VB_ERR_ARG;
    if (ws_bytes < vb_psis_filter_workspace_bytes(n_local)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    // experiment knobs: VB_EXP_MIXED=1 (fp32 series tail in the exp), VB_FILTER_OCC4=1 (64 registers, 4 CTAs per SM)
    static const bool mixed = getenv("VB_EXP_MIXED") != nullptr && atoi(getenv("VB_EXP_MIXED")) != 0;
    static const bool occ4 = getenv("VB_FILTER_OCC4") != nullptr && atoi(getenv("VB_FILTER_OCC4")) != 0;
    const int G = filter_grid(n_local);
    double* blk_body = (double*)d_ws;
    double* blk_max = (double*)((char*)d_ws + align_up((size_t)G * 8, 256));
    VB_CUDA_TRY(cudaMemsetAsync(d_out_count, 0, 8, st));
    FilterOut fo{(Cand*)d_out_cand, cap, d_out_count, blk_body, blk_max};
    const bool aligned = (reinterpret_cast<uintptr_t>(d_lw) & 15u) == 0;
#define VB_FILTER_LAUNCH(AL, MX, MB)                                                                              \
    do {                                                                                                          \
        static DeviceOnce attr;                                                                                   \
        const int dev_ = current_device();                                                                        \
        if (attr.pending(dev_)) {                                                                                 \
            VB_CUDA_TRY(cudaFuncSetAttribute(filter_kernel<AL, MX, MB>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                             (int)kFSmem));                                                       \
            attr.done(dev_);                                                                                      \
        }                                                                                                         \
        filter_kernel<AL, MX, MB><<<G, kFThreads, kFSmem, st>>>(d_lw, n_local, index0, d_pivot_x, fo);            \
    } while (0)
    if (!aligned) VB_FILTER_LAUNCH(false, 0, 3);
    else if (mixed && occ4) VB_FILTER_LAUNCH(true, 1, 4);
    else if (mixed) VB_FILTER_LAUNCH(true, 1, 3);
    else if (occ4) VB_FILTER_LAUNCH(true, 0, 4);
    else VB_FILTER_LAUNCH(true, 0, 3);
#undef VB_FILTER_LAUNCH
    VB_LAUNCH_CHECK();
    filter_finish_kernel<<<1, 256, 0, st>>>(blk_body, blk_max, G, d_out_body_max, d_out_body_max + 1);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ---- S4: keys relative to the (global) maximum; term0 = body sum rescaled to it
extern "C" int vb_psis_prepare(void* d_cand, long long count, const double* d_M, const double* d_pivot_x,
                               const double* d_body_pre, double* d_out_term0, void* stream) {
    if (!d_cand || !d_M || !d_pivot_x || !d_body_pre || !d_out_term0 || count < 0) return VB_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (count > 0) {
        rekey_kernel<<<grid_for(count, 256, kNumSMs * 4), 256, 0, st>>>((Cand*)d_cand, count, d_M);
        VB_LAUNCH_CHECK();
    }
    derive_kernel<<<1, 1, 0, st>>>(d_M, d_pivot_x, d_body_pre, nullptr, d_out_term0);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

namespace vb {
struct RefineWs { unsigned int* hist; long long* counts2; double* pivot_u; double* blk; };
static size_t refine_ws_bytes() { return (size_t)kHistBins * 4 + 256 + 256 + align_up((size_t)kSmallGrid * 8, 256); }
static RefineWs refine_ws(char* ws) {
    RefineWs r;
    r.hist = (unsigned int*)ws;
    r.counts2 = (long long*)(ws + (size_t)kHistBins * 4);
    r.pivot_u = (double*)(ws + (size_t)kHistBins * 4 + 256);
    r.blk = (double*)(ws + (size_t)kHistBins * 4 + 512);
    return r;
}
__global__ void pivot_u_kernel(const double* M, const double* px, double* out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *out = *M - *px;
}
// keep the ~`keep` largest of `count` re-keyed candidates (histogram of u); the dropped ones add exp(-u) to *out_term
static int refine_to(const Cand* in, long long count, const double* d_M, const double* d_pivot_x, long long keep,
                     Cand* out, unsigned long long* d_out_count, double* d_out_term, char* ws, cudaStream_t st,
                     const unsigned long long* cnt_ptr = nullptr) {
    const RefineWs r = refine_ws(ws);
    pivot_u_kernel<<<1, 1, 0, st>>>(d_M, d_pivot_x, r.pivot_u);
    VB_LAUNCH_CHECK();
    VB_CUDA_TRY(cudaMemsetAsync(r.hist, 0, (size_t)kHistBins * 4, st));
    VB_CUDA_TRY(cudaMemsetAsync(d_out_count, 0, 8, st));
    hist_kernel<<<grid_for(count, 512 * 8, kNumSMs), 512, 0, st>>>(in, count, r.pivot_u, r.hist, cnt_ptr);
    VB_LAUNCH_CHECK();
    hist_scan_kernel<<<1, 1024, 0, st>>>(r.hist, keep, r.counts2);
    VB_LAUNCH_CHECK();
    compact2_kernel<<<kSmallGrid, 256, 0, st>>>(in, count, r.pivot_u, r.counts2, out, d_out_count, r.blk, cnt_ptr);
    VB_LAUNCH_CHECK();
    sum_blocks_kernel<<<1, 256, 0, st>>>(r.blk, kSmallGrid, d_out_term);
    VB_LAUNCH_CHECK();
    return VB_OK;
}
}  // namespace vb

// ---- S4b (sharded callers): shrink the local candidates to about `keep` before the exchange
extern "C" size_t vb_psis_refine_workspace_bytes(void) { return refine_ws_bytes() + 256; }
extern "C" int vb_psis_refine(const void* d_cand, long long count, const double* d_M, const double* d_pivot_x,
                              long long keep, void* d_out_cand, unsigned long long* d_out_count,
                              double* d_out_term, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_cand || !d_M || !d_pivot_x || !d_out_cand || !d_out_count || !d_out_term || !d_ws || count < 0 || keep < 1)
        return VB_ERR_ARG;
    if (ws_bytes < vb_psis_refine_workspace_bytes()) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    if (count > 2 * keep && count > 8192)
        return refine_to((const Cand*)d_cand, count, d_M, d_pivot_x, keep, (Cand*)d_out_cand, d_out_count, d_out_term,
                         (char*)d_ws, st);
    if (count > 0)
        VB_CUDA_TRY(cudaMemcpyAsync(d_out_cand, d_cand, (size_t)count * sizeof(Cand), cudaMemcpyDeviceToDevice, st));
    set_u64_kernel<<<1, 1, 0, st>>>(d_out_count, (unsigned long long)count);
    VB_LAUNCH_CHECK();
    set_double_kernel<<<1, 1, 0, st>>>(d_out_term, 0.0);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ---- S4c (sharded callers, world > 1): GLOBAL refinement.  The histogram bins are the same on every rank (they
// depend on the all-reduced maximum and the common pivot only), so the ranks all-reduce the 8192 counters between
// these two calls and keep what lies above ONE global threshold: the gathered candidate set is ~keep in total
// instead of ~keep per rank, and the replicated sort does not grow with the number of ranks.
extern "C" int vb_psis_hist_bins(void) { return kHistBins; }
extern "C" int vb_psis_refine_hist(const void* d_cand, long long count, const double* d_M, const double* d_pivot_x,
                                   unsigned int* d_hist, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_cand || !d_M || !d_pivot_x || !d_hist || !d_ws || count < 0) return VB_ERR_ARG;
    if (ws_bytes < vb_psis_refine_workspace_bytes()) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const RefineWs r = refine_ws((char*)d_ws);
    pivot_u_kernel<<<1, 1, 0, st>>>(d_M, d_pivot_x, r.pivot_u);
    VB_LAUNCH_CHECK();
    VB_CUDA_TRY(cudaMemsetAsync(d_hist, 0, (size_t)kHistBins * 4, st));
    if (count > 0) {
        hist_kernel<<<grid_for(count, 512 * 8, kNumSMs), 512, 0, st>>>((const Cand*)d_cand, count, r.pivot_u, d_hist);
        VB_LAUNCH_CHECK();
    }
    return VB_OK;
}
extern "C" int vb_psis_refine_select(const void* d_cand, long long count, const double* d_M, const double* d_pivot_x,
                                     const unsigned int* d_hist_global, long long keep_global, void* d_out_cand,
                                     unsigned long long* d_out_count, double* d_out_term, void* d_ws, size_t ws_bytes,
                                     void* stream) {
    if (!d_cand || !d_M || !d_pivot_x || !d_hist_global || !d_out_cand || !d_out_count || !d_out_term || !d_ws ||
        count < 0 || keep_global < 1)
        return VB_ERR_ARG;
    if (ws_bytes < vb_psis_refine_workspace_bytes()) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const RefineWs r = refine_ws((char*)d_ws);
    pivot_u_kernel<<<1, 1, 0, st>>>(d_M, d_pivot_x, r.pivot_u);
    VB_LAUNCH_CHECK();
    VB_CUDA_TRY(cudaMemsetAsync(d_out_count, 0, 8, st));
    hist_scan_kernel<<<1, 1024, 0, st>>>(d_hist_global, keep_global, r.counts2);
    VB_LAUNCH_CHECK();
    compact2_kernel<<<kSmallGrid, 256, 0, st>>>((const Cand*)d_cand, count, r.pivot_u, r.counts2, (Cand*)d_out_cand,
                                                d_out_count, r.blk);
    VB_LAUNCH_CHECK();
    sum_blocks_kernel<<<1, 256, 0, st>>>(r.blk, kSmallGrid, d_out_term);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ---- S4 / S4b with the candidate count in device memory (sharded, device-driven protocol: viabel_b200/psis_dist.py)
// prepare + local histogram: keys relative to the global maximum, term0, this rank's 8192 bin counters
extern "C" int vb_psis_prepare_hist_dev(void* d_cand, long long cap, const unsigned long long* d_count, const double* d_M,
                                        const double* d_pivot_x, const double* d_body_pre, double* d_out_term0,
                                        unsigned int* d_hist, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_cand || !d_count || !d_M || !d_pivot_x || !d_body_pre || !d_out_term0 || !d_hist || !d_ws || cap < 1) return VB_ERR_ARG;
    if (ws_bytes < vb_psis_refine_workspace_bytes()) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const RefineWs r = refine_ws((char*)d_ws);
    rekey_kernel<<<grid_for(cap, 256, kNumSMs * 4), 256, 0, st>>>((Cand*)d_cand, cap, d_M, d_count);
    VB_LAUNCH_CHECK();
    derive_kernel<<<1, 1, 0, st>>>(d_M, d_pivot_x, d_body_pre, r.pivot_u, d_out_term0);
    VB_LAUNCH_CHECK();
    VB_CUDA_TRY(cudaMemsetAsync(d_hist, 0, (size_t)kHistBins * 4, st));
    hist_kernel<<<grid_for(cap, 512 * 8, kNumSMs), 512, 0, st>>>((const Cand*)d_cand, cap, r.pivot_u, d_hist, d_count);
    VB_LAUNCH_CHECK();
    return VB_OK;
}
// send buffer layout (rows of two 64-bit words): [0, capx) kept candidates | row capx: (term0, term1) |
// row capx + 1: (kept count, flags << 32 | 0) | row capx + 2: (shard size, filter count)
__global__ void send_header_kernel(Cand* __restrict__ send, long long capx, const double* __restrict__ terms,
                                   const unsigned long long* __restrict__ kept, const unsigned long long* __restrict__ c1,
                                   long long cap1, long long n_local) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long flags = 0;
    if (*c1 > (unsigned long long)cap1) flags = 2;                   // the filter pass overflowed this rank's buffer
    else if (*kept > (unsigned long long)capx) flags = 7;            // kept more than the exchange slot holds
    send[capx].key = (unsigned long long)__double_as_longlong(terms[0]);
    send[capx].idx = (unsigned long long)__double_as_longlong(terms[1]);
    send[capx + 1].key = *kept;
    send[capx + 1].idx = flags;
    send[capx + 2].key = (unsigned long long)n_local;
    send[capx + 2].idx = *c1;
}
extern "C" int vb_psis_refine_select_dev(const void* d_cand, long long cap, const unsigned long long* d_count,
                                         const double* d_M, const double* d_pivot_x, const unsigned int* d_hist_global,
                                         long long keep_global, void* d_send, long long capx, double* d_terms,
                                         long long n_local, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_cand || !d_count || !d_M || !d_pivot_x || !d_hist_global || !d_send || !d_terms || !d_ws || cap < 1 || capx < 1 ||
        keep_global < 1)
        return VB_ERR_ARG;
    if (ws_bytes < vb_psis_refine_workspace_bytes()) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const RefineWs r = refine_ws((char*)d_ws);
    unsigned long long* kept = (unsigned long long*)(r.counts2 + 8);          // spare words of the 256-byte counts2 block
    VB_CUDA_TRY(cudaMemsetAsync(kept, 0, 8, st));
    hist_scan_kernel<<<1, 1024, 0, st>>>(d_hist_global, keep_global, r.counts2);
    VB_LAUNCH_CHECK();
    compact2_kernel<<<kSmallGrid, 256, 0, st>>>((const Cand*)d_cand, cap, r.pivot_u, r.counts2, (Cand*)d_send, kept, r.blk,
                                                d_count, capx);
    VB_LAUNCH_CHECK();
    sum_blocks_kernel<<<1, 256, 0, st>>>(r.blk, kSmallGrid, d_terms + 1);
    VB_LAUNCH_CHECK();
    send_header_kernel<<<1, 1, 0, st>>>((Cand*)d_send, capx, d_terms, kept, d_count, cap, n_local);
    VB_LAUNCH_CHECK();
    return VB_OK;
}
// recv (world, capx + 3, 2 words) -> dense candidate list in rank order, the 2 * world body terms in rank order, the
// total count and a status: any rank's flag, a shard size that differs from the expected one, or a total that does
// not fit capx
__global__ void __launch_bounds__(256)
gather_compact_kernel(const Cand* __restrict__ recv, int world, long long capx, const long long* __restrict__ sizes,
                      Cand* __restrict__ out, unsigned long long* __restrict__ total_out, double* __restrict__ terms_out,
                      int* __restrict__ status_out) {
    const long long row = capx + 3;
    long long total = 0;
    int st = 0;
    for (int r = 0; r < world; ++r) {
        const Cand h = recv[r * row + capx + 1];
        if (h.idx && !st) st = (int)h.idx;
        if (sizes && (long long)recv[r * row + capx + 2].key != sizes[r] && !st) st = 9;
        total += (long long)h.key;
    }
    if (total > capx && !st) st = 8;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *total_out = st ? 0ull : (unsigned long long)total;
        *status_out = st;
        for (int r = 0; r < world; ++r) {
            terms_out[2 * r] = __longlong_as_double((long long)recv[r * row + capx].key);
            terms_out[2 * r + 1] = __longlong_as_double((long long)recv[r * row + capx].idx);
        }
    }
    if (st) return;
    long long base = 0;
    for (int r = 0; r < world; ++r) {
        const long long c = (long long)recv[r * row + capx + 1].key;
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < c; i += (long long)gridDim.x * blockDim.x)
            out[base + i] = recv[r * row + i];
        base += c;
    }
}
extern "C" int vb_psis_gather_compact(const void* d_recv, int world, long long capx, const long long* d_sizes_or_null,
                                      void* d_out_cands, unsigned long long* d_out_total, double* d_out_terms,
                                      int* d_out_status, void* stream) {
    if (!d_recv || !d_out_cands || !d_out_total || !d_out_terms || !d_out_status || world < 1 || capx < 1) return VB_ERR_ARG;
    gather_compact_kernel<<<kSmallGrid, 256, 0, (cudaStream_t)stream>>>((const Cand*)d_recv, world, capx, d_sizes_or_null,
                                                                      (Cand*)d_out_cands, d_out_total, d_out_terms, d_out_status);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ---- S5: replicated tail work on the (gathered) re-keyed candidates
namespace vb {
struct TailLayout {
    long long cap_sort;
    size_t off_state, off_counts, off_cnt2, off_term, off_refine, off_sort, off_x2, off_tailnew, off_bs, off_L, off_w,
        off_blk3, off_blk4, off_blk5, off_sort2, off_binoff, off_cursor, off_meta, total;
};
static TailLayout tail_layout(long long count, long long K1) {
    TailLayout L;
    L.cap_sort = pad_pow2(count);
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o += align_up(bytes, 256); return at; };
    L.off_state = take(8 * 8);
    L.off_counts = take(4 * 8);
    L.off_cnt2 = take(8);
    L.off_term = take(8);
    L.off_refine = take(refine_ws_bytes());
    L.off_sort = take((size_t)L.cap_sort * sizeof(Cand));
    L.off_x2 = take((size_t)(K1 + 1) * 8);
    L.off_tailnew = take((size_t)(K1 + 1) * 8);
    const long long m_max = 30 + (long long)sqrt((double)(K1 + 1)) + 1;
    L.off_bs = take((size_t)m_max * 8);
    L.off_L = take((size_t)m_max * 8 * (1 + kGpdSplit));
    L.off_w = take((size_t)m_max * 8);
    L.off_blk3 = take((size_t)kSmallGrid * 8);
    L.off_blk4 = take((size_t)kSmallGrid * 8);
    L.off_blk5 = take((size_t)kSmallGrid * 8);
    L.off_sort2 = take((size_t)L.cap_sort * sizeof(Cand));       // scatter target of the counting sort
    L.off_binoff = take((size_t)(kHistBins + 1) * 4);
    L.off_cursor = take((size_t)kHistBins * 4);
    L.off_meta = take(16);
    L.total = o;
    return L;
}
}  // namespace vb

extern "C" size_t vb_psis_tail_workspace_bytes(long long count, long long K1) {
    if (count < 1 || K1 < 0) return 0;
    return tail_layout(count, K1).total + 256;
}

extern "C" int vb_psis_tail(const void* d_cands, long long count, long long K1, const double* d_M,
                            const double* d_pivot_x, const double* d_terms, int n_terms,
                            double* d_out_diag, long long* d_out_tail_idx_or_null,
                            void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_cands || !d_M || !d_pivot_x || !d_terms || !d_out_diag || !d_ws || n_terms < 1 || K1 < 0) return VB_ERR_ARG;
    if (count < K1 + 1) return VB_ERR_ARG;
    if (ws_bytes < vb_psis_tail_workspace_bytes(count, K1)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const TailLayout L = tail_layout(count, K1);
    char* ws = (char*)(((uintptr_t)d_ws + 255) & ~(uintptr_t)255);
    double* state = (double*)(ws + L.off_state);
    long long* counts = (long long*)(ws + L.off_counts);
    unsigned long long* cnt2 = (unsigned long long*)(ws + L.off_cnt2);
    double* term_ref = (double*)(ws + L.off_term);
    Cand* sort_buf = (Cand*)(ws + L.off_sort);
    double* x2 = (double*)(ws + L.off_x2);
    double* tail_new = (double*)(ws + L.off_tailnew);
    double* bs = (double*)(ws + L.off_bs);
    double* Lq = (double*)(ws + L.off_L);
    double* wq = (double*)(ws + L.off_w);
    double* blk3 = (double*)(ws + L.off_blk3);
    double* blk4 = (double*)(ws + L.off_blk4);
    double* blk5 = (double*)(ws + L.off_blk5);

    Cand* sort2 = (Cand*)(ws + L.off_sort2);
    unsigned int* binoff = (unsigned int*)(ws + L.off_binoff);
    unsigned int* cursor = (unsigned int*)(ws + L.off_cursor);
    long long* meta = (long long*)(ws + L.off_meta);
    const RefineWs rw = refine_ws(ws + L.off_refine);
    static const bool no_binsort = getenv("VB_PSIS_BITONIC") != nullptr && atoi(getenv("VB_PSIS_BITONIC")) != 0;

    long long C2 = count;
    long long h[3] = {0, 0, 0};          // kept candidates, largest bin, total over the kept bins
    bool binned = false;
    if (count > kDirectSortMax) {
        int rc = refine_to((const Cand*)d_cands, count, d_M, d_pivot_x, K1 + 1, sort_buf, cnt2, term_ref,
                           ws + L.off_refine, st);
        if (rc != VB_OK) return rc;
        bin_scan_kernel<<<1, 1024, 0, st>>>(rw.hist, rw.counts2, binoff, meta);      // kept == bins <= b*
        VB_LAUNCH_CHECK();
        VB_CUDA_TRY(cudaMemcpyAsync(&h[0], cnt2, 8, cudaMemcpyDeviceToHost, st));
        VB_CUDA_TRY(cudaMemcpyAsync(&h[1], meta, 16, cudaMemcpyDeviceToHost, st));
        VB_CUDA_TRY(cudaStreamSynchronize(st));
        C2 = h[0];
        if (C2 < K1 + 1 || C2 > count) return VB_ERR_UNSUPPORTED;
        binned = true;
    } else {
        VB_CUDA_TRY(cudaMemcpyAsync(sort_buf, d_cands, (size_t)count * sizeof(Cand), cudaMemcpyDeviceToDevice, st));
        set_double_kernel<<<1, 1, 0, st>>>(term_ref, 0.0);
        VB_LAUNCH_CHECK();
        if (count > 8192 && !no_binsort) {
            // (sharded callers hand over an already refined set): bin the candidates as the refinement would
            pivot_u_kernel<<<1, 1, 0, st>>>(d_M, d_pivot_x, rw.pivot_u);
            VB_LAUNCH_CHECK();
            VB_CUDA_TRY(cudaMemsetAsync(rw.hist, 0, (size_t)kHistBins * 4, st));
            hist_kernel<<<grid_for(count, 512 * 8, kNumSMs), 512, 0, st>>>(sort_buf, count, rw.pivot_u, rw.hist);
            VB_LAUNCH_CHECK();
            bin_scan_kernel<<<1, 1024, 0, st>>>(rw.hist, nullptr, binoff, meta);
            VB_LAUNCH_CHECK();
            VB_CUDA_TRY(cudaMemcpyAsync(&h[1], meta, 16, cudaMemcpyDeviceToHost, st));
            VB_CUDA_TRY(cudaStreamSynchronize(st));
            binned = true;
        }
    }
    int rc = VB_OK;
    if (binned && !no_binsort && C2 > 8192 && h[1] <= kBinSortMax && h[2] == C2) {
        const long long* bstar = (count > kDirectSortMax) ? rw.counts2 : nullptr;
        VB_CUDA_TRY(cudaMemsetAsync(cursor, 0, (size_t)kHistBins * 4, st));
        bin_scatter_kernel<<<grid_for(C2, 256, kNumSMs * 4), 256, 0, st>>>(sort_buf, meta, rw.pivot_u, binoff, cursor, sort2);
        VB_LAUNCH_CHECK();
        bin_rank_kernel<<<kNumSMs * 4, 256, 0, st>>>(sort2, meta, bstar, binoff, sort_buf);
        VB_LAUNCH_CHECK();
    } else {
        const long long n_pad = pad_pow2(C2);
        pad_cands_kernel<<<grid_for(n_pad - C2 + 1, 256, kNumSMs * 4), 256, 0, st>>>(sort_buf, C2, n_pad);
        VB_LAUNCH_CHECK();
        rc = bitonic_sort(sort_buf, n_pad, st);
        if (rc != VB_OK) return rc;
    }

    tail_bounds_kernel<<<1, 1, 0, st>>>(sort_buf, C2, K1, state, counts);
    VB_LAUNCH_CHECK();
    tail_excess_kernel<<<kSmallGrid, 256, 0, st>>>(sort_buf, state, counts, x2);
    VB_LAUNCH_CHECK();
    rc = run_gpd_fit(x2, counts, K1 + 1, bs, Lq, wq, blk3, kSmallGrid, state, st);
    if (rc != VB_OK) return rc;
    tail_smooth_kernel<<<kSmallGrid, 256, 0, st>>>(sort_buf, state, counts, tail_new, d_out_tail_idx_or_null, blk4);
    VB_LAUNCH_CHECK();
    cand_body_kernel<<<kSmallGrid, 256, 0, st>>>(sort_buf, counts, 0, 0x7fffffffffffffffll, blk5);
    VB_LAUNCH_CHECK();
    SumList lists;
    lists.p[0] = d_terms;  lists.n[0] = n_terms;          // per-rank body sums, rank order
    lists.p[1] = term_ref; lists.n[1] = 1;                // dropped by the refinement above
    lists.p[2] = blk5;     lists.n[2] = kSmallGrid;       // sorted candidates below the cutoff
    lists.p[3] = blk4;     lists.n[3] = kSmallGrid;       // the (smoothed) tail
    psis_finish_kernel<<<1, 256, 0, st>>>(lists, d_M, counts, C2, state, d_out_diag, nullptr);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ---- S5, device-driven: the candidate count stays in device memory (d_count, at most `cap` entries in d_cands) and
// the decisions the host-driven path takes after reading counts back are replaced by psis_check_kernel: when a
// precondition fails the remaining kernels turn inert and d_out_diag[7] = -(status) tells the caller to run the
// host-driven path.  No host read, no synchronisation: ~25 dependent launches go out back to back.
static int* tail_status_ptr(char* ws, const vb::TailLayout& L) { return (int*)(ws + L.off_meta + 64); }
extern "C" int vb_psis_tail_dev(const void* d_cands, long long cap, const unsigned long long* d_count, long long K1,
                                const double* d_M, const double* d_pivot_x, const double* d_terms, int n_terms,
                                const unsigned long long* d_count1, long long cap1, const int* d_status_in_or_null,
                                double* d_out_diag, long long* d_out_tail_idx_or_null,
                                void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_cands || !d_count || !d_M || !d_pivot_x || !d_terms || !d_out_diag || !d_ws || n_terms < 1 || K1 < 0) return VB_ERR_ARG;
    if (cap < K1 + 1) return VB_ERR_ARG;
    if (ws_bytes < vb_psis_tail_workspace_bytes(cap, K1)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const TailLayout L = tail_layout(cap, K1);
    char* ws = (char*)(((uintptr_t)d_ws + 255) & ~(uintptr_t)255);
    double* state = (double*)(ws + L.off_state);
    long long* counts = (long long*)(ws + L.off_counts);
    unsigned long long* cnt2 = (unsigned long long*)(ws + L.off_cnt2);
    double* term_ref = (double*)(ws + L.off_term);
    Cand* sort_buf = (Cand*)(ws + L.off_sort);
    double* x2 = (double*)(ws + L.off_x2);
    double* tail_new = (double*)(ws + L.off_tailnew);
    double* bs = (double*)(ws + L.off_bs);
    double* Lq = (double*)(ws + L.off_L);
    double* wq = (double*)(ws + L.off_w);
    double* blk3 = (double*)(ws + L.off_blk3);
    double* blk4 = (double*)(ws + L.off_blk4);
    double* blk5 = (double*)(ws + L.off_blk5);
    Cand* sort2 = (Cand*)(ws + L.off_sort2);
    unsigned int* binoff = (unsigned int*)(ws + L.off_binoff);
    unsigned int* cursor = (unsigned int*)(ws + L.off_cursor);
    long long* meta = (long long*)(ws + L.off_meta);
    int* status = tail_status_ptr(ws, L);
    const RefineWs rw = refine_ws(ws + L.off_refine);
    int rc = refine_to((const Cand*)d_cands, cap, d_M, d_pivot_x, K1 + 1, sort_buf, cnt2, term_ref, ws + L.off_refine, st,
                       d_count);
    if (rc != VB_OK) return rc;
    bin_scan_kernel<<<1, 1024, 0, st>>>(rw.hist, rw.counts2, binoff, meta);
    VB_LAUNCH_CHECK();
    psis_check_kernel<<<1, 1, 0, st>>>(d_count1 ? d_count1 : d_count, d_count1 ? cap1 : cap, cnt2, meta, K1, kBinSortMax, status,
                                       d_status_in_or_null);
    VB_LAUNCH_CHECK();
    VB_CUDA_TRY(cudaMemsetAsync(cursor, 0, (size_t)kHistBins * 4, st));
    bin_scatter_kernel<<<grid_for(cap, 256, kNumSMs * 4), 256, 0, st>>>(sort_buf, meta, rw.pivot_u, binoff, cursor, sort2, status);
    VB_LAUNCH_CHECK();
    bin_rank_kernel<<<kNumSMs * 4, 256, 0, st>>>(sort2, meta, rw.counts2, binoff, sort_buf, status);
    VB_LAUNCH_CHECK();
    tail_bounds_kernel<<<1, 1, 0, st>>>(sort_buf, cap, K1, state, counts, cnt2, status);
    VB_LAUNCH_CHECK();
    tail_excess_kernel<<<kSmallGrid, 256, 0, st>>>(sort_buf, state, counts, x2);
    VB_LAUNCH_CHECK();
    rc = run_gpd_fit(x2, counts, K1 + 1, bs, Lq, wq, blk3, kSmallGrid, state, st);
    if (rc != VB_OK) return rc;
    tail_smooth_kernel<<<kSmallGrid, 256, 0, st>>>(sort_buf, state, counts, tail_new, d_out_tail_idx_or_null, blk4);
    VB_LAUNCH_CHECK();
    cand_body_kernel<<<kSmallGrid, 256, 0, st>>>(sort_buf, counts, 0, 0x7fffffffffffffffll, blk5);
    VB_LAUNCH_CHECK();
    SumList lists;
    lists.p[0] = d_terms;  lists.n[0] = n_terms;
    lists.p[1] = term_ref; lists.n[1] = 1;
    lists.p[2] = blk5;     lists.n[2] = kSmallGrid;
    lists.p[3] = blk4;     lists.n[3] = kSmallGrid;
    psis_finish_kernel<<<1, 256, 0, st>>>(lists, d_M, counts, cap, state, d_out_diag, nullptr, cnt2, status);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ---- S6: out = (x - M) - lse over this shard, then the smoothed tail entries it owns
extern "C" int vb_psis_write(const double* d_lw, double* d_out, long long n_local, long long index0,
                             const double* d_M, long long count, long long K1, const void* d_tail_ws, void* stream) {
    if (!d_lw || !d_out || !d_M || !d_tail_ws || n_local < 1 || d_lw == d_out) return VB_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    const TailLayout L = tail_layout(count, K1);
    const char* ws = (const char*)(((uintptr_t)d_tail_ws + 255) & ~(uintptr_t)255);
    const double* state = (const double*)(ws + L.off_state);
    const long long* counts = (const long long*)(ws + L.off_counts);
    const Cand* sort_buf = (const Cand*)(ws + L.off_sort);
    const double* tail_new = (const double*)(ws + L.off_tailnew);
    normalise_kernel<<<grid_for((n_local + 1) / 2, 256, kNumSMs * 8), 256, 0, st>>>(d_lw, d_out, n_local, d_M, state + 5);
    VB_LAUNCH_CHECK();
    tail_scatter_kernel<<<kSmallGrid, 256, 0, st>>>(sort_buf, counts, tail_new, state, d_out, index0, n_local);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// S6 with the alpha = 2 divergence partials of the RAW weights of this shard (relative to the global maximum M)
extern "C" size_t vb_psis_div_workspace_bytes(void) { return (size_t)kNumSMs * 8 * 4 * sizeof(double) + 256; }
extern "C" int vb_psis_write_div(const double* d_lw, double* d_out, long long n_local, long long index0,
                                 const double* d_M, long long count, long long K1, const void* d_tail_ws,
                                 double* d_out_div_partials, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_lw || !d_out || !d_M || !d_tail_ws || !d_out_div_partials || !d_ws || n_local < 1 || d_lw == d_out) return VB_ERR_ARG;
    if (ws_bytes < vb_psis_div_workspace_bytes()) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const TailLayout L = tail_layout(count, K1);
    const char* ws = (const char*)(((uintptr_t)d_tail_ws + 255) & ~(uintptr_t)255);
    const double* state = (const double*)(ws + L.off_state);
    const long long* counts = (const long long*)(ws + L.off_counts);
    const Cand* sort_buf = (const Cand*)(ws + L.off_sort);
    const double* tail_new = (const double*)(ws + L.off_tailnew);
    double* blocks = (double*)(((uintptr_t)d_ws + 255) & ~(uintptr_t)255);
    const int G = grid_for((n_local + 1) / 2, kNDThreads, kNumSMs * 8);
    normalise_div_kernel<<<G, kNDThreads, 0, st>>>(d_lw, d_out, n_local, d_M, state + 5, blocks);
    VB_LAUNCH_CHECK();
    normalise_div_finish_kernel<<<1, 256, 0, st>>>(d_lw, blocks, G, n_local, d_M, d_out_div_partials);
    VB_LAUNCH_CHECK();
    tail_scatter_kernel<<<kSmallGrid, 256, 0, st>>>(sort_buf, counts, tail_new, state, d_out, index0, n_local);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ---- single-GPU composite: psislw for one column
namespace vb {
struct SmoothLayout {
    long long cap1;
    size_t off_sample, off_pivot, off_count, off_bodymax, off_terms, off_filter, off_cand1, off_tail, off_div, off_pack, total;
};
static SmoothLayout smooth_layout(long long n, double reff) {
    SmoothLayout L;
    const long long K1 = psis_K1(n, reff);
    // candidate buffer: the whole array for small n, otherwise max(2 Mi, n/16) pairs
    L.cap1 = (n <= kDirectSortMax) ? n : ((n / 16 > (1ll << 21)) ? n / 16 : (1ll << 21));
    if (L.cap1 > n) L.cap1 = n;
    if (L.cap1 < K1 + 1) L.cap1 = K1 + 1;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o += align_up(bytes, 256); return at; };
    L.off_sample = take((size_t)kSampleSize * sizeof(Cand));
    L.off_pivot = take(8);
    L.off_count = take(8);
    L.off_bodymax = take(16);
    L.off_terms = take(16);
    L.off_filter = take(vb_psis_filter_workspace_bytes(n));
    L.off_cand1 = take((size_t)L.cap1 * sizeof(Cand));
    L.off_tail = take(vb_psis_tail_workspace_bytes(L.cap1, K1));
    L.off_div = take(vb_psis_div_workspace_bytes());
    L.off_pack = take(24 * sizeof(double));
    L.total = o;
    return L;
}
}  // namespace vb

extern "C" size_t vb_psis_smooth_workspace_bytes(long long n, double reff) {
    if (n < 2 || !(reff > 0.0)) return 0;
    return smooth_layout(n, reff).total + 256;
}

namespace vb {
// packed host result of vb_psis_smooth_host: [0..7] diag, [8..15] divergence statistics (alpha = 2), [16] status word
// ([8..15] are written in place by the divergence merge kernel)
__global__ void smooth_pack_kernel(const double* __restrict__ diag, const int* __restrict__ status_or_null,
                                   double* __restrict__ pack) {
    const int t = threadIdx.x;
    if (t < 8) pack[t] = diag[t];
    else if (t == 16) pack[16] = status_or_null ? (double)*status_or_null : 0.0;
}
static int smooth_impl(const double* d_lw_in, double* d_lw_out, long long n, double reff,
                       double* d_out_diag, long long* d_out_tail_idx_or_null,
                       double* d_out_div_partials_or_null, double* h_out18_or_null, void* d_ws, size_t ws_bytes, void* stream);
}  // namespace vb
extern "C" int vb_bounds_divergence_merge(const double* d_partials, int n_ranks, double alpha, double* d_out, void* stream);

extern "C" int vb_psis_smooth_div(const double* d_lw_in, double* d_lw_out, long long n, double reff,
                                  double* d_out_diag, long long* d_out_tail_idx_or_null,
                                  double* d_out_div_partials_or_null, void* d_ws, size_t ws_bytes, void* stream) {
    return smooth_impl(d_lw_in, d_lw_out, n, reff, d_out_diag, d_out_tail_idx_or_null, d_out_div_partials_or_null, nullptr,
                       d_ws, ws_bytes, stream);
}
extern "C" int vb_psis_smooth(const double* d_lw_in, double* d_lw_out, long long n, double reff,
                              double* d_out_diag, long long* d_out_tail_idx_or_null,
                              void* d_ws, size_t ws_bytes, void* stream) {
    return smooth_impl(d_lw_in, d_lw_out, n, reff, d_out_diag, d_out_tail_idx_or_null, nullptr, nullptr, d_ws, ws_bytes, stream);
}
// vb_psis_smooth_div that also hands the small results to the HOST in the one synchronisation the call has anyway:
// h_out18 = [diag (8) | divergence_bound(alpha = 2) statistics as vb_bounds_divergence writes them (8; zeros when no
// partials were requested) | path: 0 device-driven, 1 host-driven | reserved].  A caller that wants k-hat and the
// bounds as host numbers (viabel.all_bounds does) then needs no further device read.
extern "C" int vb_psis_smooth_host(const double* d_lw_in, double* d_lw_out, long long n, double reff,
                                   double* d_out_diag, long long* d_out_tail_idx_or_null,
                                   double* d_out_div_partials_or_null, double* h_out18, void* d_ws, size_t ws_bytes,
                                   void* stream) {
    if (!h_out18) return VB_ERR_ARG;
    return smooth_impl(d_lw_in, d_lw_out, n, reff, d_out_diag, d_out_tail_idx_or_null, d_out_div_partials_or_null, h_out18,
                       d_ws, ws_bytes, stream);
}
// the one-pass statistics divergence_bound(alpha = 2) needs come from the same final pass over the weights
// (layout of vb_bounds_divergence_partials; finalise with vb_bounds_divergence_merge)
static int vb::smooth_impl(const double* d_lw_in, double* d_lw_out, long long n, double reff,
                           double* d_out_diag, long long* d_out_tail_idx_or_null,
                           double* d_out_div_partials_or_null, double* h_out, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_lw_in || !d_lw_out || !d_out_diag || !d_ws || !(reff > 0.0)) return VB_ERR_ARG;
    if (n <= 1) return VB_ERR_ARG;                                    // psis.py:143-144
    if (d_lw_in == d_lw_out) return VB_ERR_ARG;                       // streaming loads assume no aliasing
    const SmoothLayout L = smooth_layout(n, reff);
    if (ws_bytes < L.total + 256) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)(((uintptr_t)d_ws + 255) & ~(uintptr_t)255);
    Cand* sample = (Cand*)(ws + L.off_sample);
    double* pivot_x = (double*)(ws + L.off_pivot);
    unsigned long long* count = (unsigned long long*)(ws + L.off_count);
    double* body_max = (double*)(ws + L.off_bodymax);
    double* terms = (double*)(ws + L.off_terms);
    Cand* cand1 = (Cand*)(ws + L.off_cand1);
    const long long K1 = psis_K1(n, reff);
    const bool direct = (n <= kDirectSortMax);
    int rc;
    if (!direct) {
        rc = vb_psis_sample(d_lw_in, n, 0, sample, stream);
        if (rc != VB_OK) return rc;
    }
    // Device-driven attempt: every stage is enqueued without reading a count back; ONE status word is read at the
    // end (the caller reads the diagnostics then anyway).  Falls through to the host-driven sequence below when a
    // precondition failed (too few candidates for the first pivot, heavy ties, tiny problems).
    static const bool host_only = getenv("VB_PSIS_HOST") != nullptr && atoi(getenv("VB_PSIS_HOST")) != 0;
    if (!direct && !host_only) {
        // first pivot at sample rank >= 10 of 4096: the candidates are then ~0.25 % of the weights instead of the 1 % the
        // host-driven loop starts from (4x fewer to stage, re-key, bin).  Fewer than K1 + 1 true values above the
        // rank-10 sample needs >= 10 sample hits where ~1 is expected (P ~ 1e-7 for exchangeable weights); the status
        // word catches it and the host-driven sequence below redoes the call from the 1 % pivot.
        double frac0 = 4.0 * (double)(K1 + 1) / (double)n;
        if (frac0 < 0.0025) frac0 = 0.0025;
        rc = vb_psis_pivot(sample, kSampleSize, kSampleSize, frac0, pivot_x, stream);
        if (rc != VB_OK) return rc;
        rc = vb_psis_filter(d_lw_in, n, 0, pivot_x, cand1, L.cap1, count, body_max, ws + L.off_filter,
                            vb_psis_filter_workspace_bytes(n), stream);
        if (rc != VB_OK) return rc;
        const double* M_dev = body_max + 1;
        rekey_kernel<<<grid_for(L.cap1, 256, kNumSMs * 4), 256, 0, st>>>(cand1, L.cap1, M_dev, count);
        VB_LAUNCH_CHECK();
        derive_kernel<<<1, 1, 0, st>>>(M_dev, pivot_x, body_max, nullptr, terms);
        VB_LAUNCH_CHECK();
        rc = vb_psis_tail_dev(cand1, L.cap1, count, K1, M_dev, pivot_x, terms, 1, nullptr, 0, nullptr, d_out_diag,
                              d_out_tail_idx_or_null, ws + L.off_tail, vb_psis_tail_workspace_bytes(L.cap1, K1), stream);
        if (rc != VB_OK) return rc;
        if (d_out_div_partials_or_null)
            rc = vb_psis_write_div(d_lw_in, d_lw_out, n, 0, M_dev, L.cap1, K1, ws + L.off_tail, d_out_div_partials_or_null,
                                   ws + L.off_div, vb_psis_div_workspace_bytes(), stream);
        else
            rc = vb_psis_write(d_lw_in, d_lw_out, n, 0, M_dev, L.cap1, K1, ws + L.off_tail, stream);
        if (rc != VB_OK) return rc;
        int h_status = 0;
        {
            const TailLayout TL = tail_layout(L.cap1, K1);
            char* tws = (char*)(((uintptr_t)(ws + L.off_tail) + 255) & ~(uintptr_t)255);
            if (h_out) {
                double* pack = (double*)(ws + L.off_pack);
                if (d_out_div_partials_or_null) {
                    rc = vb_bounds_divergence_merge(d_out_div_partials_or_null, 1, 2.0, pack + 8, stream);
                    if (rc != VB_OK) return rc;
                }
                smooth_pack_kernel<<<1, 32, 0, st>>>(d_out_diag, (const int*)tail_status_ptr(tws, TL), pack);
                VB_LAUNCH_CHECK();
                VB_CUDA_TRY(cudaMemcpyAsync(h_out, pack, 17 * sizeof(double), cudaMemcpyDeviceToHost, st));
                VB_CUDA_TRY(cudaStreamSynchronize(st));
                h_status = (int)h_out[16];
                h_out[16] = 0.0; h_out[17] = 0.0;
                if (!d_out_div_partials_or_null) for (int i = 8; i < 16; ++i) h_out[i] = 0.0;
            } else {
                VB_CUDA_TRY(cudaMemcpyAsync(&h_status, tail_status_ptr(tws, TL), sizeof(int), cudaMemcpyDeviceToHost, st));
                VB_CUDA_TRY(cudaStreamSynchronize(st));
            }
        }
        if (h_status == 0) return VB_OK;
    }
    // filter pass, widening the sample quantile until the candidates hold the K1+1 largest
    long long C1 = 0;
    double frac = 4.0 * (double)(K1 + 1) / (double)n;
    if (frac < 0.01) frac = 0.01;
    for (int attempt = 0;; ++attempt) {
        if (direct) {
            set_double_kernel<<<1, 1, 0, st>>>(pivot_x, -INFINITY);
            VB_LAUNCH_CHECK();
        } else {
            rc = vb_psis_pivot(sample, kSampleSize, kSampleSize, frac, pivot_x, stream);
            if (rc != VB_OK) return rc;
        }
        rc = vb_psis_filter(d_lw_in, n, 0, pivot_x, cand1, L.cap1, count, body_max, ws + L.off_filter,
                            vb_psis_filter_workspace_bytes(n), stream);
        if (rc != VB_OK) return rc;
        unsigned long long h = 0;
        VB_CUDA_TRY(cudaMemcpyAsync(&h, count, 8, cudaMemcpyDeviceToHost, st));
        VB_CUDA_TRY(cudaStreamSynchronize(st));
        C1 = (long long)h;
        if (C1 >= K1 + 1 && C1 <= L.cap1) break;
        if (direct || attempt >= 8) return VB_ERR_UNSUPPORTED;       // cannot happen for direct
        if (C1 > L.cap1) return VB_ERR_UNSUPPORTED;                  // more ties at the top than the buffer holds
        frac *= 4.0;                                                  // too few: widen
        if (frac > 1.0) frac = 1.0;
    }
    const double* M_ptr = body_max + 1;                               // the maximum is a candidate
    rc = vb_psis_prepare(cand1, C1, M_ptr, pivot_x, body_max, terms, stream);
    if (rc != VB_OK) return rc;
    rc = vb_psis_tail(cand1, C1, K1, M_ptr, pivot_x, terms, 1, d_out_diag, d_out_tail_idx_or_null,
                      ws + L.off_tail, vb_psis_tail_workspace_bytes(L.cap1, K1), stream);
    if (rc != VB_OK) return rc;
    if (d_out_div_partials_or_null)
        rc = vb_psis_write_div(d_lw_in, d_lw_out, n, 0, M_ptr, C1, K1, ws + L.off_tail, d_out_div_partials_or_null,
                               ws + L.off_div, vb_psis_div_workspace_bytes(), stream);
    else
        rc = vb_psis_write(d_lw_in, d_lw_out, n, 0, M_ptr, C1, K1, ws + L.off_tail, stream);
    if (rc != VB_OK || !h_out) return rc;
    double* pack = (double*)(ws + L.off_pack);
    if (d_out_div_partials_or_null) {
        rc = vb_bounds_divergence_merge(d_out_div_partials_or_null, 1, 2.0, pack + 8, stream);
        if (rc != VB_OK) return rc;
    }
    smooth_pack_kernel<<<1, 32, 0, st>>>(d_out_diag, nullptr, pack);
    VB_LAUNCH_CHECK();
    VB_CUDA_TRY(cudaMemcpyAsync(h_out, pack, 17 * sizeof(double), cudaMemcpyDeviceToHost, st));
    VB_CUDA_TRY(cudaStreamSynchronize(st));
    h_out[16] = 1.0; h_out[17] = 0.0;
    if (!d_out_div_partials_or_null) for (int i = 8; i < 16; ++i) h_out[i] = 0.0;
    return VB_OK;
}

// ------------------------------------------------------------------------------------------ standalone ops
// (vb_max / vb_max_workspace_bytes are declared in include/viabel_b200.h and defined in bounds.cu)
extern "C" size_t vb_gpd_fit_workspace_bytes(long long n) {
    if (n < 2) return 0;
    const long long m_max = 30 + (long long)sqrt((double)n) + 1;
    return 256 + 256 + (3 + kGpdSplit) * align_up((size_t)m_max * 8, 256) + align_up((size_t)kNumSMs * 2 * 8, 256);
}

extern "C" int vb_gpd_fit(const double* d_x_sorted, long long n, double* d_out_k_sigma,
                          void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_x_sorted || !d_out_k_sigma || !d_ws || n <= 1) return VB_ERR_ARG;   // psis.py:250-251
    if (ws_bytes < vb_gpd_fit_workspace_bytes(n)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)d_ws;
    const long long m_max = 30 + (long long)sqrt((double)n) + 1;
    const size_t mb = align_up((size_t)m_max * 8, 256);
    double* state = (double*)ws;
    long long* counts = (long long*)(ws + 256);
    double* bs = (double*)(ws + 512);
    double* Lq = (double*)(ws + 512 + mb);
    double* wq = (double*)(ws + 512 + (2 + kGpdSplit) * mb);
    double* blk = (double*)(ws + 512 + (3 + kGpdSplit) * mb);
    // gpdfitnew itself has no n <= 4 shortcut (that lives in psislw): always fit
    set_counts_kernel<<<1, 1, 0, st>>>(counts, n, 0, 30 + (long long)sqrt((double)n), 1);
    VB_LAUNCH_CHECK();
    int rc = run_gpd_fit(d_x_sorted, counts, n, bs, Lq, wq, blk, kNumSMs * 2, state, st);
    if (rc != VB_OK) return rc;
    VB_CUDA_TRY(cudaMemcpyAsync(d_out_k_sigma, state + 2, 16, cudaMemcpyDeviceToDevice, st));
    return VB_OK;
}

// ascending sort of a float64 array (gpdfitnew(sort=True), psis.py:254-265): monotone keys + the bitonic network
namespace vb {
__global__ void sort_fill_kernel(const double* __restrict__ x, long long n, long long n_pad, Cand* __restrict__ c) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_pad; i += (long long)gridDim.x * blockDim.x) {
        Cand e;
        e.key = i < n ? f64_to_key(x[i]) : 0xffffffffffffffffull;
        e.idx = (unsigned long long)i;
        c[i] = e;
    }
}
__global__ void sort_unpack_kernel(const Cand* __restrict__ c, long long n, double* __restrict__ out) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = key_to_f64(c[i].key);
}
}  // namespace vb
extern "C" size_t vb_sort_f64_workspace_bytes(long long n) { return n < 1 ? 0 : (size_t)pad_pow2(n) * sizeof(Cand) + 256; }
extern "C" int vb_sort_f64(const double* d_x, long long n, double* d_out, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_x || !d_out || !d_ws || n < 1) return VB_ERR_ARG;
    if (ws_bytes < vb_sort_f64_workspace_bytes(n)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    Cand* c = (Cand*)(((uintptr_t)d_ws + 255) & ~(uintptr_t)255);
    const long long n_pad = pad_pow2(n);
    sort_fill_kernel<<<grid_for(n_pad, 256, kNumSMs * 4), 256, 0, st>>>(d_x, n, n_pad, c);
    VB_LAUNCH_CHECK();
    const int rc = bitonic_sort(c, n_pad, st);
    if (rc != VB_OK) return rc;
    sort_unpack_kernel<<<grid_for(n, 256, kNumSMs * 4), 256, 0, st>>>(c, n, d_out);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// gpdfitnew(return_quadrature=True) (psis.py:314-331): the quadrature points ks_j = mean_r log1p(-b_j x_r) with the
// weak prior applied, and the UNNORMALISED weights w_j = 1 / sum_i exp(L_i - L_j) of all m points (the caller drops
// the negligible ones, psis.py:300-305, and normalises: output formatting of <= a few hundred numbers)
namespace vb {
__global__ void gpd_quadrature_kernel(const double* __restrict__ parts, const double* __restrict__ wq, long long n,
                                      long long m, double* __restrict__ ks, double* __restrict__ w) {
    for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < m; j += (long long)gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < kGpdSplit; ++k) s += parts[j * kGpdSplit + k];
        const double a = 10.0;
        ks[j] = (s / (double)n) * (double)n / ((double)n + a) + a * 0.5 / ((double)n + a);
        w[j] = wq[j];
    }
}
}  // namespace vb
extern "C" int vb_gpd_fit_quadrature(const double* d_x_sorted, long long n, double* d_out_k_sigma, double* d_out_ks,
                                     double* d_out_w, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_out_ks || !d_out_w) return VB_ERR_ARG;
    const int rc = vb_gpd_fit(d_x_sorted, n, d_out_k_sigma, d_ws, ws_bytes, stream);
    if (rc != VB_OK) return rc;
    char* ws = (char*)d_ws;
    const long long m_max = 30 + (long long)sqrt((double)n) + 1;
    const size_t mb = align_up((size_t)m_max * 8, 256);
    const double* Lq = (const double*)(ws + 512 + mb);
    const double* wq = (const double*)(ws + 512 + (2 + kGpdSplit) * mb);
    const long long m = 30 + (long long)sqrt((double)n);
    gpd_quadrature_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(Lq + m_max, wq, n, m, d_out_ks, d_out_w);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

namespace vb {
__global__ void gpd_inv_kernel(const double* __restrict__ p, long long n, double k, double sigma, double* __restrict__ out) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double pi = p[i];
        double v;
        if (!(sigma > 0.0)) v = NAN;
        else if (pi > 0.0 && pi < 1.0) v = gpinv_inside(pi, k, sigma);
        else if (pi == 0.0) v = 0.0;                                       // psis.py:372
        else if (pi == 1.0) v = (k >= 0.0) ? INFINITY : -sigma / k;        // psis.py:373-376
        else v = NAN * sigma;
        out[i] = v;
    }
}
__global__ void __launch_bounds__(256)
expsum_blocks_kernel(const double* __restrict__ x, long long n, const double* __restrict__ M_ptr, double* __restrict__ block_out) {
    __shared__ double scratch[32];
    const double M = *M_ptr;
    double s = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        s += exp(x[i] - M);
    s = block_sum(s, scratch);
    if (threadIdx.x == 0) block_out[blockIdx.x] = s;
}
__global__ void lse_finish_kernel(const double* __restrict__ blocks, int G, const double* __restrict__ M_ptr, double* __restrict__ out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double s = 0.0;
    for (int g = 0; g < G; ++g) s += blocks[g];
    *out = log(s) + *M_ptr;                                                // psis.py:387-394
}
__global__ void sub_scalar_kernel(double* __restrict__ x, long long n, double v) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        x[i] -= v;
}
}  // namespace vb

extern "C" int vb_gpd_inv(const double* d_p, long long n, double k, double sigma, double* d_out, void* stream) {
    if (!d_p || !d_out || n < 1) return VB_ERR_ARG;
    gpd_inv_kernel<<<grid_for(n, 256, kNumSMs * 8), 256, 0, (cudaStream_t)stream>>>(d_p, n, k, sigma, d_out);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" size_t vb_logsumexp_workspace_bytes(long long n) {
    return 256 + (size_t)kNumSMs * 8 * 8 + 256 + vb_max_workspace_bytes(n);
}

// dst (cols, rows) = scale * src (rows, cols)': the layout change psislw / psisloo make for 2-D input (one contiguous
// row per weight set; psisloo also negates: lw = -log_lik, psis.py:101), through 32 x 33 shared-memory tiles so that
// both the reads and the writes are coalesced.
namespace vb {
__global__ void __launch_bounds__(256)
transpose_scale_kernel(const double* __restrict__ src, long long rows, long long cols, double scale, double* __restrict__ dst) {
    __shared__ double tile[32][33];
    const long long tiles_c = (cols + 31) / 32, tiles_r = (rows + 31) / 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;          // 32 x 8 threads
    for (long long t = blockIdx.x; t < tiles_r * tiles_c; t += gridDim.x) {
        const long long r0 = (t / tiles_c) * 32, c0 = (t % tiles_c) * 32;
#pragma unroll
        for (int k = 0; k < 32; k += 8) {
            const long long r = r0 + ty + k, c = c0 + tx;
            tile[ty + k][tx] = (r < rows && c < cols) ? scale * src[r * cols + c] : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 32; k += 8) {
            const long long c = c0 + ty + k, r = r0 + tx;
            if (c < cols && r < rows) dst[c * rows + r] = tile[tx][ty + k];
        }
        __syncthreads();
    }
}
}  // namespace vb
extern "C" int vb_transpose_scale_f64(const double* d_src, long long rows, long long cols, double scale, double* d_dst,
                                      void* stream) {
    if (!d_src || !d_dst || rows < 1 || cols < 1 || d_src == d_dst) return VB_ERR_ARG;
    const long long tiles = ((rows + 31) / 32) * ((cols + 31) / 32);
    transpose_scale_kernel<<<grid_for(tiles, 1, kNumSMs * 8), 256, 0, (cudaStream_t)stream>>>(d_src, rows, cols, scale, d_dst);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" int vb_logsumexp(const double* d_x, long long n, double* d_out, void* d_ws, size_t ws_bytes, void* stream) {
    if (!d_x || !d_out || !d_ws || n < 1) return VB_ERR_ARG;
    if (ws_bytes < vb_logsumexp_workspace_bytes(n)) return VB_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)d_ws;
    double* M_ptr = (double*)ws;
    double* blocks = (double*)(ws + 256);
    void* max_ws = ws + 256 + (size_t)kNumSMs * 8 * 8 + 256;
    int rc = vb_max(d_x, n, M_ptr, (char*)max_ws, vb_max_workspace_bytes(n), stream);
    if (rc != VB_OK) return rc;
    const int G = grid_for(n, 256, kNumSMs * 8);
    expsum_blocks_kernel<<<G, 256, 0, st>>>(d_x, n, M_ptr, blocks);
    VB_LAUNCH_CHECK();
    lse_finish_kernel<<<1, 1, 0, st>>>(blocks, G, M_ptr, d_out);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

extern "C" int vb_psis_normalise(double* d_lw_out, long long n_local, double lse, void* stream) {
    if (!d_lw_out || n_local < 1) return VB_ERR_ARG;
    sub_scalar_kernel<<<grid_for(n_local, 256, kNumSMs * 8), 256, 0, (cudaStream_t)stream>>>(d_lw_out, n_local, lse);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

// ------------------------------------------------------------------------------------------ batched columns (N3)
// psislw on an (n, m) matrix (psis.py:136-137,163: one independent smoothing per column) and the psisloo
// reduction (psis.py:68-109) for the shapes PSIS-LOO has: a few thousand draws per column, many columns.
// ONE CTA per column.  Like the single-column pipeline it never sorts the column: a 256-entry strided sample gives
// a pivot with ~3 (K+1) elements above it, those are compacted into shared memory and sorted there (<= 2048
// (key, index) pairs instead of the 4096-8192 of the column), then exact cutoff / strict tail -> GPD fit (one warp
// per quadrature point) -> inverse-CDF smoothing -> log-sum-exp -> out = value - lse, with the body read from
// global memory (L2) in index order.  A column whose candidate set does not fit (heavy ties) or falls short is
// flagged and redone by a second launch of the same kernel in FULL mode (every element a candidate); short
// columns (n <= 1024) go straight to FULL mode.  Same comparisons and formulas as the single-column path above
// (tail set and order are bit-exact; k-hat and the weights differ from it by summation order only).
namespace vb {

constexpr int kColThreads = 256;
constexpr int kColMaxN = 8192;                 // FULL mode: 64 KiB of keys + 16 KiB of 16-bit indices
constexpr int kColMaxQ = 128;                  // quadrature points: 30 + sqrt(n2) <= 30 + sqrt(0.2 * 8192) = 70
constexpr int kColSelCap = 2048;               // candidate capacity of the selection mode (20 KiB)
constexpr int kColSample = 256;

struct ColSmem {
    double scratch[32];
    double bs[kColMaxQ], L[kColMaxQ], w[kColMaxQ];
    double st[8];                              // xcut, exp(xcut), k, sigma, smoothed, lse, b, spare
    long long cnt[4];                          // n2, first tail position, quadrature points, fit runs
    unsigned long long pivot_key;
    unsigned int ncand;
    int bail;
};

// ascending bitonic sort of (keys, idxs)[0, n_pad) by (key, index); n_pad a power of two
__device__ __forceinline__ void col_bitonic(unsigned long long* keys, unsigned short* idxs, int n_pad, int tid) {
    for (int k = 2; k <= n_pad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < n_pad / 2; t += kColThreads) {
                const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), q = i | j;
                const unsigned long long ka = keys[i], kb = keys[q];
                const unsigned short ia = idxs[i], ib = idxs[q];
                const bool b_less = kb < ka || (kb == ka && ib < ia);
                if (b_less == ((i & k) == 0)) { keys[i] = kb; keys[q] = ka; idxs[i] = ib; idxs[q] = ia; }
            }
            __syncthreads();
        }
    }
}


// cap: capacity (power of two) of the candidate arrays in shared memory.  SELECT: candidates = elements at or above
// a sample pivot; otherwise every element.  status (or null): SELECT launches set status[col] = 1 when the column
// must be redone; FULL launches that receive status skip the columns whose status is 0.
template <bool SELECT>
__global__ void __launch_bounds__(kColThreads)
psis_columns_kernel(const double* __restrict__ lw_cols, double* __restrict__ out_cols, int n, int cap, int K1,
                    double* __restrict__ out_k, double* __restrict__ out_loo_or_null, int* __restrict__ status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ColSmem& S = *reinterpret_cast<ColSmem*>(smem_raw);
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem_raw + ((sizeof(ColSmem) + 15) & ~(size_t)15));
    double* x2 = reinterpret_cast<double*>(keys + cap);         // [K1]: tail excesses, ascending; later the new values
    unsigned short* idxs = reinterpret_cast<unsigned short*>(x2 + (K1 > 0 ? K1 : 1));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = kColThreads / 32;
    if (!SELECT && status != nullptr && status[blockIdx.x] == 0) return;
    const double* x = lw_cols + (size_t)blockIdx.x * n;
    double* out = out_cols + (size_t)blockIdx.x * n;

    // max (psis.py:165)
    double M = -INFINITY;
    for (int i = tid; i < n; i += kColThreads) M = fmax(M, x[i]);
    M = block_max(M, S.scratch);

    int C, C_pad;
    if (SELECT) {
        // pivot from a strided sample: about 3 (K1 + 1) + 64 elements are expected at or above it
        for (int s = tid; s < kColSample; s += kColThreads) {
            keys[s] = u_to_key(M - x[(long long)s * n / kColSample]);
            idxs[s] = (unsigned short)s;
        }
        if (tid == 0) { S.ncand = 0u; S.bail = 0; }
        __syncthreads();
        col_bitonic(keys, idxs, kColSample, tid);
        if (tid == 0) {
            long long rank_top = ((long long)kColSample * (3ll * (K1 + 1) + 64) + n - 1) / n;
            if (rank_top < 4) rank_top = 4;
            if (rank_top >= kColSample) S.bail = 1;                   // the candidates would be most of the column
            else S.pivot_key = keys[kColSample - rank_top];
        }
        __syncthreads();
        if (!S.bail) {
            const unsigned long long pk = S.pivot_key;
            __syncthreads();                                          // everyone holds the pivot: the arrays are free
            for (int i = tid; i < n; i += kColThreads) {
                const unsigned long long key = u_to_key(M - x[i]);
                if (key >= pk) {                                      // larger key == larger weight
                    const unsigned int pos = atomicAdd(&S.ncand, 1u);
                    if (pos < (unsigned int)cap) { keys[pos] = key; idxs[pos] = (unsigned short)i; }
                }
            }
            __syncthreads();
            if (tid == 0 && (S.ncand > (unsigned int)cap || S.ncand < (unsigned int)(K1 + 1))) S.bail = 1;
            __syncthreads();
        }
        if (S.bail) {
            if (tid == 0) status[blockIdx.x] = 1;
            return;
        }
        C = (int)S.ncand;
        C_pad = 2;
        while (C_pad < C) C_pad <<= 1;
        for (int i = C + tid; i < C_pad; i += kColThreads) { keys[i] = ~0ull; idxs[i] = (unsigned short)0xffff; }
    } else {
        C = n;
        C_pad = cap;
        for (int i = tid; i < C_pad; i += kColThreads) {
            keys[i] = (i < n) ? u_to_key(M - x[i]) : ~0ull;
            idxs[i] = (i < n) ? (unsigned short)i : (unsigned short)0xffff;
        }
    }
    __syncthreads();
    // ascending (key, index) == ascending shifted weight, ties by ascending index (psis.py:167 with kind='stable')
    col_bitonic(keys, idxs, C_pad, tid);
    // cutoff and strict tail (psis.py:157,169-176), as tail_bounds_kernel: the K1 + 1 largest are all candidates
    if (tid == 0) {
        const double u_cut = key_to_u(keys[C - 1 - K1]);
        const double xcut = fmax(-u_cut, -kNegLogTiny);
        const double u_eff = -xcut;
        int lo = 0, hi = C;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (key_to_u(keys[mid]) < u_eff) hi = mid; else lo = mid + 1;
        }
        const int n2 = C - lo;
        S.st[0] = xcut;
        S.st[1] = exp(xcut);
        S.cnt[0] = n2;
        S.cnt[1] = lo;
        S.cnt[2] = 30 + (long long)sqrt((double)n2);
        S.cnt[3] = (n2 > 4) ? 1 : 0;
    }
    __syncthreads();
    const int n2 = (int)S.cnt[0], p0 = (int)S.cnt[1], mq = (int)S.cnt[2];
    const bool fit = S.cnt[3] != 0;
    const double ec = S.st[1], u_eff = -S.st[0];
    for (int r = tid; r < n2; r += kColThreads) x2[r] = exp(-key_to_u(keys[p0 + r])) - ec;     // psis.py:184-185
    __syncthreads();
    if (fit) {                                                                                  // psis.py:266-323
        const double x_q = x2[(int)((double)n2 / 4.0 + 0.5) - 1], x_max = x2[n2 - 1];
        for (int j = warp; j < mq; j += nw) {
            const double b = gpd_b(j, mq, x_q, x_max);
            double a = 0.0;
            for (int r = lane; r < n2; r += 32) a += log1p(-b * x2[r]);
            a = warp_sum(a);
            if (lane == 0) {
                const double ks = a / (double)n2;
                S.bs[j] = b;
                S.L[j] = (double)n2 * (log(-(b / ks)) - ks - 1.0);
            }
        }
        __syncthreads();
        for (int j = warp; j < mq; j += nw) {
            double a = 0.0;
            const double Lj = S.L[j];
            for (int i = lane; i < mq; i += 32) a += exp(S.L[i] - Lj);
            a = warp_sum(a);
            if (lane == 0) S.w[j] = 1.0 / a;
        }
        __syncthreads();
        double sw = 0.0, sbw = 0.0;
        for (int j = tid; j < mq; j += kColThreads) {
            const double w = S.w[j];
            if (w >= 10.0 * DBL_EPSILON) { sw += w; sbw += S.bs[j] * w; }
        }
        sw = block_sum(sw, S.scratch);
        sbw = block_sum(sbw, S.scratch);
        const double b = sbw / sw;
        double a = 0.0;
        for (int r = tid; r < n2; r += kColThreads) a += log1p(-b * x2[r]);
        a = block_sum(a, S.scratch);
        if (tid == 0) {
            double k = a / (double)n2;
            const double sigma = -k / b;
            k = k * (double)n2 / ((double)n2 + 10.0) + 5.0 / ((double)n2 + 10.0);
            S.st[2] = k;
            S.st[3] = sigma;
            S.st[4] = (k >= 1.0 / 3.0 && !isinf(k)) ? 1.0 : 0.0;
        }
    } else if (tid == 0) {
        S.st[2] = INFINITY; S.st[3] = NAN; S.st[4] = 0.0;                                       // psis.py:177-179
    }
    __syncthreads();
    const bool smoothed = S.st[4] != 0.0;
    const double kh = S.st[2], sigma = S.st[3];
    // new tail values (psis.py:190-198); x2 is reused for them
    double se = 0.0;
    for (int r = tid; r < n2; r += kColThreads) {
        double v;
        if (smoothed) {
            v = log(gpinv_inside(((double)r + 0.5) / (double)n2, kh, sigma) + ec);
            if (v > 0.0) v = 0.0;
        } else {
            v = -key_to_u(keys[p0 + r]);
        }
        x2[r] = v;
        se += exp(v);
    }
    // the body: every element at or below the cutoff, read back from global memory (L2) in index order
    for (int i = tid; i < n; i += kColThreads) {
        const double u = M - x[i];
        if (!(u < u_eff)) se += exp(-u);
    }
    se = block_sum(se, S.scratch);
    const double lse = log(se);                                                                  // psis.py:200
    // write back; optional leave-one-out term log sum_i exp(out_i - lw_i) (psis.py:104-107, lw = -log_lik)
    double lmax = -INFINITY;
    for (int i = tid; i < n; i += kColThreads) {
        const double u = M - x[i];
        if (!(u < u_eff)) {
            const double v = -u - lse;
            out[i] = v;
            lmax = fmax(lmax, v - x[i]);
        }
    }
    for (int r = tid; r < n2; r += kColThreads) {
        const int i = (int)idxs[p0 + r];
        const double v = x2[r] - lse;
        out[i] = v;
        lmax = fmax(lmax, v - x[i]);
    }
    if (tid == 0) out_k[blockIdx.x] = kh;
    if (out_loo_or_null) {
        lmax = block_max(lmax, S.scratch);
        double a = 0.0;
        for (int i = tid; i < n; i += kColThreads) {
            const double u = M - x[i];
            if (!(u < u_eff)) a += exp(((-u - lse) - x[i]) - lmax);
        }
        for (int r = tid; r < n2; r += kColThreads) a += exp(((x2[r] - lse) - x[(int)idxs[p0 + r]]) - lmax);
        a = block_sum(a, S.scratch);
        if (tid == 0) out_loo_or_null[blockIdx.x] = log(a) + lmax;                               // psis.py:390-395
    }
}

}  // namespace vb

extern "C" long long vb_psis_columns_max_n(void) { return vb::kColMaxN; }

extern "C" int vb_psis_smooth_columns(const double* d_lw_cols, double* d_out_cols, long long n, long long m,
                                      double reff, double* d_out_k, double* d_out_loo_or_null, int* d_status,
                                      void* stream) {
    using namespace vb;
    if (!d_lw_cols || !d_out_cols || !d_out_k || m < 1 || !(reff > 0.0)) return VB_ERR_ARG;
    if (n <= 1) return VB_ERR_ARG;                                    // psis.py:143-144
    if (n > kColMaxN) return VB_ERR_UNSUPPORTED;
    if (d_lw_cols == d_out_cols) return VB_ERR_ARG;
    long long n_pad = 2;
    while (n_pad < n) n_pad <<= 1;
    const long long K1 = psis_K1(n, reff);
    if (K1 + 1 > n) return VB_ERR_UNSUPPORTED;                        // Reff so small that the cutoff leaves the column
    cudaStream_t st = (cudaStream_t)stream;
    auto smem_for = [&](long long cap) {
        return ((sizeof(ColSmem) + 15) & ~(size_t)15) + (size_t)cap * 8 + (size_t)(K1 > 0 ? K1 : 1) * sizeof(double) +
               (size_t)cap * 2;
    };
    static DeviceMaxSize configured[2];
    const int dev = current_device();
    // selection mode when it can pay: the column is longer than the candidate buffer and the expected candidate
    // count (3 (K1 + 1) + 64) fits it with room to spare; d_status (m ints) is required for it
    static const bool no_select = getenv("VB_PSIS_COLUMNS_FULL") != nullptr && atoi(getenv("VB_PSIS_COLUMNS_FULL")) != 0;
    const bool select = !no_select && d_status != nullptr && n > kColSelCap && 3 * (K1 + 1) + 64 <= kColSelCap / 2;
    if (select) {
        const size_t smem = smem_for(kColSelCap);
        if (!configured[1].covers(dev, smem)) {
            VB_CUDA_TRY(cudaFuncSetAttribute(psis_columns_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            configured[1].set(dev, smem);
        }
        VB_CUDA_TRY(cudaMemsetAsync(d_status, 0, (size_t)m * sizeof(int), st));
        psis_columns_kernel<true><<<(unsigned)m, kColThreads, smem, st>>>(
            d_lw_cols, d_out_cols, (int)n, kColSelCap, (int)K1, d_out_k, d_out_loo_or_null, d_status);
        VB_LAUNCH_CHECK();
    }
    const size_t smem = smem_for(n_pad);
    if (!configured[0].covers(dev, smem)) {
        VB_CUDA_TRY(cudaFuncSetAttribute(psis_columns_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[0].set(dev, smem);
    }
    // FULL mode: every column, or (after a selection launch) only the flagged ones -- the other CTAs exit at once
    psis_columns_kernel<false><<<(unsigned)m, kColThreads, smem, st>>>(
        d_lw_cols, d_out_cols, (int)n, (int)n_pad, (int)K1, d_out_k, d_out_loo_or_null, select ? d_status : nullptr);
    VB_LAUNCH_CHECK();
    return VB_OK;
}
