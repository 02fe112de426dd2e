// Shared device helpers for the viabel_b200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <atomic>
#include "../../include/viabel_b200.h"
#include "log_table.cuh"

#define VB_CUDA_TRY(expr)                                   \
    do {                                                    \
        cudaError_t _e = (expr);                            \
        if (_e != cudaSuccess) return (int)_e;              \
    } while (0)

#define VB_LAUNCH_CHECK()                                   \
    do {                                                    \
        cudaError_t _e = cudaGetLastError();                \
        if (_e != cudaSuccess) return (int)_e;              \
    } while (0)

namespace vb {

constexpr int kNumSMs = 148;       // B200: 2 dies x 74 SMs
constexpr int kWarp = 32;

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---------------------------------------------------------------------------- per-device host-side caches
// cudaFuncSetAttribute and the SM count are PER DEVICE: a process that moves to another GPU after first use must
// configure the kernels there too.  One bit per device ordinal, set after the configuration succeeded; safe to
// race (two threads may both configure, which is idempotent).
constexpr int kMaxDevices = 64;
static inline int current_device() {
    int d = -1;
    return cudaGetDevice(&d) == cudaSuccess ? d : -1;
}
struct DeviceOnce {
    std::atomic<unsigned long long> bits{0};
    bool pending(int dev) const {
        return dev < 0 || dev >= kMaxDevices || !((bits.load(std::memory_order_acquire) >> dev) & 1ull);
    }
    void done(int dev) {
        if (dev >= 0 && dev < kMaxDevices) bits.fetch_or(1ull << dev, std::memory_order_release);
    }
};
// largest dynamic shared-memory size configured so far, per device (kernels whose request grows with the problem)
struct DeviceMaxSize {
    std::atomic<size_t> v[kMaxDevices];
    bool covers(int dev, size_t bytes) const { return dev >= 0 && dev < kMaxDevices && v[dev].load(std::memory_order_acquire) >= bytes; }
    void set(int dev, size_t bytes) { if (dev >= 0 && dev < kMaxDevices) v[dev].store(bytes, std::memory_order_release); }
};
static inline int device_sm_count() {
    static std::atomic<int> cache[kMaxDevices];
    const int dev = current_device();
    if (dev >= 0 && dev < kMaxDevices) {
        const int c = cache[dev].load(std::memory_order_relaxed);
        if (c > 0) return c;
    }
    int v = 0;
    if (dev < 0 || cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) return kNumSMs;
    if (dev < kMaxDevices) cache[dev].store(v, std::memory_order_relaxed);
    return v;
}

// ---------------------------------------------------------------------------- math traits
template <typename T> struct Math;
template <> struct Math<double> {
    static __device__ __forceinline__ double exp(double x) { return ::exp(x); }
    static __device__ __forceinline__ double log(double x) { return ::log(x); }
    static __device__ __forceinline__ double log1p(double x) { return ::log1p(x); }
    static __device__ __forceinline__ double tanh(double x) { return ::tanh(x); }
    static __device__ __forceinline__ double sqrt(double x) { return ::sqrt(x); }
    static __device__ __forceinline__ double rcp(double x) { return 1.0 / x; }
};
template <> struct Math<float> {
    static __device__ __forceinline__ float exp(float x) { return ::expf(x); }
    static __device__ __forceinline__ float log(float x) { return ::logf(x); }
    static __device__ __forceinline__ float log1p(float x) { return ::log1pf(x); }
    static __device__ __forceinline__ float tanh(float x) { return ::tanhf(x); }
    static __device__ __forceinline__ float sqrt(float x) { return ::sqrtf(x); }
    static __device__ __forceinline__ float rcp(float x) { return 1.0f / x; }
};

// ---------------------------------------------------------------------------- warp / block reductions
__device__ __forceinline__ double shfl_xor_f64(double v, int m) {
    return __shfl_xor_sync(0xffffffffu, v, m);
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v += shfl_xor_f64(v, m);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v = fmax(v, shfl_xor_f64(v, m));
    return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v = fmin(v, shfl_xor_f64(v, m));
    return v;
}
__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
    return v;
}

// Block-wide sum of one double; result valid in thread 0 (and broadcast through smem[0]).
// `scratch` needs >= 32 doubles.  All threads must call.
__device__ __forceinline__ double block_sum(double v, double* scratch) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) scratch[w] = v;
    __syncthreads();
    double r = (threadIdx.x < nw) ? scratch[threadIdx.x] : 0.0;
    if (w == 0) r = warp_sum(r);
    if (threadIdx.x == 0) scratch[0] = r;
    __syncthreads();
    return scratch[0];
}
__device__ __forceinline__ double block_max(double v, double* scratch) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) scratch[w] = v;
    __syncthreads();
    double r = (threadIdx.x < nw) ? scratch[threadIdx.x] : -INFINITY;
    if (w == 0) r = warp_max(r);
    if (threadIdx.x == 0) scratch[0] = r;
    __syncthreads();
    return scratch[0];
}

// exp(a * (m_old - m_new)) that is 0 (not NaN) when the old partial is empty (m_old = -inf).
__device__ __forceinline__ double rescale_factor(double alpha, double m_old, double m_new) {
    return (m_old == -INFINITY) ? 0.0 : ::exp(alpha * (m_old - m_new));
}

// ---------------------------------------------------------------------------- fast fp64 exp for x <= 0
// exp(x) = 2^(k/64) * exp(r), k = rint(x * 64/ln2), |r| <= ln2/128, degree-6 polynomial;
// 12 fp64 ops instead of the ~25 of the library routine.  Max error ~1 ulp (table entries are
// correctly rounded, polynomial truncation 2e-19).  Only for arguments <= 0 (weights relative to a
// running max); results below DBL_MIN go through the library routine.
__device__ const double kExp2Table[64] = {
    0x1.0000000000000p+0, 0x1.02c9a3e778061p+0, 0x1.059b0d3158574p+0, 0x1.0874518759bc8p+0,
    0x1.0b5586cf9890fp+0, 0x1.0e3ec32d3d1a2p+0, 0x1.11301d0125b51p+0, 0x1.1429aaea92de0p+0,
    0x1.172b83c7d517bp+0, 0x1.1a35beb6fcb75p+0, 0x1.1d4873168b9aap+0, 0x1.2063b88628cd6p+0,
    0x1.2387a6e756238p+0, 0x1.26b4565e27cddp+0, 0x1.29e9df51fdee1p+0, 0x1.2d285a6e4030bp+0,
    0x1.306fe0a31b715p+0, 0x1.33c08b26416ffp+0, 0x1.371a7373aa9cbp+0, 0x1.3a7db34e59ff7p+0,
    0x1.3dea64c123422p+0, 0x1.4160a21f72e2ap+0, 0x1.44e086061892dp+0, 0x1.486a2b5c13cd0p+0,
    0x1.4bfdad5362a27p+0, 0x1.4f9b2769d2ca7p+0, 0x1.5342b569d4f82p+0, 0x1.56f4736b527dap+0,
    0x1.5ab07dd485429p+0, 0x1.5e76f15ad2148p+0, 0x1.6247eb03a5585p+0, 0x1.6623882552225p+0,
    0x1.6a09e667f3bcdp+0, 0x1.6dfb23c651a2fp+0, 0x1.71f75e8ec5f74p+0, 0x1.75feb564267c9p+0,
    0x1.7a11473eb0187p+0, 0x1.7e2f336cf4e62p+0, 0x1.82589994cce13p+0, 0x1.868d99b4492edp+0,
    0x1.8ace5422aa0dbp+0, 0x1.8f1ae99157736p+0, 0x1.93737b0cdc5e5p+0, 0x1.97d829fde4e50p+0,
    0x1.9c49182a3f090p+0, 0x1.a0c667b5de565p+0, 0x1.a5503b23e255dp+0, 0x1.a9e6b5579fdbfp+0,
    0x1.ae89f995ad3adp+0, 0x1.b33a2b84f15fbp+0, 0x1.b7f76f2fb5e47p+0, 0x1.bcc1e904bc1d2p+0,
    0x1.c199bdd85529cp+0, 0x1.c67f12e57d14bp+0, 0x1.cb720dcef9069p+0, 0x1.d072d4a07897cp+0,
    0x1.d5818dcfba487p+0, 0x1.da9e603db3285p+0, 0x1.dfc97337b9b5fp+0, 0x1.e502ee78b3ff6p+0,
    0x1.ea4afa2a490dap+0, 0x1.efa1bee615a27p+0, 0x1.f50765b6e4540p+0, 0x1.fa7c1819e90d8p+0};

__device__ __forceinline__ double exp_nonpos(double x) {
    if (!(x > -700.0)) return ::exp(x);            // subnormal / zero / NaN / -inf: library path (rare)
    const double kMagic = 6755399441055744.0;      // 1.5 * 2^52: round-to-nearest-even to integer
    const double t = fma(x, 0x1.71547652b82fep+6, kMagic);
    const int ki = __double2loint(t);              // low word of (t) holds the integer k
    const double kd = t - kMagic;
    double r = fma(kd, -0x1.62e42fee00000p-7, x);
    r = fma(kd, -0x1.a39ef35793c76p-39, r);
    double q = fma(r, 1.0 / 720.0, 1.0 / 120.0);
    q = fma(r, q, 1.0 / 24.0);
    q = fma(r, q, 1.0 / 6.0);
    q = fma(r, q, 0.5);
    q = fma(r, q, 1.0);
    q = q * r;                                      // exp(r) - 1
    const double T = __ldg(&kExp2Table[ki & 63]);
    const double s = fma(T, q, T);
    const int e = ki >> 6;                          // arithmetic shift: floor(k / 64)
    return __hiloint2double(__double2hiint(s) + e * 1048576, __double2loint(s));   // * 2^e
}

// General-range version of the same evaluation (|x| < 700; everything else, incl. NaN, takes the library path).
__device__ __forceinline__ double fast_exp(double x) {
    if (!(fabs(x) < 700.0)) return ::exp(x);
    const double kMagic = 6755399441055744.0;
    const double t = fma(x, 0x1.71547652b82fep+6, kMagic);
    const int ki = __double2loint(t);
    const double kd = t - kMagic;
    double r = fma(kd, -0x1.62e42fee00000p-7, x);
    r = fma(kd, -0x1.a39ef35793c76p-39, r);
    double q = fma(r, 1.0 / 720.0, 1.0 / 120.0);
    q = fma(r, q, 1.0 / 24.0);
    q = fma(r, q, 1.0 / 6.0);
    q = fma(r, q, 0.5);
    q = fma(r, q, 1.0);
    q = q * r;
    const double T = __ldg(&kExp2Table[ki & 63]);
    const double s = fma(T, q, T);
    return __hiloint2double(__double2hiint(s) + (ki >> 6) * 1048576, __double2loint(s));
}

// ---------------------------------------------------------------------------- fast fp64 log
// log(y) = k ln2 + log(c_i) + log1p(r), r = z / c_i - 1 with z = y / 2^k in [0.6875, 1.375) and c_i the midpoint of
// the i-th of 128 bit-pattern intervals (|r| <= 2^-8, degree-6 series); table and a pure-Python check of the
// algorithm: scripts/gen_log_table.py (max error 1 ulp of max(|log y|, 1e-16 absolute near y = 1)).  11 fp64
// operations instead of the ~35 of the library log1p.  Non-positive, subnormal, inf and NaN arguments take the
// library path.
__device__ __forceinline__ double fast_log(double y) {
    const long long ix = __double_as_longlong(y);
    if ((unsigned long long)(ix - 0x0010000000000000ll) >= 0x7fe0000000000000ull) return ::log(y);
    const long long tmp = ix - 0x3fe6000000000000ll;
    const int i = (int)(tmp >> 45) & 127;
    const int k = (int)(tmp >> 52);
    const double z = __longlong_as_double(ix - (tmp & (long long)0xfff0000000000000ull));
    const double2 t = __ldg(reinterpret_cast<const double2*>(&kLogTable[i][0]));
    const double r = fma(z, t.x, -1.0);
    const double kd = (double)k;
    const double hi = fma(kd, 0x1.62e42fefa3800p-1, t.y);
    double p = fma(r, -1.0 / 6.0, 0.2);
    p = fma(r, p, -0.25);
    p = fma(r, p, 1.0 / 3.0);
    p = fma(r, p, -0.5);
    const double lo = fma(kd, 0x1.ef35793c76730p-45, fma(r * r, p, r));
    return hi + lo;
}
// log(1 + v) for v >= 0 with ABSOLUTE accuracy ~1e-16 (the argument is rounded to 1 + v): for sums of O(1) terms
// (log densities), not for callers that need the relative accuracy of log1p near 0.
__device__ __forceinline__ double fast_log1p_abs(double v) { return fast_log(1.0 + v); }
// type-dispatched forms for the templated model / family code (fp32 keeps the library intrinsics)
__device__ __forceinline__ double vexp(double x) { return fast_exp(x); }
__device__ __forceinline__ float vexp(float x) { return ::expf(x); }
__device__ __forceinline__ double vlog1p_abs(double v) { return fast_log1p_abs(v); }
__device__ __forceinline__ float vlog1p_abs(float v) { return ::log1pf(v); }

// Branch-free variant for streaming sums of weights relative to a maximum / pivot: the same
// 1-ulp evaluation for -700 <= x <= 0; arguments below -700 (incl. -inf) are clamped, i.e. give
// e^-700 ~ 1e-304 instead of a smaller number -- invisible in a double-precision sum that holds any
// term of normal size (every sum here contains exp(0) = 1 for the maximum); NaN stays NaN.
// `table` is the 64-entry 2^(j/64) table, normally a shared-memory copy of kExp2Table.
// Coefficients live in the constant bank (not `const`: the compiler must not fold them into
// immediates, which it would re-materialise with two MOVs in front of every DFMA); DFMA reads
// c[bank][offset] operands directly.
static __constant__ double kExpCoef[10] = {
    0x1.71547652b82fep+6,      // 64 / ln 2
    6755399441055744.0,        // 1.5 * 2^52
    -0x1.62e42fee00000p-7,     // -ln2_hi / 64
    -0x1.a39ef35793c76p-39,    // -ln2_lo / 64
    1.0 / 720.0, 1.0 / 120.0, 1.0 / 24.0, 1.0 / 6.0, 0.5, -700.0};
// MIXED = 1 evaluates the tail of the series, h = exp(r) - 1 - r (|h| <= 1.5e-5), in fp32 on the FMA pipe and adds
// it back in fp64: 7 fp64-pipe operations instead of 12, relative error <= 2e-12 (h carries fp32 rounding, 6e-8 * h).
template <int MIXED>
__device__ __forceinline__ double exp_nonpos_flush_v(double x, const double* __restrict__ table) {
    const double kMagic = kExpCoef[1];
    const double xc = !(x <= kExpCoef[9]) ? x : kExpCoef[9];     // unordered compare: NaN passes through
    const double t = fma(xc, kExpCoef[0], kMagic);
    const int ki = __double2loint(t);
    const double kd = t - kMagic;
    double r = fma(kd, kExpCoef[2], xc);
    r = fma(kd, kExpCoef[3], r);
    double q;
    if (MIXED) {
        const float rf = (float)r;
        float h = fmaf(rf, 1.0f / 120.0f, 1.0f / 24.0f);
        h = fmaf(rf, h, 1.0f / 6.0f);
        h = fmaf(rf, h, 0.5f);
        h = h * rf * rf;
        q = r + (double)h;                          // exp(r) - 1
    } else {
        q = fma(r, kExpCoef[4], kExpCoef[5]);
        q = fma(r, q, kExpCoef[6]);
        q = fma(r, q, kExpCoef[7]);
        q = fma(r, q, kExpCoef[8]);
        q = fma(r, q, 1.0);
        q = q * r;                                  // exp(r) - 1
    }
    const double T = table[ki & 63];
    // 2^floor(k/64) folded into the table value (never NaN, stays normal for k >= -64640)
    const double Ts = __hiloint2double(__double2hiint(T) + ((ki & ~63) << 14), __double2loint(T));
    return fma(Ts, q, Ts);
}
__device__ __forceinline__ double exp_nonpos_flush(double x, const double* __restrict__ table) {
    return exp_nonpos_flush_v<0>(x, table);
}

// ---------------------------------------------------------------------------- ordered keys
// Monotone map double -> uint64 (ascending).  -0.0 and +0.0 map to different keys but
// compare like numpy's sort would order equal values only through the index tie-break.
__device__ __host__ __forceinline__ uint64_t f64_to_key(double x) {
    uint64_t b;
#ifdef __CUDA_ARCH__
    b = (uint64_t)__double_as_longlong(x);
#else
    union { double d; uint64_t u; } c; c.d = x; b = c.u;
#endif
    return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __host__ __forceinline__ double key_to_f64(uint64_t k) {
    uint64_t b = (k & 0x8000000000000000ull) ? (k & 0x7fffffffffffffffull) : ~k;
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)b);
#else
    union { double d; uint64_t u; } c; c.u = b; return c.d;
#endif
}

// ---------------------------------------------------------------------------- streaming loads
__device__ __forceinline__ double ld_stream(const double* p) {
    double v;
    asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ double2 ld_stream2(const double2* p) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];"
                 : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}

// ---------------------------------------------------------------------------- mbarrier + 1-D bulk copy (TMA unit)
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                 :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy executed by the TMA unit; completion counted on `bar`.
// dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                         uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        :: "r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

}  // namespace vb
