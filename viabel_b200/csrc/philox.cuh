// Counter-based base draws (Philox4x32-10; Salmon et al., SC'11).
//
// Counter layout (128 bit): c0,c1 = global sample index (64 bit), c2 = lane (which pair of
// coordinates / which rejection attempt), c3 = call offset.  Key (64 bit) = family seed.
// Every variate is a pure function of (seed, call offset, global sample index, coordinate),
// so results are independent of grid shape and of how the sample axis is split over GPUs.
//
// Uniform -> normal by Box-Muller, gamma by Marsaglia-Tsang (2000) with a fixed budget of
// rejection attempts, all in fp32 arithmetic: the base draws carry fp32 resolution
// (|z| <= 6.66, spacing 2^-24 relative) and are then promoted to the compute type.
#pragma once
#include "common.cuh"

namespace vb {

struct Philox4 { uint32_t x, y, z, w; };

__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        c0 = hi1 ^ c1 ^ k0;
        c1 = lo1;
        c2 = hi0 ^ c3 ^ k1;
        c3 = lo0;
        k0 += W0;
        k1 += W1;
    }
    return Philox4{c0, c1, c2, c3};
}

struct PhiloxStream {
    uint32_t k0, k1;      // seed
    uint32_t offset;      // call offset
};

__device__ __forceinline__ Philox4 philox_at(const PhiloxStream& s, uint64_t sample, uint32_t lane) {
    return philox4x32_10((uint32_t)sample, (uint32_t)(sample >> 32), lane, s.offset, s.k0, s.k1);
}

// Every arithmetic step below is written with explicit round-to-nearest intrinsics: the same source inlined into
// different kernels must give the same bits (left to the compiler, a*b+c contracts to an FMA in one kernel and not in
// another, and a Marsaglia-Tsang acceptance that sits on the boundary then flips: 1 draw in ~1e7 differed between the
// stand-alone draw kernel and the fused step).

// (0,1) open interval, 2^-32 resolution before fp32 rounding; never 0, may round to 1-2^-24.
__device__ __forceinline__ float u01(uint32_t u) {
    return fminf(__fmul_rn(__fadd_rn((float)u, 0.5f), 2.3283064365386963e-10f), 0.99999994f);
}

// Two independent standard normals from two words.
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& n0, float& n1) {
    const float r = __fsqrt_rn(__fmul_rn(-2.0f, __logf(u01(a))));
    float s, c;
    __sincosf(__fmul_rn(6.2831853071795865f, __fmul_rn((float)b, 2.3283064365386963e-10f)), &s, &c);
    n0 = __fmul_rn(r, c);
    n1 = __fmul_rn(r, s);
}

// Lane numbering: coordinates use lanes [0, 2^30); rejection retries use lane | (attempt << 30)... kept
// disjoint by construction below (coordinate lanes < 2^24, attempt field in bits 24-27, chi2 marker bit 31).
constexpr uint32_t kLaneChi2 = 0x80000000u;
constexpr int kGammaAttempts = 6;   // acceptance >= 0.95 per attempt for shape >= 1

// Marsaglia-Tsang acceptance test for Gamma(shape, 1), shape >= 1, given a normal and a uniform.
__device__ __forceinline__ bool mt_accept(float d, float c, float xn, float un, float& out) {
    const float t = __fmaf_rn(c, xn, 1.0f);
    const float v = __fmul_rn(__fmul_rn(t, t), t);
    out = __fmul_rn(d, v);
    // log(u) < x^2 / 2 + d - d v + d log(v)
    const float rhs = __fmaf_rn(d, __logf(v), __fadd_rn(__fmaf_rn(__fmul_rn(0.5f, xn), xn, d), -out));
    return (t > 0.0f) && (__logf(un) < rhs);
}

// standard_t(df) variate for (sample, coord): z / sqrt(chi2/df), chi2 = 2 Gamma(df/2).
__device__ __forceinline__ float draw_student_t(const PhiloxStream& s, uint64_t sample, uint32_t coord,
                                                float df) {
    const float shape = __fmul_rn(0.5f, df);
    const float d = __fadd_rn(shape, -(1.0f / 3.0f));
    const float c = rsqrtf(__fmul_rn(9.0f, d));
    Philox4 r = philox_at(s, sample, coord);
    float z, xn;
    box_muller(r.x, r.y, z, xn);
    float g;
    bool ok = mt_accept(d, c, xn, u01(r.z), g);
#pragma unroll 1
    for (int a = 1; a < kGammaAttempts && !ok; ++a) {
        r = philox_at(s, sample, coord | ((uint32_t)a << 24));
        float x0, x1;
        box_muller(r.x, r.y, x0, x1);
        ok = mt_accept(d, c, x0, u01(r.z), g);
        if (!ok) ok = mt_accept(d, c, x1, u01(r.w), g);
    }
    if (!ok) g = d;   // probability < 1e-15; keeps the draw finite
    // chi2/df = 2 g / df
    return __fmul_rn(z, rsqrtf(__fdiv_rn(__fmul_rn(2.0f, g), df)));
}

// All D coordinates of one sample: the same variates as D calls of draw_student_t, bit for bit (same counters, same
// arithmetic), but the Marsaglia-Tsang retries of the whole row are gathered into ONE loop per thread.  A first
// attempt is rejected with probability ~5 %: written per coordinate, 80 % of the warps ran the retry body (a Philox
// block, a Box-Muller pair, two acceptance tests) for one or two lanes after EVERY coordinate; gathered, a warp runs
// it max-over-lanes times per row (2 - 3 instead of ~8 for D = 10), and the body exists once in the instruction
// stream instead of D times.
template <int D, typename T>
__device__ __forceinline__ void draw_student_t_row(const PhiloxStream& s, uint64_t sample, float df, T (&eps)[D]) {
    const float shape = __fmul_rn(0.5f, df);
    const float d = __fadd_rn(shape, -(1.0f / 3.0f));
    const float c = rsqrtf(__fmul_rn(9.0f, d));
    float zs[D];
    uint32_t pending = 0;
#pragma unroll
    for (int j = 0; j < D; ++j) {
        const Philox4 r = philox_at(s, sample, (uint32_t)j);
        float xn, g;
        box_muller(r.x, r.y, zs[j], xn);
        if (mt_accept(d, c, xn, u01(r.z), g)) zs[j] = __fmul_rn(zs[j], rsqrtf(__fdiv_rn(__fmul_rn(2.0f, g), df)));
        else pending |= 1u << j;
    }
    int a = 1;
#pragma unroll 1
    while (pending) {
        const int j = __ffs((int)pending) - 1;
        const Philox4 r = philox_at(s, sample, (uint32_t)j | ((uint32_t)a << 24));
        float x0, x1, g;
        box_muller(r.x, r.y, x0, x1);
        bool ok = mt_accept(d, c, x0, u01(r.z), g);
        if (!ok) ok = mt_accept(d, c, x1, u01(r.w), g);
        ++a;
        if (ok || a >= kGammaAttempts) {
            if (!ok) g = d;                              // probability < 1e-15; keeps the draw finite
            const float sc = rsqrtf(__fdiv_rn(__fmul_rn(2.0f, g), df));
#pragma unroll
            for (int jj = 0; jj < D; ++jj)
                if (jj == j) zs[jj] = __fmul_rn(zs[jj], sc);       // (static indices: the row stays in registers)
            pending &= pending - 1;
            a = 1;
        }
    }
#pragma unroll
    for (int j = 0; j < D; ++j) eps[j] = (T)zs[j];
}

// chisquare(df) for one sample (full-rank family, drawn "first": its own lane space).
__device__ __forceinline__ float draw_chi2(const PhiloxStream& s, uint64_t sample, float df) {
    const float shape = __fmul_rn(0.5f, df);
    const float d = __fadd_rn(shape, -(1.0f / 3.0f));
    const float c = rsqrtf(__fmul_rn(9.0f, d));
    float g = d;
    bool ok = false;
#pragma unroll 1
    for (int a = 0; a < kGammaAttempts && !ok; ++a) {
        const Philox4 r = philox_at(s, sample, kLaneChi2 | (uint32_t)a);
        float x0, x1;
        box_muller(r.x, r.y, x0, x1);
        ok = mt_accept(d, c, x0, u01(r.z), g);
        if (!ok) ok = mt_accept(d, c, x1, u01(r.w), g);
    }
    if (!ok) g = d;
    return __fmul_rn(2.0f, g);
}

// Two standard normals for coordinates (2*pair, 2*pair+1) of a sample.
__device__ __forceinline__ void draw_normal_pair(const PhiloxStream& s, uint64_t sample, uint32_t pair,
                                                 float& n0, float& n1, float& n2, float& n3) {
    const Philox4 r = philox_at(s, sample, pair);
    box_muller(r.x, r.y, n0, n1);
    box_muller(r.z, r.w, n2, n3);
}

}  // namespace vb
