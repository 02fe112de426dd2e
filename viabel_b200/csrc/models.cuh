// Elementwise model plugins: log p(x) and grad log p(x) for one sample held in registers.
// Formulas: SURVEY.md Appendix A.3; sources cited in include/viabel_b200.h.
#pragma once
#include "common.cuh"

namespace vb {

constexpr int kMaxSchools = 8;

// Host-side handle (opaque to C callers).
struct ModelHost {
    int kind;
    int dim;
    // funnel / mixture
    double p0;
    // eight schools
    int n_schools;
    double y[kMaxSchools];
    double sigma[kMaxSchools];
    // regression
    const void* d_X;
    const void* d_y;
    long long n_rows;
    double df;
    double prior_sd;
    int data_dtype;
    const float* d_X_planes;   // or null: tf32 hi/lo planes of X made once by vb_split_f32 (vb_model_regression_attach_split)
    const void* d_X_planes16;  // or null: scaled fp16 hi/lo planes + header made once by vb_split_f16 (..._attach_split_f16)
    // gauss
    const double* d_mean;
    const double* d_prec;
    const float* d_gauss_cache;   // or null: [tf32 hi/lo planes of float(P) | bias = mean' P (fp32)] made once (vb_model_gauss_attach_cache)
};

struct Funnel {
    static constexpr int D = 2;
    struct P { double inv_a, inv_a2, lp0; };       // 1/a, 1/a^2, -log(2 pi) - log a: hoisted to the host
    static inline P make(double a) { return P{1.0 / a, 1.0 / (a * a), -1.8378770664093453 - log(a)}; }
    template <typename T>
    static __device__ __forceinline__ void eval(const P& p, const T (&x)[D], T& lp, T (&g)[D]) {
        const T mu = x[0], ls = x[1];
        const T e = vexp((T)-2 * ls);
        const T mu2e = mu * mu * e;
        const T t = ls * (T)p.inv_a;
        // norm.logpdf(ls, 0, a) + norm.logpdf(mu, 0, exp(ls))
        lp = (T)p.lp0 - (T)0.5 * t * t - ls - (T)0.5 * mu2e;
        g[0] = -mu * e;
        g[1] = -ls * (T)p.inv_a2 - (T)1 + mu2e;
    }
};

struct Mixture {
    static constexpr int D = 1;
    struct P { double c; };
    template <typename T>
    static __device__ __forceinline__ void eval(const P& p, const T (&x)[D], T& lp, T (&g)[D]) {
        const T t = x[0], c = (T)p.c;
        const T a = (T)-0.5 * (t + c) * (t + c), b = (T)-0.5 * (t - c) * (t - c);
        const T hi = a > b ? a : b, lo = a > b ? b : a;
        // logaddexp(a, b) - 1/2 log 2pi - log 2
        lp = hi + Math<T>::log1p(Math<T>::exp(lo - hi)) - (T)0.9189385332046727 - (T)0.6931471805599453;
        g[0] = -t + c * Math<T>::tanh(c * t);
    }
};

template <int J>
struct EightSchools {
    static constexpr int D = J + 2;
    struct P { double y[J]; double inv_sigma[J]; };
    template <typename T>
    static __device__ __forceinline__ void eval(const P& p, const T (&u)[D], T& lp, T (&g)[D]) {
        const T mu = u[0], lt = u[1];
        const T tau = vexp(lt);
        const T tau2 = tau * tau;
        T s_rr = 0, s_ros = 0, s_tros = 0, s_tt = 0;
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const T th = u[2 + j];
            const T is = (T)p.inv_sigma[j];
            const T r = ((T)p.y[j] - mu - tau * th) * is;
            const T ros = r * is;
            s_rr += r * r;
            s_ros += ros;
            s_tros += th * ros;
            s_tt += th * th;
            g[2 + j] = -th + tau * ros;
        }
        lp = -mu * mu * (T)0.02 - vlog1p_abs(tau2 * (T)0.04) + lt - (T)0.5 * s_tt - (T)0.5 * s_rr;
        g[0] = -mu * (T)0.04 + s_ros;
        g[1] = (T)1 - (T)2 * tau2 / ((T)25 + tau2) + tau * s_tros;
    }
};

}  // namespace vb
