// Fused mean-field Monte Carlo step (KLVI / CHIVI) for the elementwise model plugins.
//
// One persistent launch: each CTA streams tiles of base draws (n, D) through shared memory
// with the TMA unit's 1-D bulk copy (cp.async.bulk + mbarrier, 3 stages), or generates them
// with Philox; one thread owns one sample per tile: x = m + exp(l) eps -> log p, grad log p
// (-> log q for CHIVI) -> fp64 accumulators with online log-sum-exp rescaling.  Block
// partials go to the workspace; the last CTA to finish merges them in block order
// (deterministic) and applies the value/gradient formulas of SURVEY.md Appendix A.1.
//
// Reference being replaced: viabel/vb.py:236-245 (black_box_klvi), :248-266 (black_box_chivi)
// on the families of :48-82 and :140-182.
#include <new>
#include "common.cuh"
#include "models.cuh"
#include "philox.cuh"

namespace vb {

constexpr int kMfThreads = 256;
constexpr int kMfStages = 3;
#ifndef VB_MF_CTAS_STREAM
#define VB_MF_CTAS_STREAM 2
#endif
#ifndef VB_MF_CTAS_PHILOX
#define VB_MF_CTAS_PHILOX 2
#endif
constexpr int kMfMinCtasStream = VB_MF_CTAS_STREAM, kMfMinCtasPhilox = VB_MF_CTAS_PHILOX;
constexpr long long kMfSingleCtaMax = 2048;   // samples: at most 8 tiles on one CTA

// Philox stream record kept in device memory by the device-resident optimiser loop: every launch of an
// iteration has identical kernel arguments (so a CUDA graph can replay it) and picks its stream up from here.
struct MfDynStream {
    unsigned long long seed;
    unsigned int offset;
    unsigned int step;
};

struct AdagradArgs {
    int P, window, n_iters, keep_from;
    double epsilon;
    double* var_param;          // [P] in/out
    const double* out;          // [1 + P]: (value, grad) written by the step
    double* ring;               // [window][P]
    const double* rates;        // [n_iters]
    const unsigned long long* seeds;   // [n_iters]
    const unsigned int* offsets;       // [n_iters]
    double* value_history;      // [n_iters]
    double* param_history;      // [n_iters - keep_from][P]
    MfDynStream* dyn;
    // update rule: VB_OPT_ADAGRAD (windowed, vb.py:368-376), VB_OPT_RMSPROP (vb.py:448-462), VB_OPT_ADAM
    // (vb.py:616-629).  The latter two keep their moment vectors in ring[0..P) (second) and ring[P..2P) (first),
    // record the PRE-update iterate of every iteration (keep_from = 0) and read the bias corrections
    // 1 - beta^(i+2) from host-made tables so that no pow() is evaluated on the device.
    int rule = 0;
    double c0 = 0.0, c1 = 0.0, c2 = 0.0, c3 = 0.0;   // rmsprop: decay, 1-decay; adam: beta1, 1-beta1, beta2, 1-beta2
    const double* bias1 = nullptr;                   // adam: 1 - beta1^(i+2), [n_iters]
    const double* bias2 = nullptr;                   // adam: 1 - beta2^(i+2), [n_iters]
};

// One AdaGrad update by one CTA (vb.py:368-376): called from the last CTA of a fused step launch, or from the
// stand-alone update kernel.  Not inlined: keeps the step kernel's register allocation untouched.
__device__ __noinline__ void adagrad_update_block(const AdagradArgs& g) {
    const unsigned int step = g.dyn->step;
    __syncthreads();
    if (step >= (unsigned int)g.n_iters) return;
    const int P = g.P, W = g.window;
    const int cnt = (int)min(step + 1u, (unsigned int)W);
    const unsigned int oldest = step + 1u - (unsigned int)cnt;
    const double rate = g.rates[step];
    if (g.rule != VB_OPT_ADAGRAD) {
        // every product / sum individually rounded, in the order numpy evaluates the reference's expressions
        for (int j = threadIdx.x; j < P; j += blockDim.x) {
            const double gr = g.out[1 + j];
            const double g2 = __dmul_rn(gr, gr);
            double stepv;
            if (g.rule == VB_OPT_RMSPROP) {
                const double second = (step == 0u) ? g2 : __dadd_rn(__dmul_rn(g.ring[j], g.c0), __dmul_rn(g.c1, g2));
                g.ring[j] = second;
                stepv = __ddiv_rn(__dmul_rn(rate, gr), __dsqrt_rn(__dadd_rn(g.epsilon, second)));
            } else {
                double second, first;
                if (step == 0u) {
                    second = __dmul_rn(0.9, g2);                     // vb.py:617-618
                    first = __dmul_rn(0.9, gr);
                } else {
                    second = __dadd_rn(__dmul_rn(g.ring[j], g.c2), __dmul_rn(g.c3, g2));
                    first = __dadd_rn(__dmul_rn(g.ring[P + j], g.c0), __dmul_rn(g.c1, gr));
                }
                g.ring[j] = second;
                g.ring[P + j] = first;
                const double m_hat = __ddiv_rn(first, g.bias1[step]);
                const double v_hat = __ddiv_rn(second, g.bias2[step]);
                stepv = __ddiv_rn(__dmul_rn(rate, m_hat), __dsqrt_rn(__dadd_rn(g.epsilon, v_hat)));
            }
            const double old = g.var_param[j];
            g.param_history[(size_t)step * P + j] = old;             // pre-update iterate, vb.py:464
            g.var_param[j] = __dsub_rn(old, stepv);
        }
    } else
    for (int j = threadIdx.x; j < P; j += blockDim.x) {
        const double gr = g.out[1 + j];
        g.ring[(size_t)(step % W) * P + j] = gr;
        double acc = 0.0;
        for (int k = 0; k < cnt; ++k) {
            const double v = (k == cnt - 1) ? gr : g.ring[(size_t)((oldest + k) % W) * P + j];
            acc = __dadd_rn(acc, __dmul_rn(v, v));
        }
        const double pn = __dsub_rn(g.var_param[j],
                                    __ddiv_rn(__dmul_rn(rate, gr), __dsqrt_rn(__dadd_rn(g.epsilon, acc))));
        g.var_param[j] = pn;
        if (step >= (unsigned int)g.keep_from) g.param_history[(size_t)(step - g.keep_from) * P + j] = pn;
    }
    if (threadIdx.x == 0) {
        g.value_history[step] = g.out[0];
        const unsigned int next = step + 1u;
        g.dyn->step = next;
        if (next < (unsigned int)g.n_iters) {
            g.dyn->seed = g.seeds[next];
            g.dyn->offset = g.offsets[next];
        }
    }
}

template <class Model>
struct MfArgs {
    typename Model::P mp;
    const double* var_param;   // [2D]
    const void* base;          // (n, D) of T or null
    long long n;
    double df;
    double alpha;
    PhiloxStream rng;
    unsigned long long sample_index0;
    double* ws_partials;       // [grid][NP]
    unsigned int* ws_ticket;
    double* out_partials;      // [NP]
    double* out_value;         // or null
    double* out_grad;          // or null
    int use_bulk;              // base pointer is 16-byte aligned
    const MfDynStream* dyn;    // or null: Philox (seed, offset) read from device memory (device-resident optimiser)
    const AdagradArgs* opt;    // or null: AdaGrad update applied by the last CTA after the finalise (fused iteration)
    int iters;                 // > 1 (single-CTA Philox launches with opt): optimiser iterations run inside this launch
    int pd;                    // VB_OBJ_KLVI_PD: the CHIVI pass with alpha = 0 (unit weights, sum of lw), finalised as KLVI
};

// log q(x(lambda); lambda) = D*c0 - sum(l) + core(eps); this returns core(eps).
template <typename T, int FAM, int D>
__device__ __forceinline__ T logq_core(const T (&eps)[D], T df) {
    T s = 0;
    if (FAM == VB_FAMILY_MF_GAUSS) {
#pragma unroll
        for (int j = 0; j < D; ++j) s += eps[j] * eps[j];
        return (T)-0.5 * s;
    } else {
        const T inv_df = Math<T>::rcp(df);
#pragma unroll
        for (int j = 0; j < D; ++j) s += vlog1p_abs(eps[j] * eps[j] * inv_df);
        return (T)-0.5 * (df + (T)1) * s;
    }
}

__host__ __device__ inline double mf_logq_const(int family, double df) {
    if (family == VB_FAMILY_MF_GAUSS) return -0.9189385332046727;   // -1/2 log 2 pi
    return lgamma(0.5 * (df + 1.0)) - lgamma(0.5 * df) - 0.5 * log(df * 3.14159265358979323846);
}

// (value, grad) from merged partials.  Appendix A.1.
__device__ __forceinline__ void mf_finalize_thread(int family, int objective, double alpha,
                                                   const double* var_param, int D,
                                                   const double* part, double n_total,
                                                   double* out_value, double* out_grad, int j) {
    const double M = part[0], sw = part[1], s2 = part[2];
    if (j < D) {
        const double sig = exp(var_param[D + j]);
        const double gm = part[4 + j], gl = part[4 + D + j] * sig;
        if (objective == VB_OBJ_KLVI || objective == VB_OBJ_KLVI_PD) {
            out_grad[j] = -gm / n_total;
            out_grad[D + j] = -(1.0 + gl / n_total);
        } else {
            out_grad[j] = alpha * gm / n_total;
            out_grad[D + j] = alpha * (gl + sw) / n_total;
        }
    }
    if (j == 0) {
        if (objective == VB_OBJ_KLVI) {
            double H = 0.0;
            for (int k = 0; k < D; ++k) H += var_param[D + k];
            if (family == VB_FAMILY_MF_GAUSS) H += 0.5 * D * (1.0 + 1.8378770664093453);
            *out_value = -(H + s2 / n_total);
        } else if (objective == VB_OBJ_KLVI_PD) {
            // vb.py:268-278: -(mean log p - mean log q(x(lambda); lambda)); s2 = sum of lw in the CHIVI pass.  The
            // total derivative of log q along the reparameterised sample is that of -entropy, so the gradient is KLVI's.
            *out_value = -(s2 / n_total);
        } else {
            *out_value = log(sw / n_total) / alpha + M;
        }
    }
}

// Merge R partial vectors in index order.  Thread j produces component j.
__device__ __forceinline__ double mf_merge_component(const double* parts, int R, int NP, int objective,
                                                     double alpha, double M, int j) {
    if (j == 0) return M;
    double acc = 0.0;
    const bool scaled = (objective == VB_OBJ_CHIVI) && (j == 1 || j >= 4);
    for (int r = 0; r < R; ++r) {
        const double f = scaled ? rescale_factor(alpha, parts[(size_t)r * NP], M) : 1.0;
        acc += parts[(size_t)r * NP + j] * f;
    }
    return acc;
}

template <typename T, int FAM, class Model, int OBJ, bool PHILOX>
__global__ void __launch_bounds__(kMfThreads, PHILOX ? kMfMinCtasPhilox : kMfMinCtasStream)
mf_step_kernel(const __grid_constant__ MfArgs<Model> a) {
    constexpr int D = Model::D;
    constexpr int NP = 2 * D + 4;
    constexpr int TILE = kMfThreads;
    constexpr uint32_t kStageBytes = TILE * D * sizeof(T);

    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ uint64_t full_bar[kMfStages];
    __shared__ double red[(kMfThreads / 32) * NP];
    __shared__ double merged[NP];
    __shared__ unsigned int s_ticket;

    const int tid = threadIdx.x;
    const long long n = a.n;
    const long long n_full_tiles = n / TILE;
    const long long n_tiles = (n + TILE - 1) / TILE;

    // location / scale live in shared memory (broadcast reads) to keep registers for the accumulators
    __shared__ T mean[D], sig[D];
    const T df = (T)a.df;
    // Small problems (one CTA) of the device-resident optimiser run `iters` whole iterations inside one launch:
    // step -> finalise -> AdaGrad update -> reload the iterate, with no kernel boundary in between.
    const int reps = (PHILOX && a.opt != nullptr && gridDim.x == 1) ? a.iters : 1;
    const double logq_c = D * mf_logq_const(FAM, a.df);
    for (int rep = 0; rep < reps; ++rep) {
    if (rep > 0) __syncthreads();      // the previous iteration's update and shared-memory reads are complete
    if (tid < D) {
        mean[tid] = (T)a.var_param[tid];
        sig[tid] = (T)exp(a.var_param[D + tid]);
    }
    __syncthreads();
    double sum_l = 0.0;
#pragma unroll
    for (int j = 0; j < D; ++j) sum_l += a.var_param[D + j];
    const double logq0 = logq_c - sum_l;

    double M = (OBJ == VB_OBJ_CHIVI) ? -INFINITY : 0.0;
    double sw = 0.0, s2 = 0.0, cnt = 0.0;
    double ag[D], age[D];
#pragma unroll
    for (int j = 0; j < D; ++j) { ag[j] = 0.0; age[j] = 0.0; }

    const bool bulk = !PHILOX && a.use_bulk;
    const T* base = reinterpret_cast<const T*>(a.base);
    T* stage_buf = reinterpret_cast<T*>(smem_raw);
    PhiloxStream rng = a.rng;
    if (a.dyn != nullptr) {
        if (PHILOX) {
            const unsigned long long sd = a.dyn->seed;
            rng.k0 = (uint32_t)sd;
            rng.k1 = (uint32_t)(sd >> 32);
            rng.offset = a.dyn->offset;
        } else {
            base += (size_t)a.dyn->step * (size_t)n * D;     // external draws: one (n, D) block per iteration
        }
    }

    if (bulk) {
        if (tid == 0) {
#pragma unroll
            for (int s = 0; s < kMfStages; ++s) mbar_init(&full_bar[s], 1);
            mbar_fence_init();
        }
        __syncthreads();
        if (tid == 0) {
#pragma unroll
            for (int s = 0; s < kMfStages; ++s) {
                const long long t = (long long)blockIdx.x + (long long)s * gridDim.x;
                if (t < n_full_tiles) {
                    mbar_expect_tx(&full_bar[s], kStageBytes);
                    bulk_g2s(stage_buf + (size_t)s * TILE * D, base + (size_t)t * TILE * D, kStageBytes,
                             &full_bar[s]);
                }
            }
        }
    }

    int it = 0;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        const long long i = tile * TILE + tid;
        const bool full_tile = tile < n_full_tiles;
        const bool active = i < n;
        T eps[D];
        if (PHILOX) {
            if (active) {
                const uint64_t gi = a.sample_index0 + (uint64_t)i;
                if (FAM == VB_FAMILY_MF_GAUSS) {
#pragma unroll
                    for (int q = 0; q < (D + 3) / 4; ++q) {
                        float z[4];
                        draw_normal_pair(rng, gi, (uint32_t)q, z[0], z[1], z[2], z[3]);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (4 * q + k < D) eps[4 * q + k] = (T)z[k];
                    }
                } else {
                    draw_student_t_row<D, T>(rng, gi, (float)a.df, eps);
                }
            }
        } else if (bulk && full_tile) {
            const int s = it % kMfStages;
            const uint32_t parity = (uint32_t)((it / kMfStages) & 1);
            mbar_wait(&full_bar[s], parity);
            const T* src = stage_buf + (size_t)s * TILE * D + (size_t)tid * D;
#pragma unroll
            for (int j = 0; j < D; ++j) eps[j] = src[j];
            __syncthreads();   // every thread has drained this stage
            if (tid == 0) {
                const long long tn = tile + (long long)kMfStages * gridDim.x;
                if (tn < n_full_tiles) {
                    mbar_expect_tx(&full_bar[s], kStageBytes);
                    bulk_g2s(stage_buf + (size_t)s * TILE * D, base + (size_t)tn * TILE * D, kStageBytes,
                             &full_bar[s]);
                }
            }
        } else if (active) {
            const T* src = base + (size_t)i * D;
#pragma unroll
            for (int j = 0; j < D; ++j) eps[j] = src[j];
        }
        if (!active) continue;

        T x[D], g[D], lp;
#pragma unroll
        for (int j = 0; j < D; ++j) x[j] = mean[j] + sig[j] * eps[j];
        Model::template eval<T>(a.mp, x, lp, g);

        cnt += 1.0;
        if (OBJ == VB_OBJ_KLVI) {
            sw += 1.0;
            s2 += (double)lp;
#pragma unroll
            for (int j = 0; j < D; ++j) {
                ag[j] += (double)g[j];
                age[j] += (double)(g[j] * eps[j]);
            }
        } else {
            const double lw = (double)lp - (logq0 + (double)logq_core<T, FAM, D>(eps, df));
            s2 += lw;
            if (lw > M) {
                const double f = rescale_factor(a.alpha, M, lw);
                sw *= f;
#pragma unroll
                for (int j = 0; j < D; ++j) { ag[j] *= f; age[j] *= f; }
                M = lw;
            }
            const double w = exp(a.alpha * (lw - M));
            sw += w;
#pragma unroll
            for (int j = 0; j < D; ++j) {
                const double wg = w * (double)g[j];
                ag[j] += wg;
                age[j] += wg * (double)eps[j];
            }
        }
    }

    // ---- block reduction ----
    const int lane = tid & 31, warp = tid >> 5;
    constexpr int NW = kMfThreads / 32;
    if (OBJ == VB_OBJ_CHIVI) {
        const double Mw = warp_max(M);
        const double f = rescale_factor(a.alpha, M, Mw);
        sw *= f;
#pragma unroll
        for (int j = 0; j < D; ++j) { ag[j] *= f; age[j] *= f; }
        M = Mw;
    }
    sw = warp_sum(sw); s2 = warp_sum(s2); cnt = warp_sum(cnt);
#pragma unroll
    for (int j = 0; j < D; ++j) { ag[j] = warp_sum(ag[j]); age[j] = warp_sum(age[j]); }
    if (lane == 0) {
        double* r = red + warp * NP;
        r[0] = M; r[1] = sw; r[2] = s2; r[3] = cnt;
#pragma unroll
        for (int j = 0; j < D; ++j) { r[4 + j] = ag[j]; r[4 + D + j] = age[j]; }
    }
    __syncthreads();
    double v = 0.0;
    if (tid < NP) {
        double Mb = red[0];
        for (int w = 1; w < NW; ++w) Mb = fmax(Mb, red[w * NP]);
        v = mf_merge_component(red, NW, NP, OBJ, a.alpha, Mb, tid);
    }
    if (gridDim.x == 1) {
        // one CTA: its partial IS the merged vector (the cross-CTA merge multiplies by exp(0) = 1: same bits)
        if (tid < NP) {
            merged[tid] = v;
            a.out_partials[tid] = v;
        }
    } else {
        if (tid < NP) a.ws_partials[(size_t)blockIdx.x * NP + tid] = v;
        __threadfence();
        __syncthreads();
        if (tid == 0) s_ticket = atomicAdd(a.ws_ticket, 1u);
        __syncthreads();
        if (s_ticket != gridDim.x - 1) return;

        // ---- last CTA: merge block partials in block order, finalise ----
        __threadfence();
        const int G = gridDim.x;
        const volatile double* parts_v = a.ws_partials;
        double Mg = -INFINITY;
        for (int b = tid; b < G; b += kMfThreads) Mg = fmax(Mg, parts_v[(size_t)b * NP]);
        Mg = block_max(Mg, red);
        if (OBJ == VB_OBJ_KLVI) Mg = 0.0;
        if (tid < NP) {
            const double vm = mf_merge_component(const_cast<const double*>(a.ws_partials), G, NP, OBJ, a.alpha, Mg, tid);
            merged[tid] = vm;
            a.out_partials[tid] = vm;
        }
        if (tid == 0) *a.ws_ticket = 0u;   // ready for the next launch on this stream
    }
    __syncthreads();
    if (a.out_value != nullptr && tid < NP)
        mf_finalize_thread(FAM, a.pd ? (int)VB_OBJ_KLVI_PD : OBJ, a.alpha, a.var_param, D, merged, merged[3], a.out_value, a.out_grad, tid);
    if (a.opt != nullptr) {            // every other CTA has retired its reads of var_param (ticket order)
        __syncthreads();
        adagrad_update_block(*a.opt);
    }
    }   // rep
}

// ------------------------------------------------------------------------------------------ dispatch
static int num_sms() { return device_sm_count(); }

struct MfCall {
    int family; double df; int objective; double alpha; int dtype;
    const double* var_param; long long n; const void* base;
    unsigned long long seed; unsigned int call_offset; unsigned long long sample_index0;
    double* out_value; double* out_grad; double* out_partials;
    void* ws; size_t ws_bytes; cudaStream_t stream;
    const MfDynStream* dyn = nullptr;
    const AdagradArgs* opt = nullptr;
    int iters = 1;
    int pd = 0;
};

template <typename T, int FAM, class Model, int OBJ, bool PHILOX>
static int launch_mf(const MfCall& c, const typename Model::P& mp) {
    constexpr int D = Model::D;
    constexpr int NP = 2 * D + 4;
    auto kern = mf_step_kernel<T, FAM, Model, OBJ, PHILOX>;
    const size_t smem = PHILOX ? 0 : (size_t)kMfStages * kMfThreads * D * sizeof(T);
    static DeviceOnce configured;        // per template instantiation, per device
    static std::atomic<int> blocks_per_sm_cached{1};
    const int dev = current_device();
    int blocks_per_sm = blocks_per_sm_cached.load(std::memory_order_relaxed);
    if (configured.pending(dev)) {
        VB_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (smem > 0)
            VB_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout,
                                             (int)cudaSharedmemCarveoutMaxShared));
        VB_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, kern, kMfThreads, smem));
        if (blocks_per_sm < 1) blocks_per_sm = 1;
        blocks_per_sm_cached.store(blocks_per_sm, std::memory_order_relaxed);      // same on every B200 of the box
        configured.done(dev);
    }
    const long long n_tiles = (c.n + kMfThreads - 1) / kMfThreads;
    long long grid = (long long)num_sms() * blocks_per_sm;
    if (grid > n_tiles) grid = n_tiles;
    // Small problems run on ONE CTA (a few tiles in sequence cost less than the cross-CTA ticket / fence / merge
    // round trips).  The rule depends on n only, so every path (single step, optimiser loop) sums in the same order.
    if (c.n <= kMfSingleCtaMax) grid = 1;
    if (grid < 1) grid = 1;
    const size_t need = align_up((size_t)grid * NP * sizeof(double), 16) + 16;
    if (c.ws_bytes < need) return VB_ERR_WORKSPACE;

    MfArgs<Model> a;
    a.mp = mp;
    a.var_param = c.var_param;
    a.base = c.base;
    a.n = c.n;
    a.df = c.df;
    a.alpha = c.alpha;
    a.rng.k0 = (uint32_t)c.seed;
    a.rng.k1 = (uint32_t)(c.seed >> 32);
    a.rng.offset = c.call_offset;
    a.sample_index0 = c.sample_index0;
    // ticket lives at the front so its address does not depend on the grid
    a.ws_ticket = reinterpret_cast<unsigned int*>(c.ws);
    a.ws_partials = reinterpret_cast<double*>(reinterpret_cast<char*>(c.ws) + 16);
    a.out_partials = c.out_partials;
    a.out_value = c.out_value;
    a.out_grad = c.out_grad;
    a.use_bulk = (c.base != nullptr) && ((reinterpret_cast<uintptr_t>(c.base) & 15u) == 0) &&
                 (c.dyn == nullptr || ((size_t)c.n * D * sizeof(T)) % 16 == 0);
    a.pd = c.pd;
    a.dyn = c.dyn;
    a.opt = c.opt;
    a.iters = (PHILOX && c.opt != nullptr && grid == 1 && c.iters > 1) ? c.iters : 1;
    kern<<<(unsigned)grid, kMfThreads, smem, c.stream>>>(a);
    VB_LAUNCH_CHECK();
    return VB_OK;
}

template <typename T, int FAM, class Model>
static int dispatch_obj(const MfCall& c, const typename Model::P& mp) {
    const bool philox = c.base == nullptr;
    if (c.objective == VB_OBJ_KLVI)
        return philox ? launch_mf<T, FAM, Model, VB_OBJ_KLVI, true>(c, mp)
                      : launch_mf<T, FAM, Model, VB_OBJ_KLVI, false>(c, mp);
    if (c.objective == VB_OBJ_CHIVI)
        return philox ? launch_mf<T, FAM, Model, VB_OBJ_CHIVI, true>(c, mp)
                      : launch_mf<T, FAM, Model, VB_OBJ_CHIVI, false>(c, mp);
    if (c.objective == VB_OBJ_KLVI_PD) {       // same kernels as CHIVI: alpha = 0 makes every weight exactly 1
        MfCall c2 = c;
        c2.objective = VB_OBJ_CHIVI;
        c2.alpha = 0.0;
        c2.pd = 1;
        return philox ? launch_mf<T, FAM, Model, VB_OBJ_CHIVI, true>(c2, mp)
                      : launch_mf<T, FAM, Model, VB_OBJ_CHIVI, false>(c2, mp);
    }
    return VB_ERR_ARG;
}

template <class Model>
static int dispatch_family(const MfCall& c, const typename Model::P& mp) {
    if (c.dtype == VB_F64) {
        if (c.family == VB_FAMILY_MF_GAUSS) return dispatch_obj<double, VB_FAMILY_MF_GAUSS, Model>(c, mp);
        if (c.family == VB_FAMILY_MF_T) return dispatch_obj<double, VB_FAMILY_MF_T, Model>(c, mp);
    } else if (c.dtype == VB_F32) {
        if (c.family == VB_FAMILY_MF_GAUSS) return dispatch_obj<float, VB_FAMILY_MF_GAUSS, Model>(c, mp);
        if (c.family == VB_FAMILY_MF_T) return dispatch_obj<float, VB_FAMILY_MF_T, Model>(c, mp);
    }
    return VB_ERR_ARG;
}

// ------------------------------------------------------------------------------------------ samples + log weights
// get_samples_and_log_weights of notebooks/experiments.py:60-63 as ONE pass for the mean-field families and
// the elementwise plugins: draw (or load) eps -> x = m + exp(l) eps -> lw = log p(x) - log q(x; lambda).
// Outputs stay on the device and feed all_bounds / psislw without a host round trip.
template <typename T, int FAM, class Model, bool PHILOX>
__global__ void __launch_bounds__(256)
mf_samples_lw_kernel(const __grid_constant__ MfArgs<Model> a, T* __restrict__ out_x, double* __restrict__ out_logp,
                     double* __restrict__ out_lw) {
    constexpr int D = Model::D;
    __shared__ T mean[D], sig[D];
    if (threadIdx.x < D) {
        mean[threadIdx.x] = (T)a.var_param[threadIdx.x];
        sig[threadIdx.x] = (T)exp(a.var_param[D + threadIdx.x]);
    }
    __syncthreads();
    double sum_l = 0.0;
#pragma unroll
    for (int j = 0; j < D; ++j) sum_l += a.var_param[D + j];
    const double logq0 = D * mf_logq_const(FAM, a.df) - sum_l;
    const T df = (T)a.df;
    const T* base = reinterpret_cast<const T*>(a.base);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < a.n; i += (long long)gridDim.x * blockDim.x) {
        T eps[D];
        if (PHILOX) {
            const uint64_t gi = a.sample_index0 + (uint64_t)i;
            if (FAM == VB_FAMILY_MF_GAUSS) {
#pragma unroll
                for (int q = 0; q < (D + 3) / 4; ++q) {
                    float z[4];
                    draw_normal_pair(a.rng, gi, (uint32_t)q, z[0], z[1], z[2], z[3]);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (4 * q + k < D) eps[4 * q + k] = (T)z[k];
                }
            } else {
                draw_student_t_row<D, T>(a.rng, gi, (float)a.df, eps);
            }
        } else {
#pragma unroll
            for (int j = 0; j < D; ++j) eps[j] = base[(size_t)i * D + j];
        }
        T x[D], g[D], lp;
#pragma unroll
        for (int j = 0; j < D; ++j) x[j] = mean[j] + sig[j] * eps[j];
        Model::template eval<T>(a.mp, x, lp, g);
#pragma unroll
        for (int j = 0; j < D; ++j) out_x[(size_t)i * D + j] = x[j];
        if (out_logp) out_logp[i] = (double)lp;
        out_lw[i] = (double)lp - (logq0 + (double)logq_core<T, FAM, D>(eps, df));
    }
}

template <typename T, int FAM, class Model>
static int launch_samples_lw(const MfCall& c, const typename Model::P& mp, void* out_x, double* out_logp, double* out_lw) {
    MfArgs<Model> a{};
    a.mp = mp;
    a.var_param = c.var_param;
    a.base = c.base;
    a.n = c.n;
    a.df = c.df;
    a.alpha = 1.0;
    a.rng.k0 = (uint32_t)c.seed;
    a.rng.k1 = (uint32_t)(c.seed >> 32);
    a.rng.offset = c.call_offset;
    a.sample_index0 = c.sample_index0;
    long long grid = (c.n + 255) / 256;
    if (grid > (long long)num_sms() * 8) grid = (long long)num_sms() * 8;
    if (c.base == nullptr)
        mf_samples_lw_kernel<T, FAM, Model, true><<<(unsigned)grid, 256, 0, c.stream>>>(a, (T*)out_x, out_logp, out_lw);
    else
        mf_samples_lw_kernel<T, FAM, Model, false><<<(unsigned)grid, 256, 0, c.stream>>>(a, (T*)out_x, out_logp, out_lw);
    VB_LAUNCH_CHECK();
    return VB_OK;
}
template <class Model>
static int dispatch_samples_lw(const MfCall& c, const typename Model::P& mp, void* out_x, double* out_logp, double* out_lw) {
    if (c.dtype == VB_F64) {
        if (c.family == VB_FAMILY_MF_GAUSS) return launch_samples_lw<double, VB_FAMILY_MF_GAUSS, Model>(c, mp, out_x, out_logp, out_lw);
        if (c.family == VB_FAMILY_MF_T) return launch_samples_lw<double, VB_FAMILY_MF_T, Model>(c, mp, out_x, out_logp, out_lw);
    } else if (c.dtype == VB_F32) {
        if (c.family == VB_FAMILY_MF_GAUSS) return launch_samples_lw<float, VB_FAMILY_MF_GAUSS, Model>(c, mp, out_x, out_logp, out_lw);
        if (c.family == VB_FAMILY_MF_T) return launch_samples_lw<float, VB_FAMILY_MF_T, Model>(c, mp, out_x, out_logp, out_lw);
    }
    return VB_ERR_ARG;
}

// ------------------------------------------------------------------------------------------ finalize (multi-rank)
__global__ void mf_finalize_kernel(int family, int objective, double alpha, const double* var_param, int D,
                                   const double* parts, int R, double* out_value, double* out_grad) {
    extern __shared__ double merged_dyn[];
    const int NP = 2 * D + 4;
    double M = 0.0;
    // path-derivative KLVI partials come from the CHIVI pass with alpha = 0: merged like CHIVI (all factors are 1)
    const int merge_obj = (objective == VB_OBJ_KLVI_PD) ? (int)VB_OBJ_CHIVI : objective;
    const double merge_alpha = (objective == VB_OBJ_KLVI_PD) ? 0.0 : alpha;
    if (merge_obj == VB_OBJ_CHIVI) {
        M = -INFINITY;
        for (int r = 0; r < R; ++r) M = fmax(M, parts[(size_t)r * NP]);
    }
    for (int j = threadIdx.x; j < NP; j += blockDim.x)
        merged_dyn[j] = mf_merge_component(parts, R, NP, merge_obj, merge_alpha, M, j);
    __syncthreads();
    for (int j = threadIdx.x; j < NP; j += blockDim.x)
        mf_finalize_thread(family, objective, alpha, var_param, D, merged_dyn, merged_dyn[3], out_value, out_grad, j);
}

// Model dispatch shared by vb_mf_step and the device-resident optimiser.
static int mf_step_call(const MfCall& c, const ModelHost* m) {
    switch (m->kind) {
        case VB_MODEL_FUNNEL: {
            const Funnel::P p = Funnel::make(m->p0);
            return dispatch_family<Funnel>(c, p);
        }
        case VB_MODEL_MIXTURE: {
            Mixture::P p{m->p0};
            return dispatch_family<Mixture>(c, p);
        }
        case VB_MODEL_EIGHT_SCHOOLS: {
            if (m->n_schools != 8) return VB_ERR_UNSUPPORTED;
            EightSchools<8>::P p;
            for (int j = 0; j < 8; ++j) { p.y[j] = m->y[j]; p.inv_sigma[j] = 1.0 / m->sigma[j]; }
            return dispatch_family<EightSchools<8>>(c, p);
        }
        default:
            return VB_ERR_UNSUPPORTED;   // regression / gauss go through vb_model_logp_grad + vb_mf_reduce
    }
}

// ------------------------------------------------------------------------------------------ device-resident AdaGrad
// adagrad_optimize of viabel/vb.py:345-389 with the whole loop on the device: per iteration one fused step
// launch (above) and one update launch; the iterate, the window of recent gradients, the value history and the
// kept iterates never leave HBM, and groups of iterations replay as one CUDA graph.  The update uses explicitly
// rounded fp64 operations (no FMA contraction) in the reference's evaluation order, so for the same gradients
// it is bit-identical to the numpy expression at vb.py:371-374.
__global__ void adagrad_init_kernel(AdagradArgs g) {
    if (threadIdx.x == 0) {
        g.dyn->seed = g.seeds[0];
        g.dyn->offset = g.offsets[0];
        g.dyn->step = 0u;
    }
}

__global__ void __launch_bounds__(128) adagrad_update_kernel(AdagradArgs g) { adagrad_update_block(g); }

}  // namespace vb

struct vb_adagrad_plan {
    vb::MfCall call;
    const vb::ModelHost* model;
    vb::AdagradArgs upd;
    int graph_iters;
    int fused;
    cudaGraph_t graph;
    cudaGraphExec_t exec;
};

namespace vb {

static int adagrad_enqueue_one(const vb_adagrad_plan* p, cudaStream_t stream) {
    MfCall c = p->call;
    c.stream = stream;
    const int rc = mf_step_call(c, p->model);
    if (rc != VB_OK || p->fused) return rc;
    adagrad_update_kernel<<<1, 128, 0, stream>>>(p->upd);
    VB_LAUNCH_CHECK();
    return VB_OK;
}
// fused single-CTA Philox problems: `iters` iterations inside ONE launch
static bool adagrad_persistent(const vb_adagrad_plan* p) {
    return p->fused && p->call.base == nullptr && p->call.n <= kMfSingleCtaMax;
}
static int adagrad_enqueue_many(const vb_adagrad_plan* p, int iters, cudaStream_t stream) {
    MfCall c = p->call;
    c.stream = stream;
    c.iters = iters;
    return mf_step_call(c, p->model);
}

}  // namespace vb

using namespace vb;

extern "C" size_t vb_mf_step_workspace_bytes(int dim) {
    if (dim < 1) return 0;
    // up to 8 resident CTAs per SM on any part this runs on
    return 16 + align_up((size_t)num_sms() * 8 * VB_MF_NPART(dim) * sizeof(double), 16) + 16;
}

extern "C" int vb_mf_step(int family, double df, const vb_model* model, int objective, double alpha,
                          int dtype, const double* d_var_param, long long n,
                          const void* d_base_or_null,
                          unsigned long long seed, unsigned int call_offset,
                          unsigned long long sample_index0,
                          double* d_out_value_or_null, double* d_out_grad_or_null,
                          double* d_out_partials, void* d_ws, size_t ws_bytes, void* stream) {
    if (!model || !d_var_param || !d_out_partials || !d_ws || n < 1) return VB_ERR_ARG;
    if ((d_out_value_or_null == nullptr) != (d_out_grad_or_null == nullptr)) return VB_ERR_ARG;
    if (family == VB_FAMILY_MF_T && !(df > 2.0)) return VB_ERR_ARG;
    if (objective == VB_OBJ_CHIVI && !(alpha > 0.0)) return VB_ERR_ARG;
    const ModelHost* m = reinterpret_cast<const ModelHost*>(model);
    MfCall c{family, df, objective, alpha, dtype, d_var_param, n, d_base_or_null, seed, call_offset,
             sample_index0, d_out_value_or_null, d_out_grad_or_null, d_out_partials, d_ws, ws_bytes,
             (cudaStream_t)stream};
    return mf_step_call(c, m);
}

extern "C" int vb_mf_finalize(int family, int objective, double alpha, const double* d_var_param,
                              int dim, const double* d_partials, int n_ranks,
                              double* d_out_value, double* d_out_grad, void* stream) {
    if (!d_var_param || !d_partials || !d_out_value || !d_out_grad || dim < 1 || n_ranks < 1) return VB_ERR_ARG;
    const int NP = VB_MF_NPART(dim);
    mf_finalize_kernel<<<1, 256, NP * sizeof(double), (cudaStream_t)stream>>>(
        family, objective, alpha, d_var_param, dim, d_partials, n_ranks, d_out_value, d_out_grad);
    VB_LAUNCH_CHECK();
    return VB_OK;
}


/* ---- device-resident AdaGrad loop (viabel/vb.py:345-389) ---- */
static size_t adagrad_state_layout(int dim, int window, size_t* off_out, size_t* off_part, size_t* off_ring) {
    const size_t P = 2 * (size_t)dim;
    size_t o = 16;                                             // MfDynStream
    o += align_up(sizeof(AdagradArgs), 16);                    // device copy of the update arguments (fused mode)
    *off_out = o;  o += align_up((1 + P) * sizeof(double), 16);
    *off_part = o; o += align_up((size_t)VB_MF_NPART(dim) * sizeof(double), 16);
    *off_ring = o; o += align_up((size_t)window * P * sizeof(double), 16);
    return o;
}

extern "C" size_t vb_mf_adagrad_state_bytes(int dim, int window) {
    if (dim < 1 || window < 1) return 0;
    size_t a, b, c;
    return adagrad_state_layout(dim, window, &a, &b, &c);
}

static int optimizer_plan_create(int rule, const double* d_bias1, const double* d_bias2,
                                 int family, double df, const vb_model* model, int objective, double alpha,
                                 int dtype, double* d_var_param, long long n_samples,
                                 const void* d_base_or_null, int n_iters, int window, double epsilon,
                                 const double* d_rates, const unsigned long long* d_seeds,
                                 const unsigned int* d_offsets, double* d_value_history,
                                 double* d_param_history, void* d_state, size_t state_bytes,
                                 void* d_ws, size_t ws_bytes, int graph_iters, int fused, void* stream,
                                 vb_adagrad_plan** out_plan) {
    if (!model || !d_var_param || !d_rates || !d_seeds || !d_offsets || !d_value_history || !d_param_history ||
        !d_state || !d_ws || !out_plan || n_samples < 1 || n_iters < 1 || window < 1 || graph_iters < 0)
        return VB_ERR_ARG;
    if (family == VB_FAMILY_MF_T && !(df > 2.0)) return VB_ERR_ARG;
    if (objective == VB_OBJ_CHIVI && !(alpha > 0.0)) return VB_ERR_ARG;
    const ModelHost* m = reinterpret_cast<const ModelHost*>(model);
    const int dim = m->dim;
    size_t off_out, off_part, off_ring;
    if (state_bytes < adagrad_state_layout(dim, window, &off_out, &off_part, &off_ring)) return VB_ERR_WORKSPACE;
    if ((reinterpret_cast<uintptr_t>(d_state) & 15u) != 0) return VB_ERR_ALIGN;
    char* st = reinterpret_cast<char*>(d_state);
    cudaStream_t s = (cudaStream_t)stream;

    vb_adagrad_plan* p = new (std::nothrow) vb_adagrad_plan();
    if (!p) return VB_ERR_ARG;
    double* out = reinterpret_cast<double*>(st + off_out);
    MfDynStream* dyn = reinterpret_cast<MfDynStream*>(st);
    p->call = MfCall{family, df, objective, alpha, dtype, d_var_param, n_samples, d_base_or_null, 0ull, 0u, 0ull,
                     out, out + 1, reinterpret_cast<double*>(st + off_part), d_ws, ws_bytes, s, dyn};
    p->model = m;
    p->upd = AdagradArgs{2 * dim, window, n_iters, 3 * n_iters / 4, epsilon, d_var_param, out,
                         reinterpret_cast<double*>(st + off_ring), d_rates, d_seeds, d_offsets, d_value_history,
                         d_param_history, dyn};
    if (rule != VB_OPT_ADAGRAD) {
        p->upd.rule = rule;
        p->upd.keep_from = 0;
        p->upd.bias1 = d_bias1;
        p->upd.bias2 = d_bias2;
        // the constants as Python evaluates them (1. - 0.9 is not 0.1)
        if (rule == VB_OPT_RMSPROP) { p->upd.c0 = 0.9; p->upd.c1 = 1.0 - 0.9; }
        else { p->upd.c0 = 0.9; p->upd.c1 = 1.0 - 0.9; p->upd.c2 = 0.999; p->upd.c3 = 1.0 - 0.999; }
    }
    p->graph_iters = graph_iters;
    p->fused = fused != 0;
    p->graph = nullptr;
    p->exec = nullptr;

    adagrad_init_kernel<<<1, 32, 0, s>>>(p->upd);
    int rc = (int)cudaGetLastError();
    AdagradArgs* d_upd = reinterpret_cast<AdagradArgs*>(st + 16);
    if (rc == VB_OK) rc = (int)cudaMemcpyAsync(d_upd, &p->upd, sizeof(AdagradArgs), cudaMemcpyHostToDevice, s);
    // one step outside any capture: configures the kernel (attributes, occupancy) and validates the arguments; it
    // only writes the (value, grad) scratch, which every iteration overwrites before reading
    if (rc == VB_OK) rc = mf_step_call(p->call, m);
    if (p->fused) p->call.opt = d_upd;
    if (rc == VB_OK && graph_iters > 0 && !adagrad_persistent(p)) {
        cudaStream_t cs = nullptr;
        rc = (int)cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking);
        if (rc == VB_OK) {
            rc = (int)cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
            if (rc == VB_OK) {
                int rc_body = VB_OK;
                for (int k = 0; k < graph_iters && rc_body == VB_OK; ++k) rc_body = adagrad_enqueue_one(p, cs);
                rc = (int)cudaStreamEndCapture(cs, &p->graph);
                if (rc_body != VB_OK) rc = rc_body;
            }
            if (rc == VB_OK) rc = (int)cudaGraphInstantiate(&p->exec, p->graph, 0ull);
            cudaStreamDestroy(cs);
        }
    }
    if (rc != VB_OK) {
        if (p->exec) cudaGraphExecDestroy(p->exec);
        if (p->graph) cudaGraphDestroy(p->graph);
        delete p;
        return rc;
    }
    *out_plan = p;
    return VB_OK;
}

extern "C" int vb_mf_adagrad_create(int family, double df, const vb_model* model, int objective, double alpha,
                                    int dtype, double* d_var_param, long long n_samples,
                                    const void* d_base_or_null, int n_iters, int window, double epsilon,
                                    const double* d_rates, const unsigned long long* d_seeds,
                                    const unsigned int* d_offsets, double* d_value_history,
                                    double* d_param_history, void* d_state, size_t state_bytes,
                                    void* d_ws, size_t ws_bytes, int graph_iters, int fused, void* stream,
                                    vb_adagrad_plan** out_plan) {
    return optimizer_plan_create(VB_OPT_ADAGRAD, nullptr, nullptr, family, df, model, objective, alpha, dtype,
                                 d_var_param, n_samples, d_base_or_null, n_iters, window, epsilon, d_rates, d_seeds,
                                 d_offsets, d_value_history, d_param_history, d_state, state_bytes, d_ws, ws_bytes,
                                 graph_iters, fused, stream, out_plan);
}

extern "C" int vb_mf_optimizer_create(int rule, int family, double df, const vb_model* model, int objective,
                                      double alpha, int dtype, double* d_var_param, long long n_samples,
                                      const void* d_base_or_null, int n_iters, double epsilon,
                                      const double* d_rates, const unsigned long long* d_seeds,
                                      const unsigned int* d_offsets, const double* d_bias1_or_null,
                                      const double* d_bias2_or_null, double* d_value_history,
                                      double* d_param_history, void* d_state, size_t state_bytes,
                                      void* d_ws, size_t ws_bytes, int graph_iters, int fused, void* stream,
                                      vb_adagrad_plan** out_plan) {
    if (rule != VB_OPT_RMSPROP && rule != VB_OPT_ADAM) return VB_ERR_ARG;
    if (rule == VB_OPT_ADAM && (!d_bias1_or_null || !d_bias2_or_null)) return VB_ERR_ARG;
    // the two moment vectors live where AdaGrad keeps its gradient window: a "window" of 2 rows
    return optimizer_plan_create(rule, d_bias1_or_null, d_bias2_or_null, family, df, model, objective, alpha, dtype,
                                 d_var_param, n_samples, d_base_or_null, n_iters, 2, epsilon, d_rates, d_seeds,
                                 d_offsets, d_value_history, d_param_history, d_state, state_bytes, d_ws, ws_bytes,
                                 graph_iters, fused, stream, out_plan);
}

extern "C" int vb_mf_adagrad_run(vb_adagrad_plan* plan, int n_steps, void* stream) {
    if (!plan || n_steps < 0) return VB_ERR_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    int left = n_steps;
    while (adagrad_persistent(plan) && left > 0) {          // no graph needed: the loop lives inside the kernel
        const int k = left < 512 ? left : 512;
        const int rc = adagrad_enqueue_many(plan, k, s);
        if (rc != VB_OK) return rc;
        left -= k;
    }
    while (plan->exec != nullptr && left >= plan->graph_iters) {
        VB_CUDA_TRY(cudaGraphLaunch(plan->exec, s));
        left -= plan->graph_iters;
    }
    for (; left > 0; --left) {
        const int rc = adagrad_enqueue_one(plan, s);
        if (rc != VB_OK) return rc;
    }
    return VB_OK;
}

extern "C" int vb_mf_adagrad_destroy(vb_adagrad_plan* plan) {
    if (!plan) return VB_OK;
    if (plan->exec) cudaGraphExecDestroy(plan->exec);
    if (plan->graph) cudaGraphDestroy(plan->graph);
    delete plan;
    return VB_OK;
}

extern "C" int vb_mf_samples_logweights(int family, double df, const vb_model* model, int dtype,
                                        const double* d_var_param, long long n, const void* d_base_or_null,
                                        unsigned long long seed, unsigned int call_offset,
                                        unsigned long long sample_index0, void* d_out_samples,
                                        double* d_out_logp_or_null, double* d_out_lw, void* stream) {
    if (!model || !d_var_param || !d_out_samples || !d_out_lw || n < 1) return VB_ERR_ARG;
    if (family == VB_FAMILY_MF_T && !(df > 2.0)) return VB_ERR_ARG;
    const ModelHost* m = reinterpret_cast<const ModelHost*>(model);
    MfCall c{family, df, VB_OBJ_KLVI, 1.0, dtype, d_var_param, n, d_base_or_null, seed, call_offset,
             sample_index0, nullptr, nullptr, nullptr, nullptr, 0, (cudaStream_t)stream};
    switch (m->kind) {
        case VB_MODEL_FUNNEL: {
            const Funnel::P p = Funnel::make(m->p0);
            return dispatch_samples_lw<Funnel>(c, p, d_out_samples, d_out_logp_or_null, d_out_lw);
        }
        case VB_MODEL_MIXTURE: {
            Mixture::P p{m->p0};
            return dispatch_samples_lw<Mixture>(c, p, d_out_samples, d_out_logp_or_null, d_out_lw);
        }
        case VB_MODEL_EIGHT_SCHOOLS: {
            if (m->n_schools != 8) return VB_ERR_UNSUPPORTED;
            EightSchools<8>::P p;
            for (int j = 0; j < 8; ++j) { p.y[j] = m->y[j]; p.inv_sigma[j] = 1.0 / m->sigma[j]; }
            return dispatch_samples_lw<EightSchools<8>>(c, p, d_out_samples, d_out_logp_or_null, d_out_lw);
        }
        default:
            return VB_ERR_UNSUPPORTED;   // regression / gauss / full rank: composed from sample + plugin + logdensity
    }
}
