"""Seeded inputs shared by scripts/make_golden.py (which runs the reference on them) and the
parity tests (which run the oracle and the CUDA path on them)."""
import numpy as np
from scipy.stats import norm

BOUNDS_CASES = ['mixture_nb', 'mixture_nb_known', 'gauss3d', 'offset1d']
PSIS_CASES = ['gpd05_n10', 'gpd05_n100', 'gpd05_n5000', 'gpd07_n20000', 'gauss_n3000', 'gpd05_n200000',
              'offset_n1000', 'gpd05_reff_n4000']
GPD_CASES = ['gpd05_2000', 'gpd02_500', 'exp_50']
LOO_CASES = ['regression_n2000_m9', 'mixed_n300_m5', 'tiny_n12_m3']
GPINV_PARAMS = [(0.5, 1.0), (-0.3, 2.0), (0.0, 1.5), (0.7, -1.0)]


def gpd_log_weights(n, k, seed, offset=0.0):
    """log of GPD(k, sigma=1) variates: heavy right tail with tail index k."""
    u = np.random.RandomState(seed).rand(n)
    return np.log(np.expm1(-k * np.log1p(-u)) / k) + offset


def bounds_case(name):
    """-> (log_weights, samples or None, kwargs for all_bounds)."""
    if name.startswith('mixture_nb'):
        # notebooks/normal-mixture.ipynb:40-43,96,113-116
        q_var = 5
        q = norm(scale=np.sqrt(q_var))
        np.random.seed(918)
        samples = q.rvs(50000)
        log_p = np.logaddexp(norm.logpdf(samples, loc=-2), norm.logpdf(samples, loc=2)) - np.log(2)
        lw = log_p - q.logpdf(samples)
        if name == 'mixture_nb':
            return lw, samples, {}
        return lw, samples, dict(q_var=q_var, log_norm_bound=0)
    if name == 'gauss3d':
        rs = np.random.RandomState(17)
        samples = rs.randn(30000, 3) * np.array([1.0, 2.0, 0.5]) + np.array([0.3, -1.0, 2.0])
        lw = -0.5 * np.sum((samples / np.array([1.2, 1.5, 0.7])) ** 2, axis=1) + 0.3 * rs.randn(30000)
        return lw, samples, {}
    if name == 'offset1d':
        rs = np.random.RandomState(23)
        samples = rs.standard_t(10, 20001)
        lw = -4000.0 + 0.4 * rs.randn(20001)
        return lw, samples, dict(p_var=2.0)
    raise KeyError(name)


def psis_case(name):
    if name == 'gpd05_n10':
        return gpd_log_weights(10, 0.5, 1)
    if name == 'gpd05_n100':
        return gpd_log_weights(100, 0.5, 2)
    if name == 'gpd05_n5000':
        return gpd_log_weights(5000, 0.5, 3)
    if name == 'gpd07_n20000':
        return gpd_log_weights(20000, 0.7, 4)
    if name == 'gauss_n3000':
        return np.random.RandomState(5).randn(3000) * 0.5          # light tail: k < 1/3, no smoothing
    if name == 'gpd05_n200000':
        return gpd_log_weights(200000, 0.5, 6)
    if name == 'offset_n1000':
        return gpd_log_weights(1000, 0.6, 7, offset=-12345.0)
    if name == 'gpd05_reff_n4000':
        return gpd_log_weights(4000, 0.5, 8)
    raise KeyError(name)


def probe_indices(n):
    return np.unique(np.linspace(0, n - 1, 257).astype(np.int64))


def gpd_case(name):
    rs = np.random.RandomState({'gpd05_2000': 11, 'gpd02_500': 12, 'exp_50': 13}[name])
    if name == 'gpd05_2000':
        u = rs.rand(2000)
        return np.sort(np.expm1(-0.5 * np.log1p(-u)) / 0.5)
    if name == 'gpd02_500':
        u = rs.rand(500)
        return np.sort(np.expm1(-0.2 * np.log1p(-u)) / 0.2) * 3.0
    return np.sort(rs.exponential(size=50))


def gpinv_probs():
    return np.concatenate([[0.0], (np.arange(0.5, 64) / 64), [1.0]])


def loo_case(name):
    """(n, m) log-likelihood draws for psisloo: n posterior draws of m pointwise terms.  Columns mix heavy tails
    (an outlying observation: -log_lik has a GPD-like right tail), light tails (no smoothing) and ties."""
    if name == 'regression_n2000_m9':
        rs = np.random.RandomState(21)
        n, m = 2000, 9
        beta = 1.0 + 0.3 * rs.randn(n)                                  # posterior draws of a slope
        sd = np.exp(0.1 * rs.randn(n))
        xs = np.linspace(-2.0, 2.0, m)
        ys = xs + np.array([0.1, -0.2, 0.05, 0.0, 3.5, -0.1, 0.2, -4.0, 0.3])     # two outliers
        r = (ys[None, :] - beta[:, None] * xs[None, :]) / sd[:, None]
        return -0.5 * r * r - np.log(sd)[:, None] - 0.9189385332046727
    if name == 'mixed_n300_m5':
        rs = np.random.RandomState(22)
        ll = -gpd_log_weights(300 * 5, 0.6, 23).reshape(300, 5)
        ll[:, 1] = 0.2 * rs.randn(300)                                  # light tail
        ll[:, 3] = np.round(ll[:, 3], 1)                                # ties
        return ll
    if name == 'tiny_n12_m3':
        return np.random.RandomState(24).randn(12, 3)
    raise KeyError(name)


# ------------------------------------------------------------------------------------------ R-hat / optimisers (N4)
RHAT_CASES = ['c4_t400_p6', 'c1_t64_p3', 'c3_t1000_p40_warm', 'const_param']
RHAT_WINDOW_CASES = ['c2_t600_p8', 'c3_t1200_p10']
AVERAGING_CASES = ['n50_k4', 'n400_k7_start100']
OPTIM_CASES = ['adagrad_const', 'adagrad_decay_lognorm', 'rmsprop_1opt', 'rmsprop_2opt_avgnorm', 'adam_1opt',
               'adam_2opt_lognorm']


def _ar_chains(n_chains, n_iters, p, seed, drift=0.0):
    """Autocorrelated chains around chain-specific levels (so R-hat is neither 1 nor huge)."""
    rs = np.random.RandomState(seed)
    x = np.zeros((n_chains, n_iters, p))
    level = 0.3 * rs.randn(n_chains, p)
    e = rs.randn(n_chains, n_iters, p)
    for t in range(1, n_iters):
        x[:, t] = 0.8 * x[:, t - 1] + e[:, t]
    return x + level[:, None, :] + drift * np.linspace(1.0, 0.0, n_iters)[None, :, None] ** 4


def rhat_case(name):
    """(chains (n_chains, n_iters, p), kwargs of compute_R_hat)."""
    if name == 'c4_t400_p6':
        return _ar_chains(4, 400, 6, 31), {'warmup': 0}
    if name == 'c1_t64_p3':
        return _ar_chains(1, 64, 3, 32), {'warmup': 0}
    if name == 'c3_t1000_p40_warm':
        return _ar_chains(3, 1000, 40, 33, drift=5.0), {}           # default warmup = 500
    if name == 'const_param':
        x = _ar_chains(2, 100, 4, 34)
        x[:, :, 2] = 1.5                                              # zero variance: only the jitter keeps W > 0
        return x, {'warmup': 20}
    raise KeyError(name)


def rhat_window_case(name):
    """(chains, window_size, interval, start) for compute_R_hat_adaptive_numpy / compute_R_hat_halfway."""
    if name == 'c2_t600_p8':
        return _ar_chains(2, 600, 8, 35, drift=3.0), 100, 100, 200
    if name == 'c3_t1200_p10':
        return _ar_chains(3, 1200, 10, 36, drift=8.0), 200, 100, 1000     # start + k*interval runs past the end
    raise KeyError(name)


def averaging_case(name):
    if name == 'n50_k4':
        return np.random.RandomState(37).randn(50, 4), 0
    if name == 'n400_k7_start100':
        return np.cumsum(np.random.RandomState(38).randn(400, 7), axis=0) * 0.05, 100
    raise KeyError(name)


class NoisyQuadratic:
    """Deterministic stand-in for a Monte Carlo objective: value 1/2 sum a (x - c)^2 + noise, gradient
    a (x - c) + noise, drawn from the objective's OWN RandomState so that the reference optimiser and the one
    under test see the same sequence (call k gets draw k).  With has_log_norm the third return value is a
    per-call log normaliser, as CHIVI-style objectives give."""

    def __init__(self, dim, seed, has_log_norm=False):
        rs = np.random.RandomState(seed)
        self.a = np.exp(0.5 * rs.randn(dim))
        self.c = rs.randn(dim)
        self.rs = np.random.RandomState(seed + 1000)
        self.has_log_norm = has_log_norm
        self.calls = 0

    def __call__(self, x):
        x = np.asarray(x, dtype=np.float64)
        self.calls += 1
        r = x - self.c
        value = 0.5 * np.sum(self.a * r * r) + 0.05 * self.rs.randn()
        grad = self.a * r + 0.1 * self.rs.randn(len(x))
        if self.has_log_norm:
            return value, grad, -0.3 * value + 0.2 * self.rs.randn()
        return value, grad


def optim_case(name):
    """(kind, n_iters, init_param, kwargs, objective seed)."""
    init6 = np.array([1.5, -0.5, 0.25, 0.0, -1.0, 0.5])
    if name == 'adagrad_const':
        return 'adagrad', 60, init6, {}, 41
    if name == 'adagrad_decay_lognorm':
        return 'adagrad', 80, init6, {'has_log_norm': True, 'window': 5, 'learning_rate': .05, 'epsilon': .2,
                                      'learning_rate_end': .005}, 42
    if name == 'rmsprop_1opt':
        return 'rmsprop', 400, init6, {'rhat_window': 100, 'tail_avg_iters': 150, 'learning_rate': .02}, 43
    if name == 'rmsprop_2opt_avgnorm':
        return 'rmsprop', 400, init6, {'rhat_window': 50, 'n_optimisers': 2, 'avg_grad_norm': True, 'window': 5,
                                       'tail_avg_iters': 100, 'learning_rate': .05, 'learning_rate_end': .01,
                                       'r_mean_threshold': 1.6, 'r_sigma_threshold': 1.7}, 44
    if name == 'adam_1opt':
        return 'adam', 400, init6, {'rhat_window': 100, 'tail_avg_iters': 150, 'learning_rate': .02}, 45
    if name == 'adam_2opt_lognorm':
        return 'adam', 400, init6, {'rhat_window': 50, 'n_optimisers': 2, 'has_log_norm': True, 'window': 4,
                                    'tail_avg_iters': 120, 'learning_rate': .05, 'learning_rate_end': .01,
                                    'r_mean_threshold': 1.6, 'r_sigma_threshold': 1.7}, 46
    raise KeyError(name)


# ------------------------------------------------------------------------------------------ vb.py objectives
# Log densities handed to the REFERENCE's black_box_klvi / black_box_chivi by scripts/make_golden_vb.py: plain numpy
# (so they also accept the complex arguments of the complex-step differentiation stand-in); the tests check that
# they equal the oracle's model plugins at real arguments.
def funnel_logp(x, a=1.35):
    mu, ls = x[:, 0], x[:, 1]
    return (-0.5 * np.log(2 * np.pi) - np.log(a) - 0.5 * (ls / a) ** 2) \
        + (-0.5 * np.log(2 * np.pi) - ls - 0.5 * mu * mu * np.exp(-2 * ls))


SCHOOLS_Y = np.array([28., 8., -3., 7., -1., 1., 18., 12.])
SCHOOLS_SIGMA = np.array([15., 10., 16., 11., 9., 11., 10., 18.])


def eight_schools_logp(u):
    mu, lt, th = u[:, 0], u[:, 1], u[:, 2:]
    tau = np.exp(lt)
    r = (SCHOOLS_Y[None, :] - mu[:, None] - tau[:, None] * th) / SCHOOLS_SIGMA[None, :]
    return -mu * mu / 50 - np.log1p(tau * tau / 25) + lt - 0.5 * np.sum(th * th, axis=1) - 0.5 * np.sum(r * r, axis=1)


VB_CASES = ['gauss_funnel', 't_schools']


def vb_case(name):
    """(family kind, dim, df, log density name, var_param, n_samples)."""
    if name == 'gauss_funnel':
        return 'mf_gauss', 2, None, 'funnel', np.array([0.1, -0.2, 0.3, -0.1]), 64
    if name == 't_schools':
        rs = np.random.RandomState(51)
        lam = np.concatenate([[4.0, 0.8], 0.2 * rs.randn(8), np.log([3.2, 0.73]), -0.05 + 0.02 * rs.randn(8)])
        return 'mf_t', 10, 7.0, 'eight_schools', lam, 48
    raise KeyError(name)


# full-rank Student-t family (vb.py:192-233): gradients of these goldens come from a 4-point central difference of
# the reference's own objective (scripts/ref_stand_ins.py, mode 'fd'), not from the complex step
VB_FR_CASES = ['fr_t_gauss3']
_FR_MEAN = np.array([0.5, -1.0, 0.25])
_FR_PREC = np.array([[2.0, 0.6, -0.3], [0.6, 1.5, 0.2], [-0.3, 0.2, 1.0]])


def gauss3_logp(x):
    r = x - _FR_MEAN
    return -0.5 * np.sum((r @ _FR_PREC) * r, axis=1)


def vb_fr_case(name):
    """(dim, df, log density name, var_param [mu, free Sigma], n_samples)."""
    if name == 'fr_t_gauss3':
        rs = np.random.RandomState(61)
        return 3, 9.0, 'gauss3', np.concatenate([0.3 * rs.randn(3), 0.25 * rs.randn(6)]), 40
    raise KeyError(name)


# ------------------------------------------------------------------------------------------ notebooks/experiments.py
EXPERIMENT_CASES = ['gauss_funnel', 't_schools']
EXPERIMENT_N = 3000


def experiment_truth(name):
    """(true_mean, true_cov) handed to improve_with_psis / check_accuracy (any fixed SPD matrix will do)."""
    kind, dim, df, logp_name, lam, n = vb_case(name)
    rs = np.random.RandomState(71)
    a = rs.randn(dim, dim)
    return 0.1 * rs.randn(dim), a @ a.T / dim + np.eye(dim)


# config C1 (funnel, MF-Gaussian, AdaGrad .01 -> .001) end to end at a size the reference finishes in seconds
C1_GOLDEN_N, C1_GOLDEN_ITERS = 100, 240
