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
