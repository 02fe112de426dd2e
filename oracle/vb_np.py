"""Closed-form numpy restatement of ``viabel/vb.py`` (oracle + timed CPU baseline).

Test infrastructure only.  The reference differentiates with ``autograd``; here every
gradient is written out analytically and is cross-checked against a
literal torch-autograd transcription in ``oracle/vb_autograd.py``.  PINNED on golden vectors
made by running the reference's own ``viabel/vb.py`` in the build container with stand-ins for
the absent ``autograd`` / ``paragami`` (``tests/golden/vb_reference.npz``,
``tests/golden/optimizers.npz``; what exactly is pinned, and the one thing that is not --
paragami's flattening layout -- is stated in ``oracle/__init__.py``).
Citations are ``/root/reference/viabel/vb.py:<line>``.
"""
from collections import namedtuple

import numpy as np
from scipy.special import gammaln

__all__ = ['VariationalFamily', 'mean_field_gaussian_variational_family',
           'mean_field_t_variational_family', 't_variational_family',
           'black_box_klvi', 'black_box_klvi_pd', 'black_box_klvi_pd2', 'black_box_chivi', 'adagrad_optimize',
           'learning_rate_schedule', 'klvi_value_and_grad', 'chivi_value_and_grad']

_Fields = namedtuple('VariationalFamily',                      # vb.py:42-45
                     ['sample', 'entropy', 'logdensity', 'mean_and_cov',
                      'pth_moment', 'var_param_dim'])


class VariationalFamily(_Fields):
    """The reference's namedtuple plus an ``impl`` attribute the closed-form objectives use."""
    impl = None


_LOG_2PI = np.log(2 * np.pi)


# --------------------------------------------------------------------------- mean field
class _MeanField:
    """lambda = [m (d), l (d)], x = m + exp(l) * eps    (vb.py:50-57, 144-151)."""

    def __init__(self, dim, df=None):
        self.dim = dim
        self.df = df                    # None -> Gaussian
        self.kind = 'mf_gauss' if df is None else 'mf_t'
        self.rs = np.random.RandomState(0)             # vb.py:49,143

    def draw_base(self, rs, n):
        if self.df is None:
            return {'eps': rs.randn(n, self.dim)}                       # vb.py:57
        return {'eps': rs.standard_t(self.df, size=(n, self.dim))}      # vb.py:151

    def transform(self, lam, base):
        m, l = lam[:self.dim], lam[self.dim:]
        return m + np.exp(l) * base['eps']

    def entropy(self, lam):
        l = lam[self.dim:]
        if self.df is None:
            return 0.5 * self.dim * (1.0 + _LOG_2PI) + np.sum(l)        # vb.py:61
        return np.sum(l)                                                # vb.py:156

    def logq_const(self):
        """Per-coordinate normalising constant of the standardised base density."""
        if self.df is None:
            return -0.5 * _LOG_2PI
        df = self.df
        return gammaln(0.5 * (df + 1)) - gammaln(0.5 * df) - 0.5 * np.log(df * np.pi)

    def logq_of_base(self, lam, base):
        """log q(x(lam); lam) expressed through the base draw (y == eps)."""
        eps = base['eps']
        l = lam[self.dim:]
        if self.df is None:
            core = -0.5 * np.sum(eps * eps, axis=1)
        else:
            core = -0.5 * (self.df + 1) * np.sum(np.log1p(eps * eps / self.df), axis=1)
        return self.dim * self.logq_const() - np.sum(l) + core

    def logdensity(self, x, lam):
        """vb.py:63-65 (Gaussian, diagonal mvn.logpdf) / vb.py:158-162 (product of t)."""
        x = np.asarray(x, dtype=float)
        squeeze = x.ndim == 1 and self.df is None
        if x.ndim == 1:
            x = x[None, :]
        m, l = lam[:self.dim], lam[self.dim:]
        yy = (x - m) * np.exp(-l)
        out = self.logq_of_base(lam, {'eps': yy})
        return out[0] if squeeze else out

    def mean_and_cov(self, lam):
        m, l = lam[:self.dim], lam[self.dim:]
        c = 1.0 if self.df is None else self.df / (self.df - 2)         # vb.py:69,166
        return m, c * np.diag(np.exp(2 * l))

    def pth_moment(self, p, lam):
        if p not in [2, 4]:                                             # vb.py:72-73
            raise ValueError('only p = 2 or 4 supported')
        s2 = np.exp(2 * lam[self.dim:])
        if self.df is None:                                             # vb.py:76-79
            return np.sum(s2) if p == 2 else 2 * np.sum(s2 ** 2) + np.sum(s2) ** 2
        df = self.df
        if df <= p:                                                     # vb.py:171-172
            raise ValueError('df must be greater than p')
        c = df / (df - 2)
        if p == 2:
            return c * np.sum(s2)
        return c ** 2 * (2 * (df - 1) / (df - 4) * np.sum(s2 ** 2) + np.sum(s2) ** 2)

    # gradient of  sum_i w_i log p(x_i)  and of  -sum_i w_i log q(x_i(lam); lam)
    def weighted_grad(self, lam, base, w, g):
        eps = base['eps']
        sig = np.exp(lam[self.dim:])
        gm = w @ g
        gl = (w @ (g * eps)) * sig
        return np.concatenate([gm, gl])

    def neg_logq_grad(self, lam, wsum):
        out = np.zeros(2 * self.dim)
        out[self.dim:] = wsum          # d(-log q)/dl_j = +1 per sample
        return out


# --------------------------------------------------------------------------- full rank t
def tril_rowmajor_indices(d):
    return np.tril_indices(d)


class _FullRankT:
    """paragami ``PatternDict(free_default=True)`` with ``mu`` then ``Sigma`` (vb.py:185-189):
    flat = [mu (d), vec_tril_rowmajor(chol(Sigma) with log on the diagonal)]."""

    def __init__(self, dim, df, scale='sqrtm'):
        # scale='chol' is NOT in the reference: x = mu + L ztilde (same distribution, different
        # per-draw map); it is the oracle of the product's documented throughput mode.
        self.dim, self.df, self.kind, self.scale = dim, df, 'fr_t', scale
        self.rs = np.random.RandomState(0)                              # vb.py:195
        self.tri = np.tril_indices(dim)
        self.diag_pos = np.cumsum(np.arange(1, dim + 1)) - 1            # flat slots of L_ii

    @property
    def n_param(self):
        return self.dim + self.dim * (self.dim + 1) // 2

    def fold(self, lam):
        d = self.dim
        mu = lam[:d]
        L = np.zeros((d, d))
        L[self.tri] = lam[d:]
        L[np.diag_indices(d)] = np.exp(np.diag(L))
        return mu, L, L @ L.T

    def sqrt_sigma(self, Sigma):
        lamb, V = np.linalg.eigh(Sigma)
        return (V * np.sqrt(lamb)) @ V.T, lamb, V

    def draw_base(self, rs, n):
        c = rs.chisquare(self.df, n)                                    # vb.py:204 (first)
        z = rs.randn(n, self.dim)                                       # vb.py:206
        return {'chi2': c, 'z': z}

    def ztilde(self, base):
        return base['z'] / np.sqrt(base['chi2'] / self.df)[:, None]

    def transform(self, lam, base):
        mu, L, Sigma = self.fold(lam)
        if self.scale == 'chol':
            return mu + self.ztilde(base) @ L.T
        S, _, _ = self.sqrt_sigma(Sigma)
        return mu + self.ztilde(base) @ S                               # vb.py:207-208

    def entropy(self, lam):
        return np.sum(lam[self.dim:][self.diag_pos])                    # vb.py:213 = sum log L_ii

    def logq_const(self):
        d, df = self.dim, self.df
        return gammaln(0.5 * (df + d)) - gammaln(0.5 * df) - 0.5 * d * np.log(np.pi * df)

    def logq_of_base(self, lam, base):
        zt = self.ztilde(base)
        maha = np.sum(zt * zt, axis=1)
        return self.logq_const() - self.entropy(lam) \
            - 0.5 * (self.df + self.dim) * np.log1p(maha / self.df)

    def logdensity(self, x, lam):
        """_distributions.py:8-38 via vb.py:198-200 (eigh pseudo-inverse, eps 1e-10)."""
        x = np.atleast_2d(np.asarray(x, dtype=float))
        mu, L, Sigma = self.fold(lam)
        s, u = np.linalg.eigh(Sigma)
        s_pinv = np.where(np.abs(s) <= 1e-10, 0.0, 1.0 / np.where(s == 0, 1.0, s))
        U = u * np.sqrt(s_pinv)
        maha = np.sum(np.square((x - mu) @ U), axis=-1)
        return self.logq_const() - 0.5 * np.sum(np.log(s)) \
            - 0.5 * (self.df + self.dim) * np.log(1 + maha / self.df)

    def mean_and_cov(self, lam):
        mu, _, Sigma = self.fold(lam)
        return mu, self.df / (self.df - 2.) * Sigma                     # vb.py:217

    def pth_moment(self, p, lam):
        if p not in [2, 4]:
            raise ValueError('only p = 2 or 4 supported')
        df = self.df
        if df <= p:
            raise ValueError('df must be greater than p')
        _, _, Sigma = self.fold(lam)
        ev = np.linalg.eigvalsh(Sigma)                                  # vb.py:225
        c = df / (df - 2)
        if p == 2:
            return c * np.sum(ev)
        return c ** 2 * (2 * (df - 1) / (df - 4) * np.sum(ev ** 2) + np.sum(ev) ** 2)

    def weighted_grad(self, lam, base, w, g):
        """d/dlam of sum_i w_i log p(mu + ztilde_i sqrtm(L L'))  (Appendix A.2)."""
        d = self.dim
        mu, L, Sigma = self.fold(lam)
        zt = self.ztilde(base)
        g_mu = w @ g
        G_S = zt.T @ (w[:, None] * g)                                   # d x d
        if self.scale == 'chol':
            phi = G_S.T[self.tri].copy()
            phi[self.diag_pos] *= np.diag(L)
            return np.concatenate([g_mu, phi])
        S, lamb, V = self.sqrt_sigma(Sigma)
        # sqrtm VJP: solve S Y + Y S = G_S  (autograd: solve_sylvester(S', S', G))
        r = np.sqrt(lamb)
        Y = V @ ((V.T @ G_S @ V) / (r[:, None] + r[None, :])) @ V.T
        G_L = (Y + Y.T) @ L
        phi = G_L[self.tri].copy()
        phi[self.diag_pos] *= np.diag(L)                                # chain through exp
        return np.concatenate([g_mu, phi])

    def neg_logq_grad(self, lam, wsum):
        out = np.zeros(self.n_param)
        out[self.dim + self.diag_pos] = wsum
        return out


# --------------------------------------------------------------------------- public builders
def _wrap(impl, var_param_dim):
    def sample(var_param, n_samples, seed=None):
        rs = impl.rs if seed is None else np.random.RandomState(seed)
        return impl.transform(np.asarray(var_param, dtype=float), impl.draw_base(rs, n_samples))
    fam = VariationalFamily(sample, impl.entropy, impl.logdensity, impl.mean_and_cov,
                            impl.pth_moment, var_param_dim)
    fam.impl = impl
    return fam


def mean_field_gaussian_variational_family(dim):
    return _wrap(_MeanField(dim), 2 * dim)


def mean_field_t_variational_family(dim, df):
    if df <= 2:                                                         # vb.py:141-142
        raise ValueError('df must be greater than 2')
    return _wrap(_MeanField(dim, df), 2 * dim)


def t_variational_family(dim, df, scale='sqrtm'):
    if df <= 2:                                                         # vb.py:193-194
        raise ValueError('df must be greater than 2')
    impl = _FullRankT(dim, df, scale)
    return _wrap(impl, impl.n_param)


# --------------------------------------------------------------------------- objectives
def klvi_value_and_grad(impl, model, lam, base):
    """vb.py:237-243 with the base draw made explicit."""
    lam = np.asarray(lam, dtype=float)
    x = impl.transform(lam, base)
    lp, g = model.logp_and_grad(x)
    n = x.shape[0]
    w = np.full(n, 1.0 / n)
    grad = -(impl.weighted_grad(lam, base, w, g) + impl.neg_logq_grad(lam, 1.0))
    return -(impl.entropy(lam) + np.mean(lp)), grad


def chivi_value_and_grad(impl, model, alpha, lam, base, return_log_weights=False):
    """vb.py:249-264 with the base draw made explicit."""
    lam = np.asarray(lam, dtype=float)
    x = impl.transform(lam, base)
    lp, g = model.logp_and_grad(x)
    lw = lp - impl.logq_of_base(lam, base)
    log_norm = np.max(lw)                                               # vb.py:260
    s = np.exp(lw - log_norm) ** alpha                                  # vb.py:261
    value = np.log(np.mean(s)) / alpha + log_norm                       # vb.py:262
    n = s.size
    grad = alpha * (impl.weighted_grad(lam, base, s, g)
                    + impl.neg_logq_grad(lam, np.sum(s))) / n           # vb.py:263
    if return_log_weights:
        return value, grad, lw
    return value, grad


def black_box_klvi(var_family, logdensity, n_samples):
    impl = var_family.impl

    def objective_and_grad(var_param):
        base = impl.draw_base(impl.rs, n_samples)       # persistent stream, vb.py:239
        return klvi_value_and_grad(impl, logdensity, var_param, base)
    return objective_and_grad


def klvi_pd_value_and_grad(impl, model, lam, base):
    """vb.py:268-278 with the base draw made explicit: value -(mean log p - mean log q(x(lam); lam)); along the
    reparameterised sample log q depends on lam only through -entropy, so the gradient equals KLVI's."""
    lam = np.asarray(lam, dtype=float)
    x = impl.transform(lam, base)
    lp = model.logp_and_grad(x)[0]
    return -(np.mean(lp) - np.mean(impl.logq_of_base(lam, base))), klvi_value_and_grad(impl, model, lam, base)[1]


def black_box_klvi_pd(var_family, logdensity, n_samples):
    impl = var_family.impl

    def objective_and_grad(var_param):
        base = impl.draw_base(impl.rs, n_samples)       # persistent stream, vb.py:271
        return klvi_pd_value_and_grad(impl, logdensity, var_param, base)
    return objective_and_grad


black_box_klvi_pd2 = black_box_klvi_pd                  # vb.py:281-296: the same function through functools.partial


def black_box_chivi(alpha, var_family, logdensity, n_samples):
    impl = var_family.impl

    def objective_grad_and_log_norm(var_param):
        seed = np.random.randint(2 ** 32)                               # vb.py:258
        base = impl.draw_base(np.random.RandomState(seed), n_samples)
        return chivi_value_and_grad(impl, logdensity, alpha, var_param, base)
    return objective_grad_and_log_norm


# --------------------------------------------------------------------------- optimiser
def learning_rate_schedule(n_iters, learning_rate, learning_rate_end):
    """vb.py:324-343."""
    if learning_rate <= 0:
        raise ValueError('learning rate must be positive')
    if learning_rate_end is not None:
        if learning_rate <= learning_rate_end:
            raise ValueError('initial learning rate must be greater than final learning rate')
        b = n_iters * learning_rate_end / (2 * (learning_rate - learning_rate_end))
        a = learning_rate * b
        lo, hi = n_iters // 4, 3 * n_iters // 4
    for i in range(n_iters):
        if learning_rate_end is None or i < lo:
            yield learning_rate
        elif i < hi:
            yield a / (b + i - lo + 1)
        else:
            yield learning_rate_end


def adagrad_optimize(n_iters, objective_and_grad, init_param, has_log_norm=False,
                     window=10, learning_rate=.01, epsilon=.1, learning_rate_end=None):
    """vb.py:345-389 without the progress bar."""
    grads, norms = [], []
    values, norm_hist, iterates = [], [], []
    lam = np.array(init_param, dtype=float)
    for i, lr in enumerate(learning_rate_schedule(n_iters, learning_rate, learning_rate_end)):
        if has_log_norm:
            val, grad, log_norm = objective_and_grad(lam)
        else:
            val, grad = objective_and_grad(lam)
            log_norm = 0
        values.append(val)
        grads.append(grad)
        norms.append(log_norm)
        norm_hist.append(log_norm)
        if len(grads) > window:
            grads.pop(0)
            norms.pop(0)
        scale = np.exp(np.min(norms) - np.array(norms))                 # vb.py:371
        accum = np.sum((scale[:, None] * np.array(grads)) ** 2, axis=0)
        lam = lam - lr * grad / np.sqrt(epsilon + accum)                # vb.py:374
        if i >= 3 * n_iters // 4:
            iterates.append(lam.copy())
    iterates = np.array(iterates)
    return np.mean(iterates, axis=0), iterates, np.array(values), np.array(norm_hist)
