"""numpy restatement of ``viabel/bounds.py`` (oracle; test infrastructure only).

Every public function keeps the reference's name, argument order, defaults, error and
warning behaviour.  Citations are ``/root/reference/viabel/bounds.py:<line>``.
"""
import warnings

import numpy as np

__all__ = ['all_bounds', 'error_bounds', 'wasserstein_bounds', 'divergence_bound',
           'mean_and_check_mc_error', 'var_bound']


def mean_and_check_mc_error(a, atol=0.01, rtol=0.0, quantity_name=None):
    """bounds.py:183-192 — mean, plus a warning when std/sqrt(n) (ddof 0) exceeds
    ``rtol*|mean| + atol``."""
    a = np.asarray(a)
    mean = np.mean(a)
    mc_sd = np.std(a) / np.sqrt(a.size)
    if mc_sd > rtol * np.abs(mean) + atol:
        what = '' if quantity_name is None else ' when computing ' + quantity_name
        warnings.warn('significant Monte Carlo error' + what +
                      ' (mean = {}, standard deviation = {})'.format(mean, mc_sd))
    return mean


def divergence_bound(log_weights, alpha=2., log_norm_bound=None,
                     return_log_norm_bound=False):
    """bounds.py:142-180 — CUBO_alpha minus (ELBO or supplied bound), scaled by a/(a-1)."""
    if alpha <= 1:                                     # bounds.py:166-167
        raise ValueError('alpha must be greater than 1')
    lw = np.asarray(log_weights)
    shift = np.max(lw)                                 # bounds.py:169
    w_alpha = np.exp(lw - shift) ** alpha              # bounds.py:170
    cubo = np.log(mean_and_check_mc_error(w_alpha, quantity_name='CUBO')) / alpha + shift
    if log_norm_bound is None:                         # bounds.py:174-176
        log_norm_bound = mean_and_check_mc_error(lw, quantity_name='ELBO')
    d_alpha = alpha / (alpha - 1) * (cubo - log_norm_bound)
    return (d_alpha, log_norm_bound) if return_log_norm_bound else d_alpha


def wasserstein_bounds(d2, samples=None, moment_bound_fn=None):
    """bounds.py:103-139 — W_p <= 2 C_{2p}^{1/(2p)} expm1(d2)^{1/(2p)}, p = 1, 2, where
    C_{2p} = mean_i sum_j (x_ij - mean_j)^{2p} unless ``moment_bound_fn`` supplies it."""
    if moment_bound_fn is None:
        if samples is None:                            # bounds.py:128-129
            raise ValueError('must provides samples if moment_bound_fn not given')
        x = np.asarray(samples)
        if x.ndim == 1:                                # bounds.py:131-132
            x = x[:, None]
        centred = x - np.mean(x, axis=0, keepdims=True)

        def moment_bound_fn(order):                    # bounds.py:135
            return np.mean(np.sum(centred ** order, axis=1))
    out = {}
    for p in (1, 2):                                   # bounds.py:136-138
        c = moment_bound_fn(2 * p)
        out['W%d' % p] = 2 * c ** (.5 / p) * np.expm1(d2) ** (.5 / p)
    return out


def _spectral_norm_if_matrix(v):
    """bounds.py:64-67."""
    return np.linalg.norm(v, ord=2) if np.asarray(v).ndim == 2 else v


def var_bound(W2, var1, var2=None):
    """bounds.py:207-213."""
    lo = var1 if var2 is None else np.min([var1, var2], axis=0)
    return 2 * (np.sqrt(lo) * W2 + W2 ** 2)


def error_bounds(W1=np.inf, W2=np.inf, q_var=np.inf, p_var=np.inf):
    """bounds.py:70-100 — mean_error = min(W1, W2); std_error = W2;
    cov_error = 2 (sqrt(min var) W2 + W2^2) with matrices replaced by their 2-norm."""
    return {'mean_error': min(W1, W2),
            'std_error': W2,
            'cov_error': var_bound(W2, _spectral_norm_if_matrix(q_var),
                                   _spectral_norm_if_matrix(p_var))}


def all_bounds(log_weights, samples=None, moment_bound_fn=None,
               q_var=None, p_var=None, log_norm_bound=None):
    """bounds.py:13-61."""
    d2, log_norm_bound = divergence_bound(log_weights, log_norm_bound=log_norm_bound,
                                          return_log_norm_bound=True)
    res = wasserstein_bounds(d2, samples, moment_bound_fn)
    if q_var is None and samples is not None:          # bounds.py:55-56 (ddof = 1)
        q_var = np.cov(np.asarray(samples).T)
    res.update(error_bounds(q_var=q_var, p_var=p_var, **res))
    res['d2'] = d2
    res['log_norm_bound'] = log_norm_bound
    return res
