"""numpy restatement of ``notebooks/psis.py`` (oracle; test infrastructure only).

Algorithm: Vehtari, Simpson, Gelman, Yao, Gabry — "Pareto smoothed importance sampling"
(arXiv:1507.02646); GPD fit: Zhang & Stephens (2009).  Citations are
``/root/reference/notebooks/psis.py:<line>``.

One deliberate knob that the reference does not have: ``sort_kind``.  The reference calls
``np.argsort`` with its default (unstable) kind (psis.py:167,182); with tied tail values its
permutation is build dependent.  ``sort_kind='stable'`` is the documented tie-break contract
of the CUDA path (ascending original index); on tie-free input both kinds agree bit for bit.
"""
import numpy as np

__all__ = ['psislw', 'psislw_detail', 'psisloo', 'gpdfitnew', 'gpinv', 'sumlogs']

_K_MIN = 1.0 / 3.0                                       # psis.py:160
_LOG_TINY = np.log(np.finfo(float).tiny)                 # psis.py:158


def sumlogs(x, axis=None, out=None):
    """psis.py:379-395 — max-shifted log-sum-exp."""
    top = x.max(axis=axis, keepdims=True)
    total = np.sum(np.exp(x - top), axis=axis, out=out)
    if isinstance(total, np.ndarray):
        np.log(total, out=total)
    else:
        total = np.log(total)
    total += np.squeeze(top)
    return total


def gpinv(p, k, sigma):
    """psis.py:334-376 — inverse CDF of the generalised Pareto distribution."""
    p = np.asarray(p, dtype=float)
    q = np.full(p.shape, np.nan)
    if sigma <= 0:                                       # psis.py:338-339
        return q
    inside = (p > 0) & (p < 1)
    tiny_k = np.abs(k) < np.finfo(float).eps

    def body(pp):                                        # psis.py:342-351 / 354-370
        t = np.log1p(-pp)
        return -t if tiny_k else np.expm1(-k * t) / k

    if np.all(inside):
        q = body(p) * sigma
    else:
        q[inside] = body(p[inside])
        q *= sigma
        q[p == 0] = 0                                    # psis.py:372
        q[p == 1] = np.inf if k >= 0 else -sigma / k     # psis.py:373-376
    return q


def gpdfitnew(x, sort=True, sort_in_place=False, return_quadrature=False):
    """psis.py:211-331 — Zhang–Stephens empirical-Bayes GPD fit; returns (k, sigma) with
    the sign convention k = -k_ZS and the weak prior k*n/(n+10) + 5/(n+10)."""
    if x.ndim != 1 or len(x) <= 1:                       # psis.py:250-251
        raise ValueError("Invalid input array.")
    if sort is True:                                     # psis.py:254-265
        if sort_in_place:
            x.sort()
            order = None
        else:
            order = np.argsort(x)
    elif sort is False:
        order = None
    else:
        order = sort
    n = len(x)
    prior = 3                                            # psis.py:267
    m = 30 + int(np.sqrt(n))                             # psis.py:268
    q_idx = int(n / 4 + 0.5) - 1                         # psis.py:276/279
    x_q = x[q_idx] if order is None else x[order[q_idx]]
    x_max = x[-1] if order is None else x[order[-1]]

    j = np.arange(1, m + 1, dtype=float) - 0.5           # psis.py:270-274
    bs = 1 - np.sqrt(m / j)
    bs /= prior * x_q
    bs += 1 / x_max

    ks = np.mean(np.log1p(-bs[:, None] * x), axis=1)     # psis.py:282-285
    L = n * (np.log(-(bs / ks)) - ks - 1)                # psis.py:287-292
    with np.errstate(over='ignore'):                     # benign overflow, psis.py:294-295
        w = 1 / np.sum(np.exp(L - L[:, None]), axis=1)   # psis.py:294-297

    keep = w >= 10 * np.finfo(float).eps                 # psis.py:300-303
    if not np.all(keep):
        w = w[keep]
        bs = bs[keep]
    w /= w.sum()                                         # psis.py:305

    b = np.sum(bs * w)                                   # psis.py:308
    k = np.mean(np.log1p(-b * x))                        # psis.py:311-313
    if return_quadrature:                                # psis.py:314-318
        ks = np.mean(np.log1p(bs[:, None] * (-x)), axis=1)
    sigma = -k / b * n / (n - 0)                         # psis.py:320 (before the shrink)
    a = 10                                               # psis.py:322-323
    k = k * n / (n + a) + a * 0.5 / (n + a)
    if return_quadrature:
        ks *= n / (n + a)
        ks += a * 0.5 / (n + a)
        return k, sigma, ks, w
    return k, sigma


def _smooth_column(x, Reff, sort_kind):
    """One column of psis.py:163-202, in place.  Returns (k, detail dict)."""
    n = x.shape[0]
    kind = None if sort_kind == 'default' else sort_kind
    cutoff_ind = -int(np.ceil(min(0.2 * n, 3 * np.sqrt(n / Reff)))) - 1   # psis.py:157
    x -= np.max(x)                                       # psis.py:165
    order = np.argsort(x, kind=kind)                     # psis.py:167
    xcutoff = max(x[order[cutoff_ind]], _LOG_TINY)       # psis.py:169-172
    exp_cut = np.exp(xcutoff)
    tailinds, = np.where(x > xcutoff)                    # psis.py:174 (strict)
    tail = x[tailinds]
    n2 = len(tail)
    sigma = np.nan
    tail_order = np.zeros(0, dtype=np.int64)
    if n2 <= 4:                                          # psis.py:177-179
        k = np.inf
    else:
        tail_order = np.argsort(tail, kind=kind)         # psis.py:182
        excess = np.exp(tail) - exp_cut                  # psis.py:184-185
        k, sigma = gpdfitnew(excess, sort=tail_order)    # psis.py:186
    smoothed = bool(k >= _K_MIN and not np.isinf(k))     # psis.py:187
    if smoothed:
        probs = np.arange(0.5, n2) / n2                  # psis.py:190-191
        qq = np.log(gpinv(probs, k, sigma) + exp_cut)    # psis.py:192-194
        x[tailinds[tail_order]] = qq                     # psis.py:196
        x[x > 0] = 0                                     # psis.py:198
    x -= sumlogs(x)                                      # psis.py:200
    return k, dict(xcutoff=xcutoff, tailinds=tailinds, tail_order=tail_order,
                   n2=n2, sigma=sigma, smoothed=smoothed, cutoff_ind=cutoff_ind)


def psislw_detail(lw, Reff=1.0, sort_kind='default'):
    """1-D ``psislw`` that also returns the intermediate tail bookkeeping (for parity
    tests of the tail index set and sort permutation)."""
    lw = np.asarray(lw, dtype=float)
    if lw.ndim != 1:
        raise ValueError("psislw_detail takes a 1-D array")
    if len(lw) <= 1:
        raise ValueError("More than one log-weight needed.")
    out = np.copy(lw)
    k, detail = _smooth_column(out, Reff, sort_kind)
    return out, k, detail


def psislw(lw, Reff=1.0, overwrite_lw=False, sort_kind='default'):
    """psis.py:112-208 — Pareto smoothed importance sampling of log weights.
    ``lw`` is (n,) or (n, m) (m independent columns).  Returns ``(lw_out, k)``."""
    if lw.ndim == 2:                                     # psis.py:136-144
        n, m = lw.shape
    elif lw.ndim == 1:
        n, m = len(lw), 1
    else:
        raise ValueError("Argument `lw` must be 1 or 2 dimensional.")
    if n <= 1:
        raise ValueError("More than one log-weight needed.")
    if overwrite_lw and lw.flags.f_contiguous:           # psis.py:146-151
        lw_out = lw
    else:
        lw_out = np.copy(lw, order='F')
    kss = np.empty(m)
    cols = lw_out.T if lw_out.ndim == 2 else lw_out[None, :]
    for i, col in enumerate(cols):
        kss[i], _ = _smooth_column(col, Reff, sort_kind)
    if lw_out.ndim == 1:                                 # psis.py:205-206
        kss = kss[0]
    return lw_out, kss


def psisloo(log_lik, **kwargs):
    """psis.py:68-109 -- PSIS leave-one-out log predictive densities from an (n, m) matrix of log-likelihood
    draws: ``lw = -log_lik`` smoothed column by column, ``loos = sumlogs(lw + log_lik, axis=0)``."""
    kwargs['overwrite_lw'] = True                        # psis.py:100
    lw = -log_lik                                        # psis.py:102
    lw, ks = psislw(lw, **kwargs)                        # psis.py:104
    lw += log_lik                                        # psis.py:106
    loos = sumlogs(lw, axis=0)                           # psis.py:107
    return loos.sum(), loos, ks                          # psis.py:108-109
