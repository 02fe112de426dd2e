"""Drop-in for ``viabel/bounds.py``: same functions, arguments, defaults, result keys,
exceptions and warning text; the passes over ``log_weights`` / ``samples`` run as CUDA
streaming reductions (``csrc/bounds.cu``).  Inputs may be numpy arrays (copied to the device)
or CUDA tensors (used in place).  Scalars come back as Python floats."""
from warnings import warn

import numpy as np
import torch

from . import _lib
from ._device import WS, all_gather_rows, dtype_code, is_host, ptr, stream_ptr, to_dev

__all__ = [
    'all_bounds',
    'error_bounds',
    'wasserstein_bounds',
    'divergence_bound',
]


def _dist_world(group):
    import torch.distributed as dist
    if group is None or group is False or not dist.is_available() or not dist.is_initialized():
        return 1, None
    g = None if group is True else group
    return dist.get_world_size(g), g


def _divergence_stats(log_weights, alpha, group=None):
    """[max, mean w, mc-sd w, mean lw, mc-sd lw, cubo, n] for w = exp(alpha (lw - max))."""
    lw = to_dev(log_weights, _lib.F64).reshape(-1)
    n = lw.numel()
    out = torch.empty(8, dtype=torch.float64, device=lw.device)
    ws = WS.get('bounds_div', _lib.lib.vb_bounds_divergence_workspace_bytes(n))
    world, g = _dist_world(group)
    # statistics psislw's final pass already took over this very tensor (alpha = 2 only): no further pass
    from .psis import _DivCache
    cached = _DivCache.lookup(log_weights, world) if (alpha == 2.0 and isinstance(log_weights, torch.Tensor)) else None
    if cached is not None and _DivCache.lookup_host(log_weights, world) is not None:
        return _DivCache.lookup_host(log_weights, world).copy()      # merged over the group by psislw already
    if world == 1 and cached is not None:
        _lib.check(_lib.lib.vb_bounds_divergence_merge(ptr(cached), 1, float(alpha), ptr(out), stream_ptr()),
                   'vb_bounds_divergence_merge')
    elif world == 1:
        _lib.check(_lib.lib.vb_bounds_divergence(ptr(lw), n, float(alpha), ptr(out), ptr(ws), ws.numel(),
                                                 stream_ptr()), 'vb_bounds_divergence')
    else:
        import torch.distributed as dist
        if cached is not None:
            part = cached
        else:
            part = torch.empty(6, dtype=torch.float64, device=lw.device)
            _lib.check(_lib.lib.vb_bounds_divergence_partials(ptr(lw), n, float(alpha), ptr(part), ptr(ws), ws.numel(),
                                                              stream_ptr()), 'vb_bounds_divergence_partials')
        gathered = all_gather_rows(part, world, g)
        _lib.check(_lib.lib.vb_bounds_divergence_merge(ptr(gathered), world, float(alpha), ptr(out), stream_ptr()),
                   'vb_bounds_divergence_merge')
    return out.cpu().numpy()


def _check_mc_error(m, s, quantity_name, atol=0.01, rtol=0.0):
    """Warning of ``mean_and_check_mc_error`` (``bounds.py:183-192``)."""
    if s > rtol * np.abs(m) + atol:
        msg = 'significant Monte Carlo error'
        if quantity_name is not None:
            msg += ' when computing ' + quantity_name
        msg += ' (mean = {}, standard deviation = {})'.format(m, s)
        warn(msg)


def divergence_bound(log_weights, alpha=2., log_norm_bound=None,
                     return_log_norm_bound=False, group=None):
    """Bound on the alpha-divergence from log weights (``bounds.py:142-180``).
    ``group``: torch.distributed group over which ``log_weights`` is sharded (default: not sharded)."""
    if alpha <= 1:
        raise ValueError('alpha must be greater than 1')
    st = _divergence_stats(log_weights, alpha, group)
    _check_mc_error(st[1], st[2], 'CUBO')
    cubo = st[5]
    if log_norm_bound is None:
        _check_mc_error(st[3], st[4], 'ELBO')
        log_norm_bound = st[3]
    dalpha = alpha / (alpha - 1) * (cubo - log_norm_bound)
    if return_log_norm_bound:
        return dalpha, log_norm_bound
    return dalpha


def _as_samples(samples):
    if isinstance(samples, torch.Tensor) and samples.is_cuda:
        code = dtype_code(samples.dtype)
    else:
        code = _lib.F64
    x = to_dev(samples, code)
    if x.ndim == 1:                                         # bounds.py:131-132
        x = x[:, None]
    return x.contiguous(), code


def _sample_moments(samples, want_cov, group=None):
    """(C2, C4, cov or None): C_{2p} = mean_i sum_j (x_ij - mean_j)^{2p}; cov with ddof = 1."""
    x, code = _as_samples(samples)
    n, d = x.shape
    world, g = _dist_world(group)
    ws = WS.get('bounds_mom', _lib.lib.vb_bounds_moments_workspace_bytes(n, d))
    colsum = torch.empty(d, dtype=torch.float64, device=x.device)
    _lib.check(_lib.lib.vb_bounds_colsum(code, ptr(x), n, d, ptr(colsum), ptr(ws), ws.numel(), stream_ptr()),
               'vb_bounds_colsum')
    n_total = n
    if world > 1:
        import torch.distributed as dist
        pack = torch.cat([colsum, torch.tensor([float(n)], dtype=torch.float64, device=x.device)])
        gathered = all_gather_rows(pack, world, g)
        tot = gathered.sum(0)
        colsum, n_total = tot[:d].contiguous(), int(round(float(tot[d])))
    mean = (colsum / n_total).contiguous()
    central = torch.empty(2, dtype=torch.float64, device=x.device)
    scatter = torch.empty((d, d), dtype=torch.float64, device=x.device) if want_cov else None
    _lib.check(_lib.lib.vb_bounds_central(code, ptr(x), n, d, ptr(mean), ptr(central), ptr(scatter), ptr(ws),
                                          ws.numel(), stream_ptr()), 'vb_bounds_central')
    if world > 1:
        import torch.distributed as dist
        pack = central if scatter is None else torch.cat([central, scatter.reshape(-1)])
        gathered = all_gather_rows(pack, world, g)
        tot = gathered.sum(0)
        central = tot[:2]
        if scatter is not None:
            scatter = tot[2:].reshape(d, d)
    c = central.cpu().numpy() / n_total
    cov = None
    if want_cov:
        cov = scatter.cpu().numpy() / (n_total - 1)                 # np.cov default ddof, bounds.py:56
        if d == 1:
            cov = cov.reshape(())                                   # np.cov of one variable is 0-d
    return c[0], c[1], cov


def wasserstein_bounds(d2, samples=None, moment_bound_fn=None, group=None):
    """1- and 2-Wasserstein bounds from a 2-divergence bound (``bounds.py:103-139``)."""
    results = dict()
    if moment_bound_fn is None:
        if samples is None:
            raise ValueError('must provides samples if moment_bound_fn not given')
        c2, c4, _ = _sample_moments(samples, False, group)
        moments = {2: c2, 4: c4}
        moment_bound_fn = moments.__getitem__
    for p in [1, 2]:
        Cp = moment_bound_fn(2 * p)
        results['W{}'.format(p)] = 2 * Cp ** (.5 / p) * np.expm1(d2) ** (.5 / p)
    return results


def _compute_norm_if_needed(var):
    if isinstance(var, torch.Tensor):
        var = var.detach().cpu().numpy()
    if np.asarray(var).ndim == 2:
        return np.linalg.norm(var, ord=2)
    return var


def mean_bound(Wp):
    return Wp


def std_bound(W2):
    return W2


def var_bound(W2, var1, var2=None):
    """``bounds.py:207-213``."""
    if var2 is not None:
        min_var = np.min([var1, var2], axis=0)
    else:
        min_var = var1
    min_std = np.sqrt(min_var)
    return 2 * (min_std * W2 + W2 ** 2)


def error_bounds(W1=np.inf, W2=np.inf, q_var=np.inf, p_var=np.inf):
    """Mean / std / covariance error bounds from Wasserstein bounds (``bounds.py:70-100``).
    O(d^3) host arithmetic on at most two (d, d) matrices."""
    results = dict()
    results['mean_error'] = mean_bound(min(W1, W2))
    results['std_error'] = std_bound(W2)
    results['cov_error'] = var_bound(W2, _compute_norm_if_needed(q_var), _compute_norm_if_needed(p_var))
    return results


def all_bounds(log_weights, samples=None, moment_bound_fn=None,
               q_var=None, p_var=None, log_norm_bound=None, group=None):
    """All error and distance bounds (``bounds.py:13-61``).  Result keys: ``W1, W2, mean_error,
    std_error, cov_error, d2, log_norm_bound``."""
    d2, log_norm_bound = divergence_bound(log_weights, log_norm_bound=log_norm_bound,
                                          return_log_norm_bound=True, group=group)
    need_cov = q_var is None and samples is not None
    if moment_bound_fn is None and samples is not None:
        c2, c4, cov = _sample_moments(samples, need_cov, group)
        moments = {2: c2, 4: c4}
        results = wasserstein_bounds(d2, moment_bound_fn=moments.__getitem__)
        if need_cov:
            q_var = cov
    else:
        results = wasserstein_bounds(d2, samples, moment_bound_fn, group=group)
        if need_cov:
            q_var = _sample_moments(samples, True, group)[2]
    results.update(error_bounds(q_var=q_var, p_var=p_var, **results))
    results['d2'] = d2
    results['log_norm_bound'] = log_norm_bound
    return results
