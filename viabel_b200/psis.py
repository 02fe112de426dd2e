"""Drop-in for the PSIS functions of ``notebooks/psis.py`` (``psislw``, ``gpdfitnew``,
``gpinv``, ``sumlogs``) on the device (``csrc/psis.cu``).

Tie-break contract: where the reference's ``np.argsort`` (default, unstable kind;
``psis.py:167,182``) leaves the order of equal tail values to the numpy build, this
implementation orders them by ascending original index (== ``kind='stable'``).
"""
import numpy as np
import torch

from . import _lib
from ._device import WS, is_host, like_input, ptr, stream_ptr, to_dev

__all__ = ['psislw', 'psislw_detail', 'psisloo', 'gpdfitnew', 'gpinv', 'sumlogs']


class _DivCache:
    """Statistics of divergence_bound(alpha = 2) taken by psislw's final pass over the same weights
    (``vb_psis_smooth_div`` / ``vb_psis_write_div``).  ``all_bounds(lw)`` / ``divergence_bound(lw)`` called on the
    SAME, unmodified CUDA tensor right after ``psislw(lw)`` -- the order of ``notebooks/experiments.py:69,159-166`` --
    then needs no further pass over the weights.  One entry; keyed on the tensor object (weak reference), its
    storage pointer, length and in-place modification counter; the world size it was computed under."""
    ref = None
    key = None
    partials = None
    host = None          # the merged statistics as host numbers (single-GPU call: they came back with k-hat)

    @classmethod
    def store(cls, t, partials, world, host=None):
        import weakref
        cls.ref, cls.key, cls.partials = weakref.ref(t), (t.data_ptr(), t.numel(), t._version, world), partials
        cls.host = host

    @classmethod
    def lookup_host(cls, t, world):
        return cls.host if cls.lookup(t, world) is not None else None

    @classmethod
    def lookup(cls, t, world):
        if cls.ref is None or cls.ref() is not t:
            return None
        if cls.key != (t.data_ptr(), t.numel(), t._version, world):
            return None
        return cls.partials


def _transposed(x, scale):
    """(rows, cols) CUDA f64 -> contiguous (cols, rows) = scale * x' in one pass (``vb_transpose_scale_f64``)."""
    x = x.contiguous()
    rows, cols = x.shape
    out = torch.empty((cols, rows), dtype=torch.float64, device=x.device)
    _lib.check(_lib.lib.vb_transpose_scale_f64(ptr(x), rows, cols, float(scale), ptr(out), stream_ptr()),
               'vb_transpose_scale_f64')
    return out


def _smooth_1d(col, Reff, want_tail=False, src=None):
    """col: CUDA f64 (n,) contiguous -> (smoothed (n,), diag (8,) numpy, tail idx or None).
    ``src``: the caller's tensor when ``col`` is it (no copy / cast happened): its divergence partials are kept."""
    n = col.numel()
    out = torch.empty_like(col)
    diag = torch.empty(8, dtype=torch.float64, device=col.device)
    tail = None
    if want_tail:
        cap = max(int(_lib.lib.vb_psis_tail_capacity(n, float(Reff))), 1)
        tail = torch.full((cap,), -1, dtype=torch.int64, device=col.device)
    ws = WS.get('psis', _lib.lib.vb_psis_smooth_workspace_bytes(n, float(Reff)))
    part = torch.empty(6, dtype=torch.float64, device=col.device) if src is not None else None
    # k-hat, the other diagnostics and the alpha = 2 divergence statistics come back to the host in the call's one
    # synchronisation (no further device read here or in all_bounds)
    host = np.empty(18)
    _lib.check(_lib.lib.vb_psis_smooth_host(ptr(col), ptr(out), n, float(Reff), ptr(diag), ptr(tail), ptr(part),
                                            host.ctypes.data, ptr(ws), ws.numel(), stream_ptr()), 'vb_psis_smooth_host')
    if src is not None:
        _DivCache.store(src, part, 1, host=host[8:16].copy())
    return out, host[:8].copy(), tail


def _smooth_columns(cols, Reff, want_loo=False):
    """cols: CUDA f64 (m, n) contiguous, one weight set per row -> (smoothed (m, n), k (m,) device, loo (m,) or None).
    One launch for all columns (``vb_psis_smooth_columns``); columns longer than the kernel holds in shared memory
    go through the single-column path one by one."""
    m, n = cols.shape
    if n <= int(_lib.lib.vb_psis_columns_max_n()):
        out = torch.empty_like(cols)
        ks = torch.empty(m, dtype=torch.float64, device=cols.device)
        loo = torch.empty(m, dtype=torch.float64, device=cols.device) if want_loo else None
        status = torch.empty(m, dtype=torch.int32, device=cols.device)
        _lib.check(_lib.lib.vb_psis_smooth_columns(ptr(cols), ptr(out), n, m, float(Reff), ptr(ks), ptr(loo),
                                                   ptr(status), stream_ptr()), 'vb_psis_smooth_columns')
        return out, ks, loo
    outs, ks = [], np.empty(m)
    for i in range(m):
        o, diag, _ = _smooth_1d(cols[i], Reff)
        outs.append(o)
        ks[i] = diag[0]
    out = torch.stack(outs, dim=0)
    loo = torch.logsumexp(out - cols, dim=1) if want_loo else None
    return out, torch.from_numpy(ks).to(cols.device), loo


def psisloo(log_lik, **kwargs):
    """PSIS leave-one-out log predictive densities (``psis.py:68-109``): ``log_lik`` is (n, m) -- n posterior
    draws of m log-likelihood terms; returns ``(loo, loos (m,), ks (m,))``.  Keyword arguments go to ``psislw``
    (``Reff``).  The smoothing of all columns and the leave-one-out sums run in one launch."""
    if log_lik.ndim != 2:
        raise ValueError("Argument `log_lik` must be 2 dimensional.")
    Reff = kwargs.pop('Reff', 1.0)
    kwargs.pop('overwrite_lw', None)                     # psis.py:100 forces it; the input is never touched here
    if kwargs:
        raise TypeError('unexpected keyword arguments: %s' % sorted(kwargs))
    if log_lik.shape[0] <= 1:
        raise ValueError("More than one log-weight needed.")
    cols = _transposed(to_dev(log_lik, _lib.F64), -1.0)             # lw = -log_lik, one column per row
    _, ks, loos = _smooth_columns(cols, Reff, want_loo=True)
    loo = loos.sum()
    if is_host(log_lik):
        return float(loo.item()), loos.cpu().numpy(), ks.cpu().numpy()
    return loo, loos, ks


def psislw(lw, Reff=1.0, overwrite_lw=False, group=None, out=None):
    """Pareto smoothed importance sampling (``psis.py:112-208``).

    ``group``: torch.distributed group (or ``True`` for the world group) over which a 1-D ``lw`` is
    sharded by contiguous index ranges; every rank passes its slice and gets its slice back, k-hat is
    the global one (``viabel_b200/psis_dist.py``).  Default: not sharded.

    ``lw``: (n,) or (n, m) log weights, numpy or CUDA tensor.  Returns ``(lw_out, kss)`` with
    ``kss`` a float for 1-D input and an (m,) numpy array otherwise.  ``overwrite_lw`` smooths
    a CUDA tensor / F-contiguous numpy array in place like the reference does.  ``out`` (1-D host input only): a
    float64 C-contiguous numpy array of the same length that receives the result; when it lives in pinned memory
    the device-to-host copy runs at link speed instead of through the driver's pageable staging."""
    if lw.ndim == 2:
        n, m = lw.shape
    elif lw.ndim == 1:
        n, m = len(lw), 1
    else:
        raise ValueError("Argument `lw` must be 1 or 2 dimensional.")
    sharded = group is not None and group is not False
    out_host, out = out, None
    if n <= 1 and not sharded:                           # a shard may hold one weight; the global check is in psis_dist
        raise ValueError("More than one log-weight needed.")
    dev = to_dev(lw, _lib.F64)
    kss = np.empty(m)
    if sharded:
        if lw.ndim != 1:
            raise ValueError('sharded psislw takes one weight column')
        from .psis_dist import psislw_sharded
        col = dev.contiguous()
        out, kss = psislw_sharded(col, Reff, group=None if group is True else group, src=lw if col is lw else None)
    elif lw.ndim == 1:
        col = dev.contiguous()
        out, diag, _ = _smooth_1d(col, Reff, src=lw if col is lw else None)
        kss = diag[0]
    else:
        cols = _transposed(dev, 1.0)                     # (m, n): one contiguous row per weight set
        out_cols, ks_dev, _ = _smooth_columns(cols, Reff)
        kss = ks_dev.cpu().numpy()
        out = _transposed(out_cols, 1.0)
    if is_host(lw):
        if out_host is not None:
            if lw.ndim != 1 or not isinstance(out_host, np.ndarray) or out_host.dtype != np.float64 \
                    or out_host.shape != (n,) or not out_host.flags.c_contiguous:
                raise ValueError('`out` must be a C-contiguous float64 numpy array of shape (n,) for 1-D host input')
            torch.from_numpy(out_host).copy_(out)
            return out_host, kss
        res = out.cpu().numpy()
        if lw.ndim == 2:
            res = np.asfortranarray(res)                 # psis.py:151 allocates F-order
        if overwrite_lw and isinstance(lw, np.ndarray) and lw.flags.f_contiguous:
            lw[...] = res                                # psis.py:146-148
            res = lw
        return res, kss
    if overwrite_lw:
        lw.copy_(out)
        out = lw
    return out, kss


def psislw_detail(lw, Reff=1.0):
    """1-D ``psislw`` that also returns the tail bookkeeping used by the parity tests:
    dict(k, sigma, xcutoff, n2, smoothed, max, lse, tail_idx_sorted) where
    ``tail_idx_sorted[r]`` is the original index of the r-th smallest tail element
    (``tailinds[x2si]`` in ``psis.py:196``)."""
    if lw.ndim != 1:
        raise ValueError('psislw_detail takes a 1-D array')
    if len(lw) <= 1:
        raise ValueError("More than one log-weight needed.")
    dev = to_dev(lw, _lib.F64).contiguous()
    out, diag, tail = _smooth_1d(dev, Reff, want_tail=True)
    n2 = int(diag[3])
    detail = dict(k=diag[0], sigma=diag[1], xcutoff=diag[2], n2=n2, smoothed=bool(diag[4]), max=diag[5],
                  lse=diag[6], n_candidates=int(diag[7]), tail_idx_sorted=tail[:n2].cpu().numpy())
    return like_input(out, lw), diag[0], detail


def gpdfitnew(x, sort=True, sort_in_place=False, return_quadrature=False):
    """Zhang–Stephens GPD fit (``psis.py:211-331``): returns ``(k, sigma)`` or, with ``return_quadrature``,
    ``(k, sigma, ks, w)`` (quadrature points and weights of the marginal posterior of k)."""
    if x.ndim != 1 or len(x) <= 1:
        raise ValueError("Invalid input array.")
    dev = to_dev(x, _lib.F64)
    if sort is False:
        xs = dev.contiguous()
    elif sort is True:
        xs = _sorted_1d(dev)
        if sort_in_place:                                 # psis.py:255-257
            if isinstance(x, np.ndarray):
                x[...] = xs.cpu().numpy()
            elif isinstance(x, torch.Tensor):
                x.copy_(xs)
    else:
        idx = torch.as_tensor(np.asarray(sort) if not isinstance(sort, torch.Tensor) else sort, device=dev.device)
        xs = dev[idx.long()].contiguous()
    n = xs.numel()
    out = torch.empty(2, dtype=torch.float64, device=xs.device)
    ws = WS.get('gpdfit', _lib.lib.vb_gpd_fit_workspace_bytes(n))
    if not return_quadrature:
        _lib.check(_lib.lib.vb_gpd_fit(ptr(xs), n, ptr(out), ptr(ws), ws.numel(), stream_ptr()), 'vb_gpd_fit')
        k, sigma = out.cpu().numpy()
        return k, sigma
    m = 30 + int(np.sqrt(n))                              # psis.py:268
    ks = torch.empty(m, dtype=torch.float64, device=xs.device)
    w = torch.empty(m, dtype=torch.float64, device=xs.device)
    _lib.check(_lib.lib.vb_gpd_fit_quadrature(ptr(xs), n, ptr(out), ptr(ks), ptr(w), ptr(ws), ws.numel(), stream_ptr()),
               'vb_gpd_fit_quadrature')
    k, sigma = out.cpu().numpy()
    ks, w = ks.cpu().numpy(), w.cpu().numpy()
    keep = w >= 10 * np.finfo(float).eps                  # psis.py:300-305: drop negligible weights, normalise
    ks, w = ks[keep], w[keep]
    return k, sigma, ks, w / w.sum()


def _sorted_1d(dev):
    """Ascending sort of a 1-D device array with the library's own bitonic network (``vb_sort_f64``)."""
    v = dev.contiguous()
    n = v.numel()
    out = torch.empty_like(v)
    ws = WS.get('sort', _lib.lib.vb_sort_f64_workspace_bytes(n))
    _lib.check(_lib.lib.vb_sort_f64(ptr(v), n, ptr(out), ptr(ws), ws.numel(), stream_ptr()), 'vb_sort_f64')
    return out


def gpinv(p, k, sigma):
    """Inverse generalised Pareto CDF (``psis.py:334-376``)."""
    dev = to_dev(p, _lib.F64).contiguous()
    out = torch.empty_like(dev)
    if dev.numel():
        _lib.check(_lib.lib.vb_gpd_inv(ptr(dev), dev.numel(), float(k), float(sigma), ptr(out), stream_ptr()),
                   'vb_gpd_inv')
    return like_input(out, p)


def sumlogs(x, axis=None, out=None):
    """``log(sum(exp(x)))`` (``psis.py:379-395``).  ``axis``: None (all elements) or an axis of a 2-D array;
    ``out``: array that receives the result (returned), as in the reference."""
    dev = to_dev(x, _lib.F64)

    def one(v):
        v = v.contiguous().reshape(-1)
        res = torch.empty(1, dtype=torch.float64, device=v.device)
        ws = WS.get('lse', _lib.lib.vb_logsumexp_workspace_bytes(v.numel()))
        _lib.check(_lib.lib.vb_logsumexp(ptr(v), v.numel(), ptr(res), ptr(ws), ws.numel(), stream_ptr()),
                   'vb_logsumexp')
        return res

    if axis is None or dev.ndim == 1:
        r = one(dev)
        res = float(r.item()) if is_host(x) else r[0]
    elif dev.ndim == 2 and axis in (0, 1, -1, -2):
        cols = dev if axis in (0, -2) else dev.t()
        r = torch.cat([one(cols[:, j]) for j in range(cols.shape[1])])
        res = like_input(r, x)
    else:
        raise ValueError('sumlogs: axis must be None or an axis of a 1-D / 2-D array')
    if out is not None:
        if isinstance(out, np.ndarray):
            out[...] = res if not isinstance(res, torch.Tensor) else res.cpu().numpy()
        else:
            out.copy_(res if isinstance(res, torch.Tensor) else torch.as_tensor(res))
        return out
    return res
