"""Drop-in for the convergence diagnostics of ``viabel/functions.py`` (``compute_R_hat``,
``compute_R_hat_adaptive_numpy``, ``compute_R_hat_halfway``, ``stochastic_iterate_averaging``): same names,
arguments, defaults, return shapes and error behaviour; the passes over the parameter history run as CUDA
reductions (``csrc/rhat.cu``).  The history may be a numpy array (copied to the device once) or a CUDA tensor
(used in place, results stay on the device)."""
import numpy as np
import torch

from . import _lib
from ._device import WS, is_host, ptr, stream_ptr, to_dev

__all__ = ['compute_R_hat', 'compute_R_hat_adaptive_numpy', 'compute_R_hat_halfway',
           'stochastic_iterate_averaging']


def _back(t, host):
    return t.cpu().numpy() if host else t


def _chains_dev(chains):
    host = is_host(chains)
    c = to_dev(chains, _lib.F64)
    if c.dim() != 3:
        raise ValueError('chains must have shape (n_chains, n_iters, n_params)')
    return c.contiguous(), host


def _rhat_windows(chains_dev, windows):
    """(var_hat, rhat), each (n_windows, K), for [(begin, length)] windows of a device history."""
    n_chains, n_iters, K = chains_dev.shape
    for b, ln in windows:
        if ln % 2 or ln < 2 or b < 0 or b + ln > n_iters:
            raise ValueError('cannot split %d iterations into two halves' % ln)
    nw = len(windows)
    dev = chains_dev.device
    win = torch.tensor(np.asarray(windows, dtype=np.int64).reshape(-1), device=dev)
    var_hat = torch.empty((nw, K), dtype=torch.float64, device=dev)
    rhat = torch.empty((nw, K), dtype=torch.float64, device=dev)
    ws = WS.get('rhat', _lib.lib.vb_rhat_workspace_bytes(n_chains, K, nw))
    _lib.check(_lib.lib.vb_rhat_windows(ptr(chains_dev), n_chains, n_iters, K, ptr(win), nw, ptr(var_hat), ptr(rhat),
                                        ptr(ws), ws.numel(), stream_ptr()), 'vb_rhat_windows')
    return var_hat, rhat


def _kept_window(warmup, n_iters):
    """The iterations compute_R_hat keeps (functions.py:12-21).  An odd count ends in the reference's own
    reshape error: its trim drops two rows instead of one."""
    begin = min(max(int(warmup), 0), n_iters) if warmup >= 0 else max(n_iters + int(warmup), 0)
    kept = n_iters - begin
    if kept % 2 == 1:
        raise ValueError('cannot reshape array: compute_R_hat needs an even number of iterations after warmup '
                         '(got %d)' % kept)
    return begin, kept


def compute_R_hat(chains, warmup=500):
    """Split R-hat per parameter of a (n_chains, n_iters, K) history (``functions.py:8-31``)
    -> ``(var_hat, R_hat)``, each ``(K,)``."""
    c, host = _chains_dev(chains)
    var_hat, rhat = _rhat_windows(c, [_kept_window(warmup, c.shape[1])])
    return _back(var_hat[0], host), _back(rhat[0], host)


def compute_R_hat_adaptive_numpy(chains, window_size=100):
    """R-hat of every consecutive window of ``window_size`` iterations (``functions.py:44-52``) -> (n_windows, K)."""
    c, host = _chains_dev(chains)
    n_iters = c.shape[1]
    n_windows = n_iters // window_size
    if n_windows * window_size != n_iters:
        raise ValueError('cannot reshape array: n_iters = %d is not a multiple of window_size = %d'
                         % (n_iters, window_size))
    _kept_window(0, window_size)
    return _back(_rhat_windows(c, [(i * window_size, window_size) for i in range(n_windows)])[1], host)


def compute_R_hat_halfway(chains, interval=100, start=1000):
    """R-hat of the second half of the growing prefixes ``chains[:, :start + (i + 1) * interval]``
    (``functions.py:54-65``) -> (n_iters // interval, K)."""
    c, host = _chains_dev(chains)
    n_iters = c.shape[1]
    windows = []
    for i in range(n_iters // interval):
        length = min(start + (i + 1) * interval, n_iters)
        windows.append(_kept_window(length // 2, length))
    if not windows:
        return _back(torch.empty((0,), dtype=torch.float64, device=c.device), host)
    return _back(_rhat_windows(c, windows)[1], host)


def stochastic_iterate_averaging(estimate, start):
    """Running mean of ``estimate[start:]`` along the iterations (``functions.py:68-77``)
    -> ``(estimate_iters (N - start, K), estimate_mean (K,))``."""
    host = is_host(estimate)
    if not host and estimate.dtype == torch.float64 and estimate.dim() == 2 and estimate.stride(1) == 1:
        e = estimate                  # a column block of a device-resident history is read in place (row stride)
    else:
        e = to_dev(estimate, _lib.F64)
    if e.dim() != 2:
        raise ValueError('estimate must have shape (n_iters, n_params)')
    N, K = e.shape
    start = int(start)
    if N - start <= 0:
        # the reference raises a str here, which Python reports as this TypeError
        raise TypeError('exceptions must derive from BaseException '
                        '(Start of stationary distribution must be lower than number of iterates)')
    if start < 0:
        start += N
    out = torch.empty((N - start, K), dtype=torch.float64, device=e.device)
    _lib.check(_lib.lib.vb_iterate_average(ptr(e), N, K, e.stride(0), start, ptr(out), stream_ptr()),
               'vb_iterate_average')
    return _back(out, host), _back(out[-1], host)
