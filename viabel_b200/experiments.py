"""The callers either side of the hot path in the reference's notebooks:
``get_samples_and_log_weights``, ``psis_correction``, ``improve_with_psis`` and ``check_accuracy`` of
``notebooks/experiments.py:25-89``, with the samples and log weights left on the device between the
stages (sample -> log p - log q -> psislw -> weighted mean / covariance never leave HBM).

Same names, argument order and return kinds as the reference; ``device=True`` returns CUDA tensors
instead of numpy arrays.
"""
import numpy as np
import torch

from . import _lib
from ._device import WS, like_input, ptr, stream_ptr, to_dev, torch_dtype, device as _device
from .models import Model
from .psis import psislw

__all__ = ['get_samples_and_log_weights', 'psis_correction', 'improve_with_psis', 'check_accuracy', 'check_approx_accuracy',
           'weighted_mean_and_cov']


def _samples_and_log_weights_device(logdensity, var_family, var_param, n_samples, base_draws=None):
    impl = getattr(var_family, 'impl', None)
    if impl is None or not isinstance(logdensity, Model):
        raise TypeError('needs a viabel_b200 variational family and a registered model plugin (no CPU fallback)')
    n = int(n_samples)
    vp = to_dev(var_param, _lib.F64)
    if impl.kind != _lib.FAMILY_FR_T and logdensity.fused:
        # one fused pass (vb_mf_samples_logweights): draw -> x -> log p - log q
        if base_draws is not None:
            base = impl.coerce_base(base_draws)['eps']
            seed = off = 0
        else:
            base = None
            seed, off = impl.next_stream(None)          # var_family.sample(var_param, n) advances the stream
        x = torch.empty((n, impl.dim), dtype=torch_dtype(impl.code), device=_device())
        lw = torch.empty(n, dtype=torch.float64, device=_device())
        _lib.check(_lib.lib.vb_mf_samples_logweights(impl.kind, impl.df, logdensity.handle, impl.code, ptr(vp), n,
                                                     ptr(base), seed, off, 0, ptr(x), 0, ptr(lw), stream_ptr()),
                   'vb_mf_samples_logweights')
        return x, lw
    # composed path (full-rank family, regression / gauss plugins): the same device ops the closures use
    if base_draws is not None:
        raise ValueError('base_draws is only supported on the fused mean-field path')
    x = var_family.sample(vp, n)                        # CUDA in -> CUDA out
    lw = logdensity(x) - var_family.logdensity(x, vp)
    return x, lw


def get_samples_and_log_weights(logdensity, var_family, var_param, n_samples, device=False, base_draws=None):
    """``notebooks/experiments.py:60-63``: ``samples (n, d)`` and ``log p(samples) - log q(samples)``."""
    x, lw = _samples_and_log_weights_device(logdensity, var_family, var_param, n_samples, base_draws)
    if device:
        return x, lw
    return like_input(x, var_param), like_input(lw, var_param)


def psis_correction(logdensity, var_family, var_param, n_samples, device=False, base_draws=None):
    """``notebooks/experiments.py:66-70``: ``(samples.T, smoothed_log_weights, khat)``."""
    x, lw = _samples_and_log_weights_device(logdensity, var_family, var_param, n_samples, base_draws)
    slw, khat = psislw(lw)
    if device:
        return x.t(), slw, khat
    return like_input(x, var_param).T, like_input(slw, var_param), khat


def weighted_mean_and_cov(samples, log_weights, want_cov=True):
    """Self-normalised importance-weighted mean and covariance of ``samples (n, d)`` with ``log_weights
    (n,)`` (``experiments.py:81-86``; ``np.cov(samples.T, aweights=w, ddof=0)``).  numpy or CUDA in, numpy out."""
    x = samples if isinstance(samples, torch.Tensor) and samples.is_cuda else to_dev(samples, _lib.F64)
    if x.ndim == 1:
        x = x[:, None]
    x = x.contiguous()
    code = _lib.F64 if x.dtype == torch.float64 else _lib.F32
    lw = to_dev(log_weights, _lib.F64).contiguous()
    n, d = x.shape
    if lw.shape != (n,):
        raise ValueError('log_weights must have shape (n,)')
    mean = torch.empty(d, dtype=torch.float64, device=x.device)
    cov = torch.empty((d, d), dtype=torch.float64, device=x.device) if want_cov else None
    ws = WS.get('wmoments', _lib.lib.vb_weighted_moments_workspace_bytes(n, d))
    _lib.check(_lib.lib.vb_weighted_moments(code, ptr(x), ptr(lw), n, d, ptr(mean), ptr(cov), ptr(ws), ws.numel(),
                                            stream_ptr()), 'vb_weighted_moments')
    return mean.cpu().numpy(), (cov.cpu().numpy() if want_cov else None)


def check_accuracy(true_mean, true_cov, approx_mean, approx_cov, verbose=False, method=None):
    """``notebooks/experiments.py:25-48`` (host arithmetic on d-vectors / d x d matrices)."""
    true_std = np.sqrt(np.diag(true_cov))
    approx_std = np.sqrt(np.diag(approx_cov))
    results = dict(mean_error=np.linalg.norm(true_mean - approx_mean),
                   cov_error_2=np.linalg.norm(true_cov - approx_cov, ord=2),
                   cov_norm_2=np.linalg.norm(true_cov, ord=2),
                   cov_error_nuc=np.linalg.norm(true_cov - approx_cov, ord='nuc'),
                   cov_norm_nuc=np.linalg.norm(true_cov, ord='nuc'),
                   std_error=np.linalg.norm(true_std - approx_std),
                   rel_std_error=np.linalg.norm(approx_std / true_std - 1))
    if method is not None:
        results['method'] = method
    if verbose:
        print('mean   =', approx_mean)
        print('stdevs =', approx_std)
        print()
        print('mean error             = {:.3g}'.format(results['mean_error']))
        print('stdev error            = {:.3g}'.format(results['std_error']))
        print('||cov error||_2^{{1/2}}  = {:.3g}'.format(np.sqrt(results['cov_error_2'])))
        print('||true cov||_2^{{1/2}}   = {:.3g}'.format(np.sqrt(results['cov_norm_2'])))
    return results


def check_approx_accuracy(var_family, var_param, true_mean, true_cov, verbose=False, name=None):
    """``notebooks/experiments.py:51-55``: accuracy of the family's own mean and covariance."""
    return check_accuracy(true_mean, true_cov, *var_family.mean_and_cov(var_param), verbose, name)


def improve_with_psis(logdensity, var_family, var_param, n_samples, true_mean, true_cov, transform=None,
                      verbose=False, base_draws=None):
    """``notebooks/experiments.py:73-89``: PSIS-corrected mean / covariance estimates and their errors.
    ``transform`` (if given) is applied to the ``(d, n)`` samples on the host, like the reference does."""
    x, lw = _samples_and_log_weights_device(logdensity, var_family, var_param, n_samples, base_draws)
    slw, khat = psislw(lw)
    if verbose:
        print('khat = {:.3g}'.format(khat))
        print()
    if transform is not None:
        xt = np.ascontiguousarray(transform(x.t().cpu().numpy()).T)       # reference transforms samples.T
        approx_mean, approx_cov = weighted_mean_and_cov(xt, slw)
    else:
        approx_mean, approx_cov = weighted_mean_and_cov(x, slw)
    res = check_accuracy(true_mean, true_cov, approx_mean, approx_cov, verbose)
    res['khat'] = khat
    return res, approx_mean, approx_cov
