"""PyTorch custom ops over the C ABI (``torch.ops.viabel_b200.*``).

BASELINE.json's north star asks for the CUDA path to be reachable as PyTorch custom ops.  Each op
below is a thin shim: it takes tensors, allocates outputs / workspace with torch, and forwards raw
pointers + sizes + the current stream to one ``extern "C"`` entry point of ``libviabel_b200.so``
(``include/viabel_b200.h``).  The ops are registered for the CUDA dispatch key only -- calling them
with CPU tensors raises (there is no CPU fallback).  Model plugins are passed as their integer
handle (``Model.handle.value``).
"""
import torch

from . import _lib
from ._device import Workspace, ptr, stream_ptr

_WS = Workspace()
_lib_def = torch.library.Library('viabel_b200', 'DEF')
_lib_def.define('mf_step(int family, float df, int model, int objective, float alpha, Tensor var_param, int n, '
                'Tensor? base, int seed, int call_offset, int sample_index0) -> (Tensor, Tensor)')
_lib_def.define('mf_finalize(int family, int objective, float alpha, Tensor var_param, Tensor partials) -> Tensor')
_lib_def.define('model_logp_grad(int model, Tensor x, bool want_grad) -> (Tensor, Tensor)')
_lib_def.define('bounds_divergence(Tensor lw, float alpha) -> Tensor')
_lib_def.define('psis_smooth(Tensor lw, float reff) -> (Tensor, Tensor)')
_lib_def.define('psis_smooth_columns(Tensor lw_cols, float reff, bool want_loo) -> (Tensor, Tensor, Tensor)')
_lib_def.define('mf_adagrad(int family, float df, int model, int objective, float alpha, Tensor init_param, int n, '
                'Tensor rates, Tensor seeds, Tensor offsets, int window, float epsilon) -> (Tensor, Tensor, Tensor)')
_lib_def.define('logsumexp(Tensor x) -> Tensor')
_lib_def.define('rhat_windows(Tensor chains, Tensor windows) -> (Tensor, Tensor)')
_lib_def.define('iterate_average(Tensor estimate, int start) -> Tensor')
_lib_def.define('gemm_tf32x3(Tensor a, Tensor b) -> Tensor')


def _code(t):
    if t.dtype == torch.float64:
        return _lib.F64
    if t.dtype == torch.float32:
        return _lib.F32
    raise TypeError('viabel_b200 ops take float64 or float32 tensors')


def _mf_step(family, df, model, objective, alpha, var_param, n, base, seed, call_offset, sample_index0):
    """(out, partials): out = [value, grad (2d)], partials = VB_MF_NPART(d) doubles (vb_mf_step)."""
    d = var_param.numel() // 2
    out = torch.empty(1 + 2 * d, dtype=torch.float64, device=var_param.device)
    partials = torch.empty(2 * d + 4, dtype=torch.float64, device=var_param.device)
    ws = _WS.get('mf_step', _lib.lib.vb_mf_step_workspace_bytes(d))
    code = _lib.F64 if base is None else _code(base)
    _lib.check(_lib.lib.vb_mf_step(family, df, model, objective, alpha, code, ptr(var_param), n, ptr(base), seed,
                                   call_offset, sample_index0, ptr(out), ptr(out) + 8, ptr(partials), ptr(ws),
                                   ws.numel(), stream_ptr()), 'vb_mf_step')
    return out, partials


def _mf_finalize(family, objective, alpha, var_param, partials):
    d = var_param.numel() // 2
    out = torch.empty(1 + 2 * d, dtype=torch.float64, device=var_param.device)
    _lib.check(_lib.lib.vb_mf_finalize(family, objective, alpha, ptr(var_param), d, ptr(partials),
                                       partials.numel() // (2 * d + 4), ptr(out), ptr(out) + 8, stream_ptr()),
               'vb_mf_finalize')
    return out


def _model_logp_grad(model, x, want_grad):
    n = x.shape[0]
    code = _code(x)
    logp = torch.empty(n, dtype=torch.float64, device=x.device)
    grad = torch.empty_like(x) if want_grad else torch.empty(0, dtype=x.dtype, device=x.device)
    ws = _WS.get('model', _lib.lib.vb_model_logp_grad_workspace_bytes(model, n, code))
    _lib.check(_lib.lib.vb_model_logp_grad(model, code, ptr(x), n, ptr(logp), ptr(grad) if want_grad else 0,
                                           ptr(ws), ws.numel(), stream_ptr()), 'vb_model_logp_grad')
    return logp, grad


def _bounds_divergence(lw, alpha):
    out = torch.empty(8, dtype=torch.float64, device=lw.device)
    ws = _WS.get('div', _lib.lib.vb_bounds_divergence_workspace_bytes(lw.numel()))
    _lib.check(_lib.lib.vb_bounds_divergence(ptr(lw), lw.numel(), alpha, ptr(out), ptr(ws), ws.numel(), stream_ptr()),
               'vb_bounds_divergence')
    return out


def _psis_smooth(lw, reff):
    out = torch.empty_like(lw)
    diag = torch.empty(8, dtype=torch.float64, device=lw.device)
    ws = _WS.get('psis', _lib.lib.vb_psis_smooth_workspace_bytes(lw.numel(), reff))
    _lib.check(_lib.lib.vb_psis_smooth(ptr(lw), ptr(out), lw.numel(), reff, ptr(diag), 0, ptr(ws), ws.numel(),
                                       stream_ptr()), 'vb_psis_smooth')
    return out, diag


def _psis_smooth_columns(lw_cols, reff, want_loo):
    """(smoothed (m, n), k-hat (m,), loo (m,) or empty): one CTA per row of lw_cols (vb_psis_smooth_columns)."""
    m, n = lw_cols.shape
    out = torch.empty_like(lw_cols)
    ks = torch.empty(m, dtype=torch.float64, device=lw_cols.device)
    loo = torch.empty(m if want_loo else 0, dtype=torch.float64, device=lw_cols.device)
    status = torch.empty(m, dtype=torch.int32, device=lw_cols.device)
    _lib.check(_lib.lib.vb_psis_smooth_columns(ptr(lw_cols), ptr(out), n, m, reff, ptr(ks), ptr(loo) if want_loo else 0,
                                               ptr(status), stream_ptr()), 'vb_psis_smooth_columns')
    return out, ks, loo


def _mf_adagrad(family, df, model, objective, alpha, init_param, n, rates, seeds, offsets, window, epsilon):
    """Device-resident AdaGrad (vb_mf_adagrad_*): (final iterate (2d,), kept iterates (n_iters - 3 n_iters // 4, 2d),
    value history (n_iters,)).  rates float64, seeds int64, offsets int32, all (n_iters,) on the device."""
    import ctypes
    P, n_iters = init_param.numel(), rates.numel()
    d = P // 2
    dev = init_param.device
    param = init_param.to(torch.float64).clone()
    values = torch.zeros(n_iters, dtype=torch.float64, device=dev)
    kept = torch.zeros((n_iters - 3 * n_iters // 4, P), dtype=torch.float64, device=dev)
    state = torch.zeros(_lib.lib.vb_mf_adagrad_state_bytes(d, window), dtype=torch.uint8, device=dev)
    ws = torch.zeros(_lib.lib.vb_mf_step_workspace_bytes(d), dtype=torch.uint8, device=dev)
    plan = ctypes.c_void_p()
    _lib.check(_lib.lib.vb_mf_adagrad_create(family, df, model, objective, alpha, _lib.F64, ptr(param), n, 0, n_iters,
                                             window, epsilon, ptr(rates), ptr(seeds), ptr(offsets), ptr(values),
                                             ptr(kept), ptr(state), state.numel(), ptr(ws), ws.numel(),
                                             50 if n_iters >= 100 else 0, 1, stream_ptr(), ctypes.byref(plan)),
               'vb_mf_adagrad_create')
    try:
        _lib.check(_lib.lib.vb_mf_adagrad_run(plan, n_iters, stream_ptr()), 'vb_mf_adagrad_run')
        torch.cuda.current_stream().synchronize()        # the plan owns graph / state until the work has drained
    finally:
        _lib.lib.vb_mf_adagrad_destroy(plan)
    return param, kept, values


def _logsumexp(x):
    out = torch.empty(1, dtype=torch.float64, device=x.device)
    ws = _WS.get('lse', _lib.lib.vb_logsumexp_workspace_bytes(x.numel()))
    _lib.check(_lib.lib.vb_logsumexp(ptr(x), x.numel(), ptr(out), ptr(ws), ws.numel(), stream_ptr()), 'vb_logsumexp')
    return out


def _rhat_windows(chains, windows):
    """(var_hat, R-hat), each (n_windows, K): split R-hat of the iteration windows [(begin, even length)] (int64,
    (n_windows, 2), on the device) of a contiguous (n_chains, n_iters, K) float64 history (vb_rhat_windows)."""
    n_chains, n_iters, K = chains.shape
    nw = windows.shape[0]
    var_hat = torch.empty((nw, K), dtype=torch.float64, device=chains.device)
    rhat = torch.empty((nw, K), dtype=torch.float64, device=chains.device)
    ws = _WS.get('rhat', _lib.lib.vb_rhat_workspace_bytes(n_chains, K, nw))
    _lib.check(_lib.lib.vb_rhat_windows(ptr(chains), n_chains, n_iters, K, ptr(windows), nw, ptr(var_hat), ptr(rhat),
                                        ptr(ws), ws.numel(), stream_ptr()), 'vb_rhat_windows')
    return var_hat, rhat


def _iterate_average(estimate, start):
    """Running mean of estimate[start:] along dim 0 (vb_iterate_average); estimate (N, K) float64, unit column stride."""
    N, K = estimate.shape
    out = torch.empty((N - start, K), dtype=torch.float64, device=estimate.device)
    _lib.check(_lib.lib.vb_iterate_average(ptr(estimate), N, K, estimate.stride(0), start, ptr(out), stream_ptr()),
               'vb_iterate_average')
    return out


def _gemm_tf32x3(a, b):
    """C (m, n) fp32 = a (m, k) @ b (k, n) with 3xTF32 tcgen05 arithmetic (vb_gemm, dtype f32)."""
    m, k = a.shape
    n = b.shape[1]
    c = torch.empty((m, n), dtype=torch.float32, device=a.device)
    ws = _WS.get('gemm', _lib.lib.vb_gemm_workspace_bytes(_lib.F32, m, n, k))
    _lib.check(_lib.lib.vb_gemm(_lib.F32, 0, 0, m, n, k, 1.0, ptr(a), a.stride(0), ptr(b), b.stride(0), 0.0, ptr(c), n,
                                ptr(ws), ws.numel(), stream_ptr()), 'vb_gemm')
    return c


for _name, _fn in (('mf_step', _mf_step), ('mf_finalize', _mf_finalize), ('model_logp_grad', _model_logp_grad),
                   ('bounds_divergence', _bounds_divergence), ('psis_smooth', _psis_smooth),
                   ('psis_smooth_columns', _psis_smooth_columns), ('mf_adagrad', _mf_adagrad),
                   ('logsumexp', _logsumexp), ('gemm_tf32x3', _gemm_tf32x3), ('rhat_windows', _rhat_windows),
                   ('iterate_average', _iterate_average)):
    _lib_def.impl(_name, _fn, 'CUDA')

OPS = ('mf_step', 'mf_finalize', 'model_logp_grad', 'bounds_divergence', 'psis_smooth', 'psis_smooth_columns',
       'mf_adagrad', 'logsumexp', 'gemm_tf32x3', 'rhat_windows', 'iterate_average')
