"""Registered model plugins: ``logdensity(x) -> (n,)`` callables backed by CUDA kernels.

These take the place of the arbitrary autograd-traceable ``logdensity`` functions that the
reference passes to ``black_box_klvi`` / ``black_box_chivi`` (``viabel/vb.py:236,248``) and of
``make_stan_log_density`` (``viabel/vb.py:314-321``): each plugin supplies ``log p`` and
``grad log p`` on the unconstrained space.  Sources:

* ``Funnel``             ``notebooks/funnel-distribution.ipynb:48-54``
* ``NormalMixture``      ``notebooks/normal-mixture.ipynb:43``
* ``EightSchoolsNCP``    ``notebooks/eight_schools_ncp.stan:1-23`` (data ``notebooks/eight-schools.ipynb:173-174``)
* ``StudentTRegression`` ``notebooks/robust-regression.ipynb:73-86``
* ``CorrelatedGaussian`` BASELINE.json config 3 (no reference source)
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._device import WS, dtype_code, is_host, like_input, ptr, stream_ptr, to_dev, torch_dtype

__all__ = ['Model', 'Funnel', 'NormalMixture', 'EightSchoolsNCP', 'StudentTRegression',
           'CorrelatedGaussian', 'EIGHT_SCHOOLS_Y', 'EIGHT_SCHOOLS_SIGMA']

EIGHT_SCHOOLS_Y = (28., 8., -3., 7., -1., 1., 18., 12.)
EIGHT_SCHOOLS_SIGMA = (15., 10., 16., 11., 9., 11., 10., 18.)


class Model:
    """A model plugin handle.  ``model(x)`` returns ``log p`` like the reference's
    ``logdensity`` argument; ``model.logp_and_grad(x)`` also returns the gradient."""

    fused = False          # True when vb_mf_step evaluates this plugin inside the fused kernel
    name = None

    def __init__(self):
        self._handle = C.c_void_p()
        self._keep = []        # device tensors the C handle points into
        self._ws = None        # scratch of logp_and_grad_device: owned by the model, freed with it

    def __del__(self):
        h = getattr(self, '_handle', None)
        lib = getattr(_lib, 'lib', None) if _lib is not None else None     # module globals may be gone at exit
        if h and lib is not None:
            lib.vb_model_destroy(h)
            self._handle = None

    @property
    def handle(self):
        return self._handle

    @property
    def dim(self):
        return _lib.lib.vb_model_dim(self._handle)

    def logp_and_grad_device(self, x_dev, code, want_grad=True):
        """x_dev: CUDA (n, dim) tensor of dtype `code` -> (logp f64 (n,), grad (n, dim) or None)."""
        import torch
        n = x_dev.shape[0]
        logp = torch.empty(n, dtype=torch.float64, device=x_dev.device)
        grad = torch.empty_like(x_dev) if want_grad else None
        nbytes = _lib.lib.vb_model_logp_grad_workspace_bytes(self._handle, n, code)
        if self._ws is None:
            from ._device import Workspace
            self._ws = Workspace()
        ws = self._ws.get('logp_grad', nbytes)
        _lib.check(_lib.lib.vb_model_logp_grad(self._handle, code, ptr(x_dev), n, ptr(logp), ptr(grad),
                                               ptr(ws), ws.numel(), stream_ptr()), 'vb_model_logp_grad')
        return logp, grad

    def _prep(self, x, dtype):
        import torch
        if dtype is None and isinstance(x, torch.Tensor):
            dtype = x.dtype
        code = dtype_code(dtype)
        xd = to_dev(x, code)
        if xd.ndim == 1:
            xd = xd[None, :] if self.dim > 1 else xd[:, None]
        if xd.ndim != 2 or xd.shape[1] != self.dim:
            raise ValueError('expected samples of shape (n, %d)' % self.dim)
        return xd.contiguous(), code

    def logp_and_grad(self, x, dtype=None):
        xd, code = self._prep(x, dtype)
        lp, g = self.logp_and_grad_device(xd, code, True)
        return like_input(lp, x), like_input(g, x)

    def __call__(self, x, dtype=None):
        xd, code = self._prep(x, dtype)
        lp, _ = self.logp_and_grad_device(xd, code, False)
        return like_input(lp, x)


class Funnel(Model):
    name, fused = 'funnel', True

    def __init__(self, log_sigma_stdev=1.35):
        super().__init__()
        self.log_sigma_stdev = float(log_sigma_stdev)
        _lib.check(_lib.lib.vb_model_create_funnel(self.log_sigma_stdev, C.byref(self._handle)), 'funnel')


class NormalMixture(Model):
    name, fused = 'mixture', True

    def __init__(self, loc=2.0):
        super().__init__()
        self.loc = float(loc)
        _lib.check(_lib.lib.vb_model_create_mixture(self.loc, C.byref(self._handle)), 'mixture')


class EightSchoolsNCP(Model):
    name, fused = 'eight_schools', True

    def __init__(self, y=EIGHT_SCHOOLS_Y, sigma=EIGHT_SCHOOLS_SIGMA):
        super().__init__()
        self.y = np.ascontiguousarray(y, dtype=np.float64)
        self.sigma = np.ascontiguousarray(sigma, dtype=np.float64)
        if self.y.shape != self.sigma.shape or self.y.ndim != 1:
            raise ValueError('y and sigma must be 1-D arrays of equal length')
        dp = C.POINTER(C.c_double)
        _lib.check(_lib.lib.vb_model_create_eight_schools(
            self.y.ctypes.data_as(dp), self.sigma.ctypes.data_as(dp), len(self.y), C.byref(self._handle)),
            'eight_schools')


class StudentTRegression(Model):
    name = 'regression'

    def __init__(self, X, y, df=40.0, prior_sd=10.0, dtype='f64'):
        super().__init__()
        code = dtype_code(dtype)
        Xd, yd = to_dev(X, code), to_dev(y, code)
        if Xd.ndim != 2 or yd.ndim != 1 or yd.shape[0] != Xd.shape[0]:
            raise ValueError('X must be (N, d) and y (N,)')
        self._keep = [Xd, yd]
        self.df, self.prior_sd, self.data_dtype = float(df), float(prior_sd), code
        _lib.check(_lib.lib.vb_model_create_regression(ptr(Xd), ptr(yd), Xd.shape[0], Xd.shape[1], self.df,
                                                       self.prior_sd, code, C.byref(self._handle)), 'regression')
        if code == _lib.F32 and Xd.shape[1] % 4 == 0:
            # tensor path: X feeds tcgen05 as tf32 (hi, lo) planes; it is model data, so split it once here
            N, d = Xd.shape
            planes = torch.empty(_lib.lib.vb_split_bytes(N, d) // 4, dtype=torch.float32, device=Xd.device)
            _lib.check(_lib.lib.vb_split_f32(ptr(Xd), N, d, d, ptr(planes), stream_ptr()), 'vb_split_f32')
            _lib.check(_lib.lib.vb_model_regression_attach_split(self._handle, ptr(planes)), 'attach_split')
            self._keep.append(planes)
        if code == _lib.F32 and Xd.shape[1] <= 128:
            # fused kernel: X as scaled fp16 (hi, lo) planes, made once (model data)
            N, d = Xd.shape
            p16 = torch.empty(_lib.lib.vb_split_f16_bytes(N, d), dtype=torch.uint8, device=Xd.device)
            _lib.check(_lib.lib.vb_split_f16(ptr(Xd), N, d, d, ptr(p16), stream_ptr()), 'vb_split_f16')
            _lib.check(_lib.lib.vb_model_regression_attach_split_f16(self._handle, ptr(p16)), 'attach_split_f16')
            self._keep.append(p16)


class CorrelatedGaussian(Model):
    name = 'gauss'

    def __init__(self, mean, precision):
        super().__init__()
        md, Pd = to_dev(mean, _lib.F64), to_dev(precision, _lib.F64)
        if Pd.ndim != 2 or Pd.shape[0] != Pd.shape[1] or md.shape != (Pd.shape[0],):
            raise ValueError('mean must be (d,) and precision (d, d)')
        self._keep = [md, Pd]
        _lib.check(_lib.lib.vb_model_create_gauss(ptr(md), ptr(Pd), Pd.shape[0], C.byref(self._handle)), 'gauss')
        # constants of the fp32 tensor path (tf32 planes of P, bias mean' P): model data, made once
        import torch
        cache = torch.empty(int(_lib.lib.vb_model_gauss_cache_bytes(Pd.shape[0])), dtype=torch.uint8, device=Pd.device)
        _lib.check(_lib.lib.vb_model_gauss_attach_cache(self._handle, ptr(cache), stream_ptr()), 'gauss_attach_cache')
        self._keep.append(cache)
