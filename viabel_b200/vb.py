"""Drop-in for the Monte Carlo variational path of ``viabel/vb.py``.

Same names, argument order, defaults, return types and error behaviour as the reference for

* ``mean_field_gaussian_variational_family`` (``vb.py:48-82``)
* ``mean_field_t_variational_family``        (``vb.py:140-182``)
* ``t_variational_family``                   (``vb.py:192-233``, ``_distributions.py:8-38``)
* ``black_box_klvi`` / ``black_box_chivi``   (``vb.py:236-266``)
* ``adagrad_optimize`` / ``learning_rate_schedule`` (``vb.py:324-389``)

What differs by design: ``logdensity`` must be a registered model plugin
(``viabel_b200.models``) instead of an arbitrary autograd-traceable function, gradients are
closed-form CUDA reductions instead of an autograd tape, and base draws come from a Philox
counter stream (seed 0 per family, advancing once per ``seed=None`` call, mirroring the
reference's persistent ``RandomState(0)``) unless supplied through ``base_draws=``.
There is no CPU fallback.
"""
from collections import namedtuple
import ctypes
import math

import numpy as np
import torch
import tqdm

from . import _lib
from ._device import WS, Workspace, all_gather_rows, device, dtype_code, is_host, like_input, ptr, stream_ptr, to_dev, torch_dtype
from .models import Model

__all__ = [
    'VariationalFamily',
    'mean_field_gaussian_variational_family',
    'mean_field_t_variational_family',
    't_variational_family',
    'black_box_klvi',
    'black_box_klvi_pd',
    'black_box_klvi_pd2',
    'black_box_chivi',
    'adagrad_optimize',
    'rmsprop_IA_optimize_with_rhat',
    'adam_IA_optimize_with_rhat',
    'learning_rate_schedule',
    'multivariate_t_logpdf',
]

_Fields = namedtuple('VariationalFamily',
                     ['sample', 'entropy', 'logdensity', 'mean_and_cov', 'pth_moment', 'var_param_dim'])


class VariationalFamily(_Fields):
    """The reference's namedtuple (``vb.py:42-45``); ``impl`` carries what the fused objectives need."""
    impl = None


_LOG_2PI = math.log(2.0 * math.pi)


def _host_param(var_param):
    if isinstance(var_param, torch.Tensor):
        return var_param.detach().cpu().numpy().astype(np.float64)
    return np.asarray(var_param, dtype=np.float64)


class _FamilyImpl:
    """Host-side state of one family instance: kind, shape, compute dtype and the Philox stream."""

    def __init__(self, kind, dim, df, dtype):
        self.kind, self.dim, self.df = kind, int(dim), (0.0 if df is None else float(df))
        self.code = dtype_code(dtype)
        self.seed0, self.offset = 0, 0        # persistent stream, like RandomState(0) at vb.py:49,143,195
        self.ws = Workspace()

    def next_stream(self, seed):
        """(seed, call offset) for one draw; ``seed=None`` advances the family's own stream."""
        if seed is None:
            off = self.offset
            self.offset += 1
            return self.seed0, off
        return int(seed) & 0xFFFFFFFFFFFFFFFF, 0

    def base_draws(self, n, seed, offset, index0=0):
        """Philox base draws on the device: dict(eps=...) or dict(z=..., chi2=...)."""
        dt = torch_dtype(self.code)
        eps = torch.empty((n, self.dim), dtype=dt, device=device())
        chi2 = torch.empty(n, dtype=dt, device=device()) if self.kind == _lib.FAMILY_FR_T else None
        if self.kind == _lib.FAMILY_FR_T and self.code == _lib.F32:
            # fp32 storage: the draw kernel also writes the 3xTF32 operand planes of z that the weighted outer product
            # of the step consumes (no separate split pass over z)
            planes = torch.empty(_lib.lib.vb_split_bytes(n, self.dim) // 4, dtype=torch.float32, device=device())
            _lib.check(_lib.lib.vb_base_draws_fr_split(self.df, n, self.dim, seed, offset, index0, ptr(eps), ptr(chi2),
                                                       ptr(planes), stream_ptr()), 'vb_base_draws_fr_split')
            return {'z': eps, 'chi2': chi2, 'z_planes': planes}
        _lib.check(_lib.lib.vb_base_draws(self.kind, self.df, self.code, n, self.dim, seed, offset, index0,
                                          ptr(eps), ptr(chi2), stream_ptr()), 'vb_base_draws')
        if self.kind == _lib.FAMILY_FR_T:
            return {'z': eps, 'chi2': chi2}
        return {'eps': eps}

    def coerce_base(self, base):
        """External base draws (numpy or torch) -> device tensors of the compute dtype."""
        if self.kind == _lib.FAMILY_FR_T:
            z, c = to_dev(base['z'], self.code), to_dev(base['chi2'], self.code)
            if z.ndim != 2 or z.shape[1] != self.dim or c.shape != (z.shape[0],):
                raise ValueError('base draws must be z (n, dim) and chi2 (n,)')
            out = {'z': z, 'chi2': c}
            planes = base.get('z_planes') if isinstance(base, dict) else None
            if (planes is not None and self.code == _lib.F32 and isinstance(planes, torch.Tensor) and planes.is_cuda
                    and z is base['z'] and planes.numel() * 4 == _lib.lib.vb_split_bytes(z.shape[0], self.dim)):
                out['z_planes'] = planes                 # made by base_draws() together with this very z
            return out
        eps = base['eps'] if isinstance(base, dict) else base
        eps = to_dev(eps, self.code)
        if eps.ndim != 2 or eps.shape[1] != self.dim:
            raise ValueError('base draws must have shape (n, dim)')
        return {'eps': eps}


# =========================================================================== mean-field families
def _mean_field_family(dim, df, dtype):
    kind = _lib.FAMILY_MF_GAUSS if df is None else _lib.FAMILY_MF_T
    impl = _FamilyImpl(kind, dim, df, dtype)

    def unpack(var_param):
        lam = _host_param(var_param)
        return lam[:dim], lam[dim:]

    def sample(var_param, n_samples, seed=None):
        seed_, off = impl.next_stream(seed)
        base = impl.base_draws(int(n_samples), seed_, off)
        vp = to_dev(var_param, _lib.F64)
        x = torch.empty_like(base['eps'])
        _lib.check(_lib.lib.vb_mf_sample(impl.code, ptr(vp), int(n_samples), dim, ptr(base['eps']), ptr(x),
                                         stream_ptr()), 'vb_mf_sample')
        return like_input(x, var_param)

    def entropy(var_param):
        _, log_s = unpack(var_param)
        if df is None:
            return 0.5 * dim * (1.0 + _LOG_2PI) + np.sum(log_s)          # vb.py:61
        return np.sum(log_s)                                             # vb.py:156 (df terms dropped)

    def logdensity(x, var_param):
        host = is_host(x)
        xd = to_dev(x, impl.code if host else dtype_code(x.dtype))
        code = dtype_code(xd.dtype)
        one_d = xd.ndim == 1
        if one_d:
            xd = xd[None, :]
        if xd.ndim != 2 or xd.shape[1] != dim:
            raise ValueError('expected x of shape (n, %d)' % dim)
        xd = xd.contiguous()
        vp = to_dev(var_param, _lib.F64)
        out = torch.empty(xd.shape[0], dtype=torch.float64, device=xd.device)
        _lib.check(_lib.lib.vb_mf_logdensity(kind, impl.df, code, ptr(vp), xd.shape[0], dim, ptr(xd), ptr(out),
                                             stream_ptr()), 'vb_mf_logdensity')
        if one_d and df is None:          # mvn.logpdf of a single point is a scalar (vb.py:65)
            out = out[0]
        return like_input(out, x)

    def mean_and_cov(var_param):
        mean, log_s = unpack(var_param)
        c = 1.0 if df is None else df / (df - 2)                        # vb.py:69,166
        return mean, c * np.diag(np.exp(2 * log_s))

    def pth_moment(p, var_param):
        if p not in [2, 4]:
            raise ValueError('only p = 2 or 4 supported')
        if df is not None and df <= p:
            raise ValueError('df must be greater than p')
        _, log_s = unpack(var_param)
        s2 = np.exp(2 * log_s)
        if df is None:                                                   # vb.py:76-79
            return np.sum(s2) if p == 2 else 2 * np.sum(s2 ** 2) + np.sum(s2) ** 2
        c = df / (df - 2)                                                # vb.py:175-179
        if p == 2:
            return c * np.sum(s2)
        return c ** 2 * (2 * (df - 1) / (df - 4) * np.sum(s2 ** 2) + np.sum(s2) ** 2)

    fam = VariationalFamily(sample, entropy, logdensity, mean_and_cov, pth_moment, 2 * dim)
    fam.impl = impl
    return fam


def mean_field_gaussian_variational_family(dim, dtype='f64'):
    return _mean_field_family(dim, None, dtype)


def mean_field_t_variational_family(dim, df, dtype='f64'):
    if df <= 2:
        raise ValueError('df must be greater than 2')
    return _mean_field_family(dim, df, dtype)


# =========================================================================== full-rank Student t
class _FullRank(_FamilyImpl):
    """Parameter chain of paragami's free (mu, Sigma) pattern (vb.py:185-189):
    flat = [mu (d), row-major lower triangle of chol(Sigma) with log on the diagonal]."""

    def __init__(self, dim, df, dtype, scale):
        super().__init__(_lib.FAMILY_FR_T, dim, df, dtype)
        if scale not in ('sqrtm', 'chol'):
            raise ValueError("scale must be 'sqrtm' (reference) or 'chol'")
        self.scale = scale
        self.n_param = dim + dim * (dim + 1) // 2
        self._tri = None

    def tri(self):
        if self._tri is None or self._tri[0].device != device():
            idx = torch.tril_indices(self.dim, self.dim, device=device())
            diag_pos = torch.cumsum(torch.arange(1, self.dim + 1, device=device()), 0) - 1
            self._tri = (idx[0], idx[1], diag_pos)
        return self._tri

    def gemm(self, A, B, transa=False, transb=False):
        """fp64 d x d product through the library's own GEMM."""
        A, B = A.contiguous(), B.contiguous()      # eigh returns column-major eigenvectors
        m = A.shape[1] if transa else A.shape[0]
        k = A.shape[0] if transa else A.shape[1]
        n = B.shape[0] if transb else B.shape[1]
        Cm = torch.empty((m, n), dtype=torch.float64, device=A.device)
        _lib.check(_lib.lib.vb_gemm(_lib.F64, int(transa), int(transb), m, n, k, 1.0, ptr(A), A.shape[1],
                                    ptr(B), B.shape[1], 0.0, ptr(Cm), n, 0, 0, stream_ptr()), 'vb_gemm')
        return Cm

    def fold(self, vp):
        """vp: device f64 (P,) -> mu, L, Sigma."""
        d = self.dim
        r, c, _ = self.tri()
        L = torch.zeros((d, d), dtype=torch.float64, device=vp.device)
        L[r, c] = vp[d:]
        L.diagonal().exp_()
        return vp[:d].contiguous(), L, self.gemm(L, L, transb=True)

    def use_eigh(self):
        import os
        return os.environ.get('VB_FR_EIGH', '0') not in ('', '0')

    def chain_kernel(self, vp):
        """Forward half of the O(d^3) chain in ONE cooperative launch (csrc/fr_chain.cu): fold, then the symmetric
        square root as the polar factor of L by Newton-Schulz -- no eigendecomposition."""
        d = self.dim
        dev = vp.device
        L = torch.empty((d, d), dtype=torch.float64, device=dev)
        S = torch.empty((d, d), dtype=torch.float64, device=dev)
        S32 = torch.empty((d, d), dtype=torch.float32, device=dev) if self.code != _lib.F64 else None
        scal = torch.empty(4, dtype=torch.float64, device=dev)
        ws = self.ws.get('chain', _lib.lib.vb_fr_chain_workspace_bytes(d))
        _lib.check(_lib.lib.vb_fr_chain_forward(ptr(vp), d, 0 if self.scale == 'sqrtm' else 1, ptr(L), ptr(S), ptr(S32),
                                                ptr(scal), ptr(ws), ws.numel(), stream_ptr()), 'vb_fr_chain_forward')
        return {'mu': vp[:d], 'L': L, 'S': S, 'S32': S32, 'scal': scal, 'kernel': True}

    def chain_info(self):
        """[polar, inverse, smith iterations, status, forward GEMMs, backward GEMMs, symmetric GEMMs among them: forward,
        backward] of the last chain kernels (synchronises)."""
        out = (ctypes.c_int * 8)()
        ws = self.ws.get('chain', _lib.lib.vb_fr_chain_workspace_bytes(self.dim))
        _lib.check(_lib.lib.vb_fr_chain_info(ptr(ws), self.dim, out, stream_ptr()), 'vb_fr_chain_info')
        return list(out)

    def chain(self, vp):
        """Everything of O(d^3) that a step needs."""
        if not self.use_eigh():
            return self.chain_kernel(vp)
        mu, L, Sigma = self.fold(vp)
        out = {'mu': mu, 'L': L, 'Sigma': Sigma}
        if self.scale == 'sqrtm':
            # symmetric square root (scipy sqrtm in the reference, vb.py:207); the one library
            # call on this path: cuSOLVER syevd through torch
            lam, V = torch.linalg.eigh(Sigma)
            root = lam.sqrt()
            out.update(lam=lam, V=V, root=root, S=self.gemm(V * root, V, transb=True))
        else:
            out['S'] = L.T.contiguous()          # row form: x = mu + ztilde L'
        return out

    def scale_matrix(self, ch):
        S = ch['S']
        if self.code == _lib.F64:
            return S
        return ch['S32'] if ch.get('S32') is not None else S.to(torch.float32)

    def pack_grad(self, ch, G_S):
        """d/dphi of sum_i w_i log p(x_i) from G_S = Ztilde' diag(w) G  (SURVEY Appendix A.2)."""
        r, c, diag_pos = self.tri()
        L = ch['L']
        if self.scale == 'sqrtm':
            V, root = ch['V'], ch['root']
            inner = self.gemm(self.gemm(V, G_S, transa=True), V) / (root[:, None] + root[None, :])
            Y = self.gemm(self.gemm(V, inner), V, transb=True)          # Sylvester: S Y + Y S = G_S
            G_L = self.gemm(Y + Y.T, L)
        else:
            G_L = G_S.T
        phi = G_L[r, c].clone()
        phi[diag_pos] *= L.diagonal()
        return phi


def multivariate_t_logpdf(x, m, S, df=np.inf, _impl=None):
    """``viabel/_distributions.py:8-38``: log density of the multivariate Student t (any positive semidefinite
    ``S``: pseudo-inverse and pseudo-determinant through ``eigh``, eigenvalues with ``abs(s) <= 1e-10`` dropped,
    ``:27-31``).  ``x``: (n, d) or (d,) numpy / CUDA tensor; returns (n,) of the same kind.  ``df = inf`` is the
    multivariate normal (``:24-25``)."""
    host = is_host(x)
    code = (_impl.code if _impl is not None else _lib.F64) if host else dtype_code(x.dtype)
    xd = to_dev(x, code)
    if xd.ndim == 1:
        xd = xd[None, :]
    xd = xd.contiguous()
    n, d = xd.shape
    mu, Sigma = to_dev(m, _lib.F64).contiguous(), to_dev(S, _lib.F64).contiguous()
    s, u = torch.linalg.eigh(Sigma)                                      # _distributions.py:27 (off the step path)
    s_pinv = torch.where(s.abs() <= 1e-10, torch.zeros_like(s), 1.0 / s)
    U = (u * s_pinv.sqrt()).contiguous()
    log_pdet = float(torch.log(s).sum())
    out = torch.empty(n, dtype=torch.float64, device=xd.device)
    ws_owner = _impl.ws if _impl is not None else WS
    ws = ws_owner.get('logdensity', _lib.lib.vb_fr_logdensity_workspace_bytes(n, d))
    _lib.check(_lib.lib.vb_fr_logdensity(dtype_code(xd.dtype), ptr(mu), ptr(U), log_pdet, float(df), n, d, ptr(xd),
                                         ptr(out), ptr(ws), ws.numel(), stream_ptr()), 'vb_fr_logdensity')
    return like_input(out, x)


def t_variational_family(dim, df, dtype='f64', scale='sqrtm'):
    if df <= 2:
        raise ValueError('df must be greater than 2')
    impl = _FullRank(dim, df, dtype, scale)

    def sample_device(vp, base):
        ch = impl.chain(vp)
        n = base['z'].shape[0]
        x = torch.empty_like(base['z'])
        S = impl.scale_matrix(ch)
        ws = impl.ws.get('fr_sample', _lib.lib.vb_fr_sample_workspace_bytes(impl.code, n, dim))
        _lib.check(_lib.lib.vb_fr_sample(impl.code, ptr(ch['mu']), ptr(S), impl.df, n, dim, ptr(base['z']),
                                         ptr(base['chi2']), ptr(x), ptr(ws), ws.numel(), stream_ptr()), 'vb_fr_sample')
        return x, ch
    impl.sample_device = sample_device

    def sample(var_param, n_samples, seed=None):
        seed_, off = impl.next_stream(seed)
        base = impl.base_draws(int(n_samples), seed_, off)
        x, _ = sample_device(to_dev(var_param, _lib.F64), base)
        return like_input(x, var_param)

    def entropy(var_param):
        lam = _host_param(var_param)
        diag_pos = np.cumsum(np.arange(1, dim + 1)) - 1
        return np.sum(lam[dim:][diag_pos])               # 1/2 log det Sigma, vb.py:213

    def _fold_host(var_param):
        lam = _host_param(var_param)
        L = np.zeros((dim, dim))
        L[np.tril_indices(dim)] = lam[dim:]
        L[np.diag_indices(dim)] = np.exp(np.diag(L))
        return lam[:dim], L @ L.T

    def logdensity(x, var_param):
        vp = to_dev(var_param, _lib.F64)
        mu, _, Sigma = impl.fold(vp)
        return multivariate_t_logpdf(x, mu, Sigma, df, _impl=impl)

    def mean_and_cov(var_param):
        mu, Sigma = _fold_host(var_param)
        return mu, df / (df - 2.) * Sigma                                # vb.py:217

    def pth_moment(p, var_param):
        if p not in [2, 4]:
            raise ValueError('only p = 2 or 4 supported')
        if df <= p:
            raise ValueError('df must be greater than p')
        _, Sigma = _fold_host(var_param)
        ev = np.linalg.eigvalsh(Sigma)                                   # vb.py:225
        c = df / (df - 2)
        if p == 2:
            return c * np.sum(ev)
        return c ** 2 * (2 * (df - 1) / (df - 4) * np.sum(ev ** 2) + np.sum(ev) ** 2)

    fam = VariationalFamily(sample, entropy, logdensity, mean_and_cov, pth_moment, impl.n_param)
    fam.impl = impl
    return fam


# =========================================================================== objectives
def _require_plugin(logdensity):
    if not isinstance(logdensity, Model):
        raise TypeError('logdensity must be a registered viabel_b200.models plugin (Funnel, NormalMixture, '
                        'EightSchoolsNCP, StudentTRegression, CorrelatedGaussian); arbitrary Python '
                        'callables have no CUDA path and there is no CPU fallback')
    return logdensity


def _dist_info(group):
    """(rank, world, group) of the sample-sharding group; (0, 1, None) when not sharded.  Sharding is opt-in, as in
    bounds.py and psis.py: ``group=None`` / ``False`` never shards, ``True`` is the world group."""
    import torch.distributed as dist
    if group is None or group is False or not dist.is_available() or not dist.is_initialized():
        return 0, 1, None
    g = None if group is True else group
    return dist.get_rank(g), dist.get_world_size(g), g


def _shard(n, rank, world):
    """Contiguous split of n samples: (first global index, count) of this rank."""
    base, rem = divmod(n, world)
    count = base + (1 if rank < rem else 0)
    start = rank * base + min(rank, rem)
    return start, count


# --------------------------------------------------------------------------- full-rank multi-rank exchange
# (CHIVI row).  Device-agnostic torch code: the world-2 gloo test drives these two functions with CPU
# tensors, the NCCL check (scripts/check_multi_gpu.py) with the real kernels around them.
def fr_exchange_max(rows, w, alpha, group):
    """rows = [local max lw, sum w (relative to the local max), sum log p, sum lw]; w (n_local,) relative to the
    local max.  All-reduce the maximum and rescale to the GLOBAL one: s_i = exp(alpha (lw_i - max)) (vb.py:259-260)."""
    import torch.distributed as dist
    gm = rows[0:1].clone()
    dist.all_reduce(gm, op=dist.ReduceOp.MAX, group=group)
    scale = torch.exp(alpha * (rows[0] - gm[0]))
    return w * scale, torch.stack([gm[0], rows[1] * scale, rows[2], rows[3]])


def fr_exchange_sums(rows, g_mu, G_S, world, group, n_local=None):
    """One all-gather of [sum w, sum log p, sum lw, local sample count, g_mu (d), G_S (d*d)], summed over the ranks in
    one fixed-order reduction (every rank computes the same bits).  Returns rows (5,) = [max lw, sum w, sum log p,
    sum lw, global sample count], g_mu, G_S.  The count rides along so that uneven external shards need no separate
    collective and no host read."""
    d = g_mu.numel()
    cnt = torch.full((1,), float(n_local if n_local is not None else 0), dtype=torch.float64, device=rows.device)
    pack = torch.cat([rows[1:4], cnt, g_mu, G_S.reshape(-1)])
    tot = all_gather_rows(pack, world, group).sum(dim=0)
    return torch.cat([rows[0:1], tot[0:4]]), tot[4:4 + d], tot[4 + d:].view(d, d)


class _Objective:
    """One KLVI or CHIVI objective bound to (family, model, n_samples)."""

    def __init__(self, objective, alpha, var_family, model, n_samples, group):
        impl = getattr(var_family, 'impl', None)
        if not isinstance(impl, _FamilyImpl):
            raise TypeError('var_family must come from viabel_b200.vb')
        self.obj, self.alpha = objective, float(alpha)
        self.impl, self.model, self.n = impl, _require_plugin(model), int(n_samples)
        if self.model.dim != impl.dim:
            raise ValueError('model dimension %d != family dimension %d' % (self.model.dim, impl.dim))
        self.group = group
        self.P = var_family.var_param_dim
        self.ws = Workspace()
        self.last_partials = None

    # ------------------------------------------------------------------ mean field
    def _mf(self, vp, base, seed, off, index0, n_local, finalize):
        impl, d = self.impl, self.impl.dim
        NP = 2 * d + 4
        out = torch.empty(1 + 2 * d, dtype=torch.float64, device=vp.device)
        partials = torch.empty(NP, dtype=torch.float64, device=vp.device)
        val_ptr = ptr(out) if finalize else 0
        grad_ptr = ptr(out) + 8 if finalize else 0
        if self.model.fused:
            ws = self.ws.get('mf_step', _lib.lib.vb_mf_step_workspace_bytes(d))
            _lib.check(_lib.lib.vb_mf_step(
                impl.kind, impl.df, self.model.handle, self.obj, self.alpha, impl.code, ptr(vp), n_local,
                ptr(base['eps']) if base is not None else 0, seed, off, index0,
                val_ptr, grad_ptr, ptr(partials), ptr(ws), ws.numel(), stream_ptr()), 'vb_mf_step')
        else:
            if base is None:
                base = impl.base_draws(n_local, seed, off, index0)
            x = torch.empty_like(base['eps'])
            _lib.check(_lib.lib.vb_mf_sample(impl.code, ptr(vp), n_local, d, ptr(base['eps']), ptr(x),
                                             stream_ptr()), 'vb_mf_sample')
            logp, g = self.model.logp_and_grad_device(x, impl.code, True)
            ws = self.ws.get('mf_reduce', _lib.lib.vb_mf_reduce_workspace_bytes(n_local, d))
            _lib.check(_lib.lib.vb_mf_reduce(
                impl.kind, impl.df, self.obj, self.alpha, impl.code, ptr(vp), n_local, d, ptr(base['eps']),
                ptr(logp), ptr(g), val_ptr, grad_ptr, ptr(partials), ptr(ws), ws.numel(), stream_ptr()),
                'vb_mf_reduce')
        return out, partials

    def _mf_merge(self, vp, partials, world, g):
        import torch.distributed as dist
        d = self.impl.dim
        gathered = all_gather_rows(partials, world, g)
        out = torch.empty(1 + 2 * d, dtype=torch.float64, device=vp.device)
        _lib.check(_lib.lib.vb_mf_finalize(self.impl.kind, self.obj, self.alpha, ptr(vp), d, ptr(gathered), world,
                                           ptr(out), ptr(out) + 8, stream_ptr()), 'vb_mf_finalize')
        return out

    # ------------------------------------------------------------------ full rank
    def _fr(self, vp, base, n_total, rank, world, g, entropy_host=None):
        import torch.distributed as dist
        impl, d = self.impl, self.impl.dim
        n = base['z'].shape[0]
        x, ch = impl.sample_device(vp, base)
        logp, grad = self.model.logp_and_grad_device(x, impl.code, True)
        const = (math.lgamma(0.5 * (impl.df + d)) - math.lgamma(0.5 * impl.df)
                 - 0.5 * d * math.log(math.pi * impl.df))
        if ch.get('kernel'):
            # sum of the log-diagonal = 1/2 log det Sigma (vb.py:213): from the host copy of var_param when there is one
            # (no device read-back), else from the forward kernel's scalar
            entropy = entropy_host if entropy_host is not None else ch['scal'][0]
        else:
            entropy = ch['L'].diagonal().log().sum()
        lw = torch.empty(n, dtype=torch.float64, device=vp.device)
        w = torch.empty(n, dtype=torch.float64, device=vp.device)
        rows = torch.empty(4, dtype=torch.float64, device=vp.device)
        ws = self.ws.get('fr_rows', _lib.lib.vb_fr_rows_workspace_bytes(n))
        # path-derivative KLVI: unit weights like KLVI; its value uses the sum of lw that vb_fr_rows also returns
        rows_obj = _lib.OBJ_CHIVI if self.obj == _lib.OBJ_CHIVI else _lib.OBJ_KLVI
        _lib.check(_lib.lib.vb_fr_rows(rows_obj, self.alpha, impl.df, impl.code, n, d, ptr(base['z']),
                                       ptr(base['chi2']), ptr(logp), float(const) - float(entropy),
                                       ptr(lw), ptr(w), ptr(rows), ptr(ws), ws.numel(), stream_ptr()), 'vb_fr_rows')
        if world > 1 and self.obj == _lib.OBJ_CHIVI:
            w, rows = fr_exchange_max(rows, w, self.alpha, g)
        g_mu = torch.empty(d, dtype=torch.float64, device=vp.device)
        G_S = torch.empty((d, d), dtype=torch.float64, device=vp.device)
        ws2 = self.ws.get('fr_weighted', _lib.lib.vb_fr_weighted_workspace_bytes(n, d))
        if base.get('z_planes') is not None:
            _lib.check(_lib.lib.vb_fr_weighted_split(impl.df, n, d, ptr(base['z']), ptr(base['z_planes']), ptr(base['chi2']),
                                                     ptr(w), ptr(grad), ptr(g_mu), ptr(G_S), ptr(ws2), ws2.numel(),
                                                     stream_ptr()), 'vb_fr_weighted_split')
        else:
            _lib.check(_lib.lib.vb_fr_weighted(impl.code, impl.df, n, d, ptr(base['z']), ptr(base['chi2']), ptr(w),
                                               ptr(grad), ptr(g_mu), ptr(G_S), ptr(ws2), ws2.numel(), stream_ptr()),
                       'vb_fr_weighted')
        if world > 1:
            rows, g_mu, G_S = fr_exchange_sums(rows, g_mu, G_S, world, g, n)
        self.last_log_weights = lw
        if ch.get('kernel'):
            # backward half of the chain + assembly of (value, gradient) in ONE cooperative launch
            out = torch.empty(1 + self.P, dtype=torch.float64, device=vp.device)
            cws = impl.ws.get('chain', _lib.lib.vb_fr_chain_workspace_bytes(d))
            # n_total None: the global count is rows[4] on the device (0.0 tells the kernel to read it there)
            _lib.check(_lib.lib.vb_fr_chain_backward(d, 0 if impl.scale == 'sqrtm' else 1, self.obj, self.alpha,
                                                     0.0 if n_total is None else float(n_total), ptr(ch['L']),
                                                     ptr(ch['S']), ptr(G_S.contiguous()), ptr(g_mu.contiguous()),
                                                     ptr(rows.contiguous()), ptr(ch['scal']), ptr(out), ptr(cws),
                                                     cws.numel(), stream_ptr()), 'vb_fr_chain_backward')
            return out
        phi = impl.pack_grad(ch, G_S)
        _, _, diag_pos = impl.tri()
        out = torch.empty(1 + self.P, dtype=torch.float64, device=vp.device)
        nt = float(n_total)
        if self.obj != _lib.OBJ_CHIVI:
            # KLVI: analytic entropy (vb.py:240); path-derivative form: -(mean log p - mean log q) (vb.py:272)
            out[0] = -(entropy + rows[2] / nt) if self.obj == _lib.OBJ_KLVI else -(rows[3] / nt)
            out[1:1 + d] = -g_mu / nt
            phi = -phi / nt
            phi[diag_pos] -= 1.0
            out[1 + d:] = phi
        else:
            out[0] = torch.log(rows[1] / nt) / self.alpha + rows[0]
            out[1:1 + d] = self.alpha * g_mu / nt
            phi[diag_pos] += rows[1]
            out[1 + d:] = self.alpha * phi / nt
        self.last_log_weights = lw
        return out

    # ------------------------------------------------------------------ call
    def __call__(self, var_param, base_draws=None):
        vp = to_dev(var_param, _lib.F64)
        if vp.shape != (self.P,):
            raise ValueError('var_param must have shape (%d,)' % self.P)
        rank, world, g = _dist_info(self.group)
        impl = self.impl
        # base-draw stream: KLVI advances the family's persistent stream (vb.py:239); CHIVI takes a
        # fresh seed from numpy's global RNG on every call (vb.py:258)
        if base_draws is not None:
            base = impl.coerce_base(base_draws)
            n_local = (base['eps'] if 'eps' in base else base['z']).shape[0]
            seed = off = index0 = 0
            n_total = n_local
            if world > 1 and impl.kind == _lib.FAMILY_FR_T and not impl.use_eigh():
                # external shards may be uneven: the global sample count is the sum of the local ones.  It travels
                # with the step's one exchange of sums and stays on the device (no extra collective, no host read)
                n_total = None
            elif world > 1:
                import torch.distributed as dist
                cnt = torch.tensor([n_local], dtype=torch.int64, device=vp.device)
                dist.all_reduce(cnt, group=g)
                n_total = int(cnt.item())
        else:
            base = None
            if self.obj != _lib.OBJ_CHIVI:
                seed, off = impl.next_stream(None)
            else:
                seed, off = int(np.random.randint(2 ** 32)), 0          # every rank consumes one draw per call (vb.py:258)
                if world > 1:
                    # the ranks need ONE seed per call.  Rank 0's first draw is broadcast once per objective (one host
                    # read, first call only); call k then uses base + k * golden-ratio increment (mod 2^64), so the
                    # steady state has no collective and no host read for the seed.  Every rank must make the same
                    # sequence of calls -- which the step's own collectives require anyway.
                    if getattr(self, '_shard_seed', None) is None:
                        import torch.distributed as dist
                        st = torch.tensor([seed], dtype=torch.int64, device=vp.device)
                        dist.broadcast(st, src=dist.get_global_rank(g, 0) if g is not None else 0, group=g)
                        self._shard_seed, self._shard_calls = int(st.item()), 0
                    seed = (self._shard_seed + 0x9E3779B97F4A7C15 * self._shard_calls) & 0xFFFFFFFFFFFFFFFF
                    self._shard_calls += 1
            index0, n_local = _shard(self.n, rank, world)
            n_total = self.n
        if impl.kind == _lib.FAMILY_FR_T:
            if base is None:
                base = impl.base_draws(n_local, seed, off, index0)
            entropy_host = None
            if is_host(var_param):
                lam_h = _host_param(var_param)
                entropy_host = float(np.sum(lam_h[impl.dim:][np.cumsum(np.arange(1, impl.dim + 1)) - 1]))
            out = self._fr(vp, base, n_total, rank, world, g, entropy_host)
        else:
            out, partials = self._mf(vp, base, seed, off, index0, n_local, finalize=(world == 1))
            self.last_partials = partials
            if world > 1:
                out = self._mf_merge(vp, partials, world, g)
        if is_host(var_param):
            h = out.cpu().numpy()
            return h[0], h[1:].copy()
        return out[0], out[1:]


def black_box_klvi(var_family, logdensity, n_samples, group=None):
    """``vb.py:236-245``.  Returns ``objective_and_grad(var_param) -> (value, grad)`` with
    value ``-(entropy + mean log p)``.  ``group``: torch.distributed group to shard the sample
    axis over (``True``: the world group; default ``None`` / ``False``: not sharded, every rank runs the whole
    objective -- the same convention as ``bounds`` and ``psis``)."""
    return _Objective(_lib.OBJ_KLVI, 1.0, var_family, logdensity, n_samples, group)


def black_box_klvi_pd(var_family, logdensity, n_samples, group=None):
    """``vb.py:268-278``: KLVI whose value is ``-(mean log p(x) - mean log q(x; var_param))`` on the drawn samples
    (a Monte Carlo entropy instead of the analytic one).  Differentiated as the reference does (through the samples
    AND the explicit parameter) the gradient is KLVI's: along a reparameterised sample ``log q`` depends on the
    parameter only through ``-entropy``.  One fused launch: the CHIVI pass with unit weights."""
    return _Objective(_lib.OBJ_KLVI_PD, 1.0, var_family, logdensity, n_samples, group)


def black_box_klvi_pd2(var_family, logdensity, n_samples, group=None):
    """``vb.py:281-296``: the same function as ``black_box_klvi_pd`` written with ``functools.partial``; autograd
    still traces ``var_param`` through the partial, so value and gradient are identical."""
    return _Objective(_lib.OBJ_KLVI_PD, 1.0, var_family, logdensity, n_samples, group)


def black_box_chivi(alpha, var_family, logdensity, n_samples, group=None):
    """``vb.py:248-266``.  Returns a callable giving ``(CUBO_alpha estimate, gradient)`` — a
    2-tuple, and the gradient is ``(alpha/n) sum_i s_i grad lw_i`` with
    ``s_i = exp(alpha (lw_i - max lw))`` (not normalised by ``mean s``), as in the reference."""
    return _Objective(_lib.OBJ_CHIVI, alpha, var_family, logdensity, n_samples, group)


# =========================================================================== optimiser
def learning_rate_schedule(n_iters, learning_rate, learning_rate_end):
    """``vb.py:324-343``: constant, then a/(b+i) over the middle half, then constant."""
    if learning_rate <= 0:
        raise ValueError('learning rate must be positive')
    decay = learning_rate_end is not None
    if decay:
        if learning_rate <= learning_rate_end:
            raise ValueError('initial learning rate must be greater than final learning rate')
        b = n_iters * learning_rate_end / (2 * (learning_rate - learning_rate_end))
        a = learning_rate * b
        first, last = n_iters // 4, 3 * n_iters // 4
    for i in range(n_iters):
        if not decay or i < first:
            yield learning_rate
        elif i < last:
            yield a / (b + i - first + 1)
        else:
            yield learning_rate_end


def _device_loop_ok(objective_and_grad, has_log_norm):
    """The device-resident loop serves single-rank objectives of the fused mean-field path."""
    o = objective_and_grad
    if has_log_norm or not isinstance(o, _Objective):
        return False
    if o.impl.kind == _lib.FAMILY_FR_T or not o.model.fused:
        return False
    return _dist_info(o.group)[1] == 1


def _adagrad_device(n_iters, o, init_param, window, learning_rate, epsilon, learning_rate_end, base_draws,
                    graph_iters, progress, fused=True):
    """``vb.py:345-389`` with the loop on the device (``vb_mf_adagrad_*``): one fused step launch and one
    update launch per iteration, replayed in CUDA graphs of ``graph_iters`` iterations; the host only reads
    the value history back between groups of iterations to refresh the progress bar."""
    impl, d = o.impl, o.impl.dim
    P = 2 * d
    rates = np.fromiter(learning_rate_schedule(n_iters, learning_rate, learning_rate_end), dtype=np.float64,
                        count=n_iters)
    if base_draws is not None:
        base = impl.coerce_base(base_draws.reshape(-1, d) if hasattr(base_draws, 'reshape') else base_draws)['eps']
        if base.shape[0] % n_iters != 0 or base.shape[0] // n_iters < 1:
            raise ValueError('base_draws must have shape (n_iters, n_samples, dim)')
        n_samples = base.shape[0] // n_iters
        seeds, offsets = np.zeros(n_iters, dtype=np.uint64), np.zeros(n_iters, dtype=np.uint32)
    else:
        base, n_samples = None, o.n
        if o.obj != _lib.OBJ_CHIVI:                     # the family's persistent stream, one call offset per step
            seeds = np.full(n_iters, impl.seed0, dtype=np.uint64)
            offsets = (impl.offset + np.arange(n_iters)).astype(np.uint32)
            impl.offset += n_iters
        else:                                           # CHIVI: a fresh global-RNG seed per call (vb.py:258)
            seeds = np.array([np.random.randint(2 ** 32) for _ in range(n_iters)], dtype=np.uint64)
            offsets = np.zeros(n_iters, dtype=np.uint32)
    dev = device()
    param = torch.from_numpy(np.array(init_param, dtype=np.float64, copy=True)).to(dev)
    if param.shape != (P,):
        raise ValueError('init_param must have shape (%d,)' % P)
    keep_from = 3 * n_iters // 4
    d_rates = torch.from_numpy(rates).to(dev)
    d_seeds = torch.from_numpy(seeds.view(np.int64)).to(dev)
    d_offsets = torch.from_numpy(offsets.view(np.int32)).to(dev)
    values = torch.zeros(n_iters, dtype=torch.float64, device=dev)
    kept = torch.zeros((n_iters - keep_from, P), dtype=torch.float64, device=dev)
    state = torch.zeros(_lib.lib.vb_mf_adagrad_state_bytes(d, window), dtype=torch.uint8, device=dev)
    ws = torch.zeros(_lib.lib.vb_mf_step_workspace_bytes(d), dtype=torch.uint8, device=dev)
    plan = ctypes.c_void_p()
    _lib.check(_lib.lib.vb_mf_adagrad_create(
        impl.kind, impl.df, o.model.handle, o.obj, o.alpha, impl.code, ptr(param), n_samples, ptr(base), n_iters,
        window, float(epsilon), ptr(d_rates), ptr(d_seeds), ptr(d_offsets), ptr(values), ptr(kept), ptr(state),
        state.numel(), ptr(ws), ws.numel(), int(graph_iters), int(bool(fused)), stream_ptr(), ctypes.byref(plan)),
        'vb_mf_adagrad_create')
    try:
        group = max(int(graph_iters), 1) * max(1, 500 // max(int(graph_iters), 1))
        done = 0
        while done < n_iters:
            k = min(group, n_iters - done)
            _lib.check(_lib.lib.vb_mf_adagrad_run(plan, k, stream_ptr()), 'vb_mf_adagrad_run')
            done += k
            if progress is not None:
                i = (done - 1) // 10 * 10                # last iteration of this group that refreshes the bar
                host = values[max(0, i - 1000):i + 1].cpu().numpy()
                progress.update(k)
                progress.set_description('Average Loss = {:,.5g}'.format(np.mean(host)))
        kept_host = kept.cpu().numpy()
        values_host = values.cpu().numpy()
    finally:
        torch.cuda.current_stream().synchronize()
        _lib.lib.vb_mf_adagrad_destroy(plan)
    return np.mean(kept_host, axis=0), kept_host, values_host, np.array([0] * n_iters)


def adagrad_optimize(n_iters, objective_and_grad, init_param,
                     has_log_norm=False, window=10, learning_rate=.01,
                     epsilon=.1, learning_rate_end=None, device_loop=None, base_draws=None, graph_iters=None):
    """Windowed AdaGrad with iterate averaging over the last quarter (``vb.py:345-389``).

    ``device_loop`` (extension): ``None`` runs the whole loop on the device when the objective is a single-rank
    fused mean-field objective of this package (same iterates as the host loop, bit for bit, because both consume
    the same Philox streams and the update is evaluated in the same order), ``False`` forces the host loop of the
    reference, ``True`` raises if the device loop does not apply; ``'two-launch'`` is ``True`` with the update in its
    own kernel instead of the step kernel's last CTA (ablation).  ``base_draws`` (n_iters, n_samples, dim):
    external base draws, one block per iteration (validation against the CPU oracle).  ``graph_iters``:
    iterations per CUDA-graph replay (default 50; 0 = plain launches)."""
    if device_loop is None:
        device_loop = _device_loop_ok(objective_and_grad, has_log_norm)
    elif device_loop and not _device_loop_ok(objective_and_grad, has_log_norm):
        raise TypeError('device_loop=True needs a single-rank fused mean-field objective of viabel_b200.vb and '
                        'has_log_norm=False')
    if device_loop:
        # learning-rate validation errors surface before the progress bar opens, as in the reference (vb.py:356)
        if graph_iters is None:
            graph_iters = 50 if n_iters >= 100 else 0
        with tqdm.trange(n_iters) as progress:
            try:
                return _adagrad_device(n_iters, objective_and_grad, init_param, window, learning_rate, epsilon,
                                       learning_rate_end, base_draws, graph_iters, progress,
                                       fused=device_loop != 'two-launch')
            finally:
                progress.close()
    if base_draws is not None:
        base_draws = np.asarray(base_draws)
    recent_grads, recent_norms = [], []
    values, norms, kept = [], [], []
    param = np.array(init_param, dtype=float, copy=True)
    keep_from = 3 * n_iters // 4
    with tqdm.trange(n_iters) as progress:
        try:
            rates = learning_rate_schedule(n_iters, learning_rate, learning_rate_end)
            for i, rate in zip(progress, rates):
                extra = {} if base_draws is None else {'base_draws': base_draws[i]}
                if has_log_norm:
                    value, grad, log_norm = objective_and_grad(param, **extra)
                else:
                    value, grad = objective_and_grad(param, **extra)
                    log_norm = 0
                values.append(value)
                norms.append(log_norm)
                recent_grads.append(grad)
                recent_norms.append(log_norm)
                if len(recent_grads) > window:                       # pop after append, vb.py:368-370
                    del recent_grads[0], recent_norms[0]
                rescale = np.exp(np.min(recent_norms) - np.array(recent_norms))
                accum = np.sum((rescale[:, None] * np.array(recent_grads)) ** 2, axis=0)
                param = param - rate * grad / np.sqrt(epsilon + accum)  # epsilon inside the sqrt
                if i >= keep_from:
                    kept.append(param.copy())                        # post-update iterate
                if i % 10 == 0:
                    avg_loss = np.mean(values[max(0, i - 1000):i + 1])
                    progress.set_description('Average Loss = {:,.5g}'.format(avg_loss))
        except (KeyboardInterrupt, StopIteration):  # pragma: no cover
            progress.close()
        finally:
            progress.close()
    kept = np.array(kept)
    return np.mean(kept, axis=0), kept, np.array(values), np.array(norms)


# --------------------------------------------------------------------------- R-hat-monitored optimisers (fork)
def _first_converged_window(rhat_windows, threshold, rhat_window, default):
    """First window that, together with its successor, has every R-hat below the threshold (``vb.py:499-515``)."""
    for ee in range(rhat_windows.shape[0] - 1):
        if (rhat_windows[ee] < threshold).all() and (rhat_windows[ee + 1] < threshold).all():
            return ee * rhat_window
    return default


def _restart_device(rule, n_iters, o, param0, learning_rate, epsilon, learning_rate_end, progress, graph_iters=None):
    """One restart of the RMSProp / Adam loop on the device (``vb_mf_optimizer_create``): step, finalise and update
    in one launch per iteration (or many iterations per launch for n <= 2048 samples), moments / iterate / histories
    in HBM.  Returns (final param, pre-update iterates (n_iters, P), values (n_iters,)), bit-identical to the host
    loop over the same objective: same Philox streams, every operation of the update individually rounded in
    numpy's order, bias corrections ``1 - beta**(i + 2)`` tabulated by numpy itself."""
    impl, d = o.impl, o.impl.dim
    P = 2 * d
    rates = np.fromiter(learning_rate_schedule(n_iters, learning_rate, learning_rate_end), dtype=np.float64,
                        count=n_iters)
    if o.obj != _lib.OBJ_CHIVI:
        seeds = np.full(n_iters, impl.seed0, dtype=np.uint64)
        offsets = (impl.offset + np.arange(n_iters)).astype(np.uint32)
        impl.offset += n_iters
    else:
        seeds = np.array([np.random.randint(2 ** 32) for _ in range(n_iters)], dtype=np.uint64)
        offsets = np.zeros(n_iters, dtype=np.uint32)
    if graph_iters is None:
        graph_iters = 50 if n_iters >= 100 else 0
    dev = device()
    steps = np.arange(n_iters)
    d_bias1 = torch.from_numpy(1 - np.power(0.9, steps + 2)).to(dev)
    d_bias2 = torch.from_numpy(1 - np.power(0.999, steps + 2)).to(dev)
    param = torch.from_numpy(np.array(param0, dtype=np.float64, copy=True)).to(dev)
    d_rates = torch.from_numpy(rates).to(dev)
    d_seeds = torch.from_numpy(seeds.view(np.int64)).to(dev)
    d_offsets = torch.from_numpy(offsets.view(np.int32)).to(dev)
    values = torch.zeros(n_iters, dtype=torch.float64, device=dev)
    history = torch.zeros((n_iters, P), dtype=torch.float64, device=dev)
    state = torch.zeros(_lib.lib.vb_mf_adagrad_state_bytes(d, 2), dtype=torch.uint8, device=dev)
    ws = torch.zeros(_lib.lib.vb_mf_step_workspace_bytes(d), dtype=torch.uint8, device=dev)
    plan = ctypes.c_void_p()
    _lib.check(_lib.lib.vb_mf_optimizer_create(
        _lib.OPT_RMSPROP if rule == 'rmsprop' else _lib.OPT_ADAM, impl.kind, impl.df, o.model.handle, o.obj, o.alpha,
        impl.code, ptr(param), o.n, 0, n_iters, float(epsilon), ptr(d_rates), ptr(d_seeds), ptr(d_offsets),
        ptr(d_bias1), ptr(d_bias2), ptr(values), ptr(history), ptr(state), state.numel(), ptr(ws), ws.numel(),
        int(graph_iters), 1, stream_ptr(), ctypes.byref(plan)), 'vb_mf_optimizer_create')
    try:
        group = max(int(graph_iters), 1) * max(1, 500 // max(int(graph_iters), 1))
        done = 0
        while done < n_iters:
            k = min(group, n_iters - done)
            _lib.check(_lib.lib.vb_mf_adagrad_run(plan, k, stream_ptr()), 'vb_mf_adagrad_run')
            done += k
            if progress is not None:
                i = (done - 1) // 100 * 100
                progress.update(k)
                progress.set_description('Average Loss = {:,.6g}'.format(
                    np.mean(values[max(0, i - 1000):i + 1].cpu().numpy())))
        return param.cpu().numpy(), history.cpu().numpy(), values.cpu().numpy()
    finally:
        torch.cuda.current_stream().synchronize()
        _lib.lib.vb_mf_adagrad_destroy(plan)


def _optimize_with_rhat(rule, n_iters, objective_and_grad, init_param, K, has_log_norm, window, learning_rate,
                        epsilon, rhat_window, n_optimisers, r_mean_threshold, r_sigma_threshold, tail_avg_iters,
                        avg_grad_norm, learning_rate_end, device_loop=None):
    """Shared body of ``rmsprop_IA_optimize_with_rhat`` (``vb.py:392-548``) and ``adam_IA_optimize_with_rhat``
    (``vb.py:554-712``): n_optimisers restarts of the update rule, the iterate history of every restart, then the
    windowed / halfway split R-hat of that history and the iterate averages from the first converged window on.
    The objective is this package's CUDA step; the diagnostics run on the device over the uploaded history
    (``functions.py``).  The reference's windowed-gradient bookkeeping (``vb.py:456-459``) feeds nothing and is
    not restated."""
    import scipy.stats as stats
    from . import functions as fn
    init_param = np.array(init_param, dtype=float, copy=True)
    param = init_param.copy()
    values, norms, chains = [], [], []
    decay, beta1, beta2 = 0.9, 0.9, 0.999
    spread = 0.5 if rule == 'rmsprop' else 0.2                       # vb.py:425, :592
    on_device = _device_loop_ok(objective_and_grad, has_log_norm) and not avg_grad_norm
    if device_loop is None:
        device_loop = on_device
    elif device_loop and not on_device:
        raise TypeError('device_loop=True needs a single-rank fused mean-field objective of viabel_b200.vb, '
                        'has_log_norm=False and avg_grad_norm=False')
    for o in range(n_optimisers):
        history = []
        np.random.seed(seed=o)                                       # the reference reseeds the GLOBAL numpy stream
        if o >= 1:
            param = init_param + stats.norm.rvs(size=len(init_param)) * (o + 1) * spread
        if device_loop:
            learning_rate_schedule(1, learning_rate, learning_rate_end).__next__()   # rate errors before the bar opens
            with tqdm.trange(n_iters) as progress:
                try:
                    param, hist, vals = _restart_device(rule, n_iters, objective_and_grad, param, learning_rate,
                                                        epsilon, learning_rate_end, progress)
                finally:
                    progress.close()
            values.extend(vals.tolist())
            norms.extend([0] * n_iters)
            chains.append(hist[-100 * window:] if n_iters > 100 * window else hist)
            continue
        with tqdm.trange(n_iters) as progress:
            try:
                rates = learning_rate_schedule(n_iters, learning_rate, learning_rate_end)
                for i, rate in zip(progress, rates):
                    if has_log_norm:
                        value, grad, log_norm = objective_and_grad(param)
                    else:
                        value, grad = objective_and_grad(param)
                        log_norm = 0
                    grad = np.asarray(grad, dtype=float)
                    values.append(value)
                    norms.append(log_norm)
                    if rule == 'rmsprop':
                        grad_norm = np.exp(log_norm) if has_log_norm else np.sum(grad ** 2, axis=0)
                        if i == 0:
                            second = grad_norm if avg_grad_norm else grad ** 2
                        elif avg_grad_norm:
                            second = grad_norm * decay + (1. - decay) * grad_norm      # vb.py:452, as written
                        else:
                            second = second * decay + (1. - decay) * grad ** 2
                        step = rate * grad / np.sqrt(epsilon + second)
                    else:
                        if i == 0:
                            second, first = 0.9 * grad ** 2, 0.9 * grad               # vb.py:617-618
                        else:
                            second = second * beta2 + (1. - beta2) * grad ** 2
                            first = first * beta1 + (1. - beta1) * grad
                        m_hat = first / (1 - np.power(beta1, i + 2))
                        v_hat = second / (1 - np.power(beta2, i + 2))
                        step = rate * m_hat / np.sqrt(epsilon + v_hat)
                    history.append(param.copy())                     # pre-update iterate, vb.py:464
                    param = param - step
                    if len(history) > 100 * window:
                        del history[0]
                    if i % 100 == 0:
                        progress.set_description('Average Loss = {:,.6g}'.format(np.mean(values[max(0, i - 1000):i + 1])))
            except (KeyboardInterrupt, StopIteration):  # pragma: no cover
                progress.close()
            finally:
                progress.close()
        chains.append(np.array(history))
    chains = np.stack(chains, axis=0)
    chains_dev = to_dev(chains, _lib.F64)                             # one upload; every diagnostic reads it in place
    rhats = fn.compute_R_hat_adaptive_numpy(chains_dev, window_size=rhat_window).cpu().numpy()
    halfway = fn.compute_R_hat_halfway(chains_dev, interval=100, start=200).cpu().numpy()
    start_m = _first_converged_window(rhats[:, :K], r_mean_threshold, rhat_window, n_iters - tail_avg_iters)
    start_s = _first_converged_window(rhats[:, K:], r_sigma_threshold, rhat_window, n_iters - tail_avg_iters)
    avg_means, avg_sigmas = [], []
    for o in range(n_optimisers):
        avg_means.append(fn.stochastic_iterate_averaging(chains_dev[o, :, :K], start_m)[0].cpu().numpy())
        avg_sigmas.append(fn.stochastic_iterate_averaging(chains_dev[o, :, K:], start_s)[0].cpu().numpy())
    log = {'start_avg_mean_iters': start_m, 'start_avg_sigma_iters': start_s,
           'r_hat_mean': rhats[:, :K], 'r_hat_sigma': rhats[:, K:],
           'r_hat_mean_halfway': halfway[:, :K], 'r_hat_sigma_halfway': halfway[:, K:]}
    return param, chains, avg_means, avg_sigmas, np.array(values), np.array(norms), log


def rmsprop_IA_optimize_with_rhat(n_iters, objective_and_grad, init_param, K,
                                  has_log_norm=False, window=500, learning_rate=.01,
                                  epsilon=.000001, rhat_window=500, averaging=True, n_optimisers=1,
                                  r_mean_threshold=1.15, r_sigma_threshold=1.20, tail_avg_iters=2000,
                                  avg_grad_norm=False, learning_rate_end=None, device_loop=None):
    """RMSProp with iterate averaging started where the windowed split R-hat says the iterates are stationary
    (``vb.py:392-548``).  ``K`` = number of location parameters (the first K entries of the variational parameter;
    the rest are log scales).  Returns ``(param, history (n_optimisers, kept, P), averaged means per optimiser,
    averaged log scales per optimiser, value_history, log_norm_history, optimisation_log)``.

    ``device_loop`` (extension, as in ``adagrad_optimize``): ``None`` runs every restart on the device when the
    objective is a single-rank fused mean-field objective of this package (bit-identical iterates), ``False`` forces
    the reference's host loop, ``True`` raises if the device loop does not apply."""
    return _optimize_with_rhat('rmsprop', n_iters, objective_and_grad, init_param, K, has_log_norm, window,
                               learning_rate, epsilon, rhat_window, n_optimisers, r_mean_threshold,
                               r_sigma_threshold, tail_avg_iters, avg_grad_norm, learning_rate_end, device_loop)


def adam_IA_optimize_with_rhat(n_iters, objective_and_grad, init_param, K,
                               has_log_norm=False, window=500, learning_rate=.01,
                               epsilon=.000001, rhat_window=500, averaging=True, n_optimisers=1,
                               r_mean_threshold=1.15, r_sigma_threshold=1.20, tail_avg_iters=2000,
                               learning_rate_end=None, device_loop=None):
    """Adam (bias correction with exponent i + 2, first step seeded with 0.9 g, as ``vb.py:617-629``) with the same
    R-hat-triggered iterate averaging (``vb.py:554-712``); same return tuple as the RMSProp variant."""
    return _optimize_with_rhat('adam', n_iters, objective_and_grad, init_param, K, has_log_norm, window,
                               learning_rate, epsilon, rhat_window, n_optimisers, r_mean_threshold,
                               r_sigma_threshold, tail_avg_iters, False, learning_rate_end, device_loop)
