"""Literal torch-CPU-fp64-autograd transcription of the reference's VI path (oracle).

Test infrastructure only.  ``autograd`` (HIPS) is not installable here, so
``torch.autograd`` stands in for ``value_and_grad`` / ``vector_jacobian_product``
(``viabel/vb.py:3,243,255``).  The computation graph is the reference's, operation by
operation — dense-covariance ``mvn.logpdf`` (vb.py:65), per-coordinate ``t.logpdf``
(vb.py:162), ``sqrtm`` with autograd's Sylvester VJP (vb.py:207), the eigh-based
``multivariate_t_logpdf`` (``viabel/_distributions.py:8-38``), paragami's free PSD folding
(vb.py:185-189) and CHIVI evaluated as forward → weights → VJP (vb.py:257-264).  Its only
job is to validate the closed-form gradients in ``oracle/vb_np.py``, which is itself pinned on
golden vectors of the reference (see ``oracle/__init__.py``).
"""
import math

import numpy as np
import torch

_LOG_2PI = math.log(2 * math.pi)
F64 = torch.float64


class _Sqrtm(torch.autograd.Function):
    """Symmetric PSD square root; backward = autograd's ``solve_sylvester(S', S', g)``."""

    @staticmethod
    def forward(ctx, A):
        lam, V = torch.linalg.eigh(A)
        r = lam.sqrt()
        S = (V * r) @ V.T
        ctx.save_for_backward(r, V)
        return S

    @staticmethod
    def backward(ctx, g):
        r, V = ctx.saved_tensors
        return V @ ((V.T @ g @ V) / (r[:, None] + r[None, :])) @ V.T


def sqrtm(A):
    return _Sqrtm.apply(A)


# ----------------------------------------------------------------------------- models (torch)
def funnel_logp(x, a=1.35):
    mu, ls = x[:, 0], x[:, 1]
    sig_d = -0.5 * _LOG_2PI - math.log(a) - 0.5 * (ls / a) ** 2
    mu_d = -0.5 * _LOG_2PI - ls - 0.5 * (mu / torch.exp(ls)) ** 2
    return sig_d + mu_d


def mixture_logp(x, c=2.0):
    t = x[:, 0]
    return torch.logaddexp(-0.5 * (t + c) ** 2 - 0.5 * _LOG_2PI,
                           -0.5 * (t - c) ** 2 - 0.5 * _LOG_2PI) - math.log(2)


def regression_logp(b, X, y, df=40.0, prior_sd=10.0):
    X = torch.as_tensor(X, dtype=F64)
    y = torch.as_tensor(y, dtype=F64)
    r = y[None, :] - b @ X.T
    return -0.5 * (b * b).sum(1) / prior_sd ** 2 - 0.5 * (df + 1) * torch.log1p(r * r / df).sum(1)


def eight_schools_logp(u, y, sigma):
    y = torch.as_tensor(y, dtype=F64)
    sigma = torch.as_tensor(sigma, dtype=F64)
    mu, lt, th = u[:, 0], u[:, 1], u[:, 2:]
    tau = torch.exp(lt)
    theta = mu[:, None] + tau[:, None] * th
    return (-0.5 * (mu / 5) ** 2 - torch.log1p((tau / 5) ** 2) + lt
            - 0.5 * (th ** 2).sum(1) - 0.5 * (((y[None, :] - theta) / sigma[None, :]) ** 2).sum(1))


def gauss_logp(x, m, P):
    m = torch.as_tensor(m, dtype=F64)
    P = torch.as_tensor(P, dtype=F64)
    dv = x - m[None, :]
    return -0.5 * ((dv @ P) * dv).sum(1)


# ----------------------------------------------------------------------------- families
def _fold_psd(lam, d):
    """paragami free PSD pattern: row-major lower triangle, log on the diagonal."""
    mu = lam[:d]
    idx = torch.tril_indices(d, d)
    L = torch.zeros(d, d, dtype=F64)
    L = L.index_put((idx[0], idx[1]), lam[d:])
    dg = torch.diagonal(L)
    L = L - torch.diag(dg) + torch.diag(torch.exp(dg))
    return mu, L @ L.T


def sample(kind, lam, base, d, df=None):
    if kind in ('mf_gauss', 'mf_t'):
        mean, log_s = lam[:d], lam[d:]
        eps = torch.as_tensor(base['eps'], dtype=F64)
        return eps * torch.exp(log_s) + mean                            # vb.py:57 / 151
    s = torch.sqrt(torch.as_tensor(base['chi2'], dtype=F64) / df)      # vb.py:204
    mu, Sigma = _fold_psd(lam, d)
    z = torch.as_tensor(base['z'], dtype=F64)
    return mu + (z @ sqrtm(Sigma)) / s[:, None]                         # vb.py:207-208


def entropy(kind, lam, d):
    if kind == 'mf_gauss':
        return 0.5 * d * (1.0 + _LOG_2PI) + lam[d:].sum()               # vb.py:61
    if kind == 'mf_t':
        return lam[d:].sum()                                            # vb.py:156
    _, Sigma = _fold_psd(lam, d)
    return 0.5 * torch.logdet(Sigma)                                    # vb.py:213


def logdensity(kind, x, lam, d, df=None):
    if kind == 'mf_gauss':                                              # vb.py:63-65
        mean, log_s = lam[:d], lam[d:]
        cov = torch.diag(torch.exp(2 * log_s))
        mvn = torch.distributions.MultivariateNormal(mean, covariance_matrix=cov)
        return mvn.log_prob(x)
    if kind == 'mf_t':                                                  # vb.py:158-162
        mean, log_s = lam[:d], lam[d:]
        yv = (x - mean) / torch.exp(log_s)
        c = math.lgamma(0.5 * (df + 1)) - math.lgamma(0.5 * df) - 0.5 * math.log(df * math.pi)
        return (c - log_s - 0.5 * (df + 1) * torch.log1p(yv * yv / df)).sum(-1)
    mu, Sigma = _fold_psd(lam, d)                                       # _distributions.py:27-37
    s, u = torch.linalg.eigh(Sigma)
    s_pinv = torch.where(s.abs() <= 1e-10, torch.zeros_like(s), 1 / s)
    U = u * torch.sqrt(s_pinv)
    log_pdet = torch.log(s).sum()
    out = math.lgamma(0.5 * (df + d)) - math.lgamma(0.5 * df) - 0.5 * d * math.log(math.pi * df)
    maha = ((x - mu) @ U).square().sum(-1)
    return out - 0.5 * log_pdet - 0.5 * (df + d) * torch.log(1 + maha / df)


# ----------------------------------------------------------------------------- objectives
def klvi(kind, logp, lam_np, base, d, df=None):
    lam = torch.tensor(np.asarray(lam_np, dtype=float), dtype=F64, requires_grad=True)
    x = sample(kind, lam, base, d, df)
    val = -(entropy(kind, lam, d) + logp(x).mean())                     # vb.py:240-241
    val.backward()
    return val.item(), lam.grad.numpy().copy()


def chivi(kind, logp, alpha, lam_np, base, d, df=None):
    lam = torch.tensor(np.asarray(lam_np, dtype=float), dtype=F64, requires_grad=True)

    def log_weights():
        x = sample(kind, lam, base, d, df)
        return logp(x) - logdensity(kind, x, lam, d, df)                # vb.py:252

    with torch.no_grad():
        lw = log_weights()                                              # forward #1, vb.py:259
        log_norm = lw.max()
        scaled = torch.exp(lw - log_norm) ** alpha                      # vb.py:261
        value = torch.log(scaled.mean()) / alpha + log_norm
    lw2 = log_weights()                                                 # forward #2 under the tape
    (g,) = torch.autograd.grad(lw2, lam, grad_outputs=scaled)           # vb.py:255
    return value.item(), (alpha * g / scaled.numel()).numpy().copy(), lw.numpy().copy()
