"""Stand-ins that let the REFERENCE's own `viabel/vb.py` and `viabel/functions.py` be imported and run in the
build container, where `autograd` and `paragami` are not installed (no network).  Used only by the golden-vector
generators under scripts/; nothing under viabel_b200/, tests/ or bench.py imports this.

What stands in for what:
  * `autograd.numpy`, `autograd.numpy.random`   -> numpy itself (the reference uses no other property of the
    wrappers), except that every `RandomState` the reference creates is tracked so that a differentiation
    stand-in can replay a stochastic objective on the same draws.
  * `autograd.value_and_grad`, `autograd.vector_jacobian_product` -> COMPLEX-STEP differentiation of the very
    same Python function: d f / d x_j = Im f(x + i h e_j) / h with h = 1e-30, exact to rounding for analytic f
    (no subtraction, so no cancellation).  Each evaluation is replayed on the random-stream state the real call
    saw; afterwards the streams are left where ONE evaluation leaves them, as with autograd.
  * `autograd.scipy.stats.multivariate_normal.logpdf`, `autograd.scipy.stats.t.logpdf` -> the textbook formulas
    in plain numpy (so that they accept complex arguments); `check_stand_ins()` verifies them against
    scipy.stats at real arguments to 1e-13.
  * `paragami`, `autograd.scipy.linalg.sqrtm`, `autograd.extend` -> placeholders (the full-rank family is not
    exercised through these stand-ins: eigh / sqrtm are not complex-analytic as called).
"""
import importlib
import os
import sys
import types

import numpy as np
import scipy.linalg
import scipy.special
import scipy.stats

REF = os.environ.get('VIABEL_REFERENCE', '/root/reference')
_STREAMS = []            # every RandomState the reference creates
_H = 1e-30


class TrackedRandomState(np.random.RandomState):
    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        _STREAMS.append(self)


def _states():
    return [s.get_state() for s in _STREAMS] + [np.random.get_state()]


def _restore(states):
    for s, st in zip(_STREAMS, states[:-1]):
        s.set_state(st)
    np.random.set_state(states[-1])


def value_and_grad(f):
    def vg(x, *args):
        x = np.asarray(x, dtype=float)
        before = _states()
        value = f(x, *args)
        after = _states()
        grad = np.empty_like(x)
        for j in range(x.size):
            _restore(before)
            xc = x.astype(complex)
            xc[j] += 1j * _H
            grad[j] = np.imag(f(xc, *args)) / _H
        _restore(after)
        return float(np.real(value)), grad
    return vg


def vector_jacobian_product(f):
    def vjp(x, *rest):
        *args, g = rest
        x = np.asarray(x, dtype=float)
        before = _states()
        f(x, *args)
        after = _states()
        out = np.empty_like(x)
        for j in range(x.size):
            _restore(before)
            xc = x.astype(complex)
            xc[j] += 1j * _H
            out[j] = np.sum(np.real(g) * np.imag(f(xc, *args))) / _H
        _restore(after)
        return out
    return vjp


def mvn_logpdf(x, mean, cov):
    x = np.asarray(x)
    d = np.shape(mean)[0]
    r = x - mean
    sol = np.linalg.solve(cov, r.T).T if r.ndim == 2 else np.linalg.solve(cov, r)
    maha = np.sum(r * sol, axis=-1)
    return -0.5 * (d * np.log(2 * np.pi) + np.log(np.linalg.det(cov)) + maha)


def t_logpdf(x, df, loc=0, scale=1):
    y = (x - loc) / scale
    c = scipy.special.gammaln(0.5 * (df + 1)) - scipy.special.gammaln(0.5 * df) - 0.5 * np.log(df * np.pi)
    return c - np.log(scale) - 0.5 * (df + 1) * np.log1p(y * y / df)


def check_stand_ins():
    rs = np.random.RandomState(99)
    x = rs.randn(7, 3)
    m, s = rs.randn(3), np.exp(0.3 * rs.randn(3))
    np.testing.assert_allclose(mvn_logpdf(x, m, np.diag(s ** 2)),
                               scipy.stats.multivariate_normal.logpdf(x, m, np.diag(s ** 2)), rtol=1e-13)
    np.testing.assert_allclose(t_logpdf(x, 7.5, m, s), scipy.stats.t.logpdf(x, 7.5, m, s), rtol=1e-13)
    # complex step against a closed form: d/dx sum log1p(x^2) = 2x / (1 + x^2)
    g = value_and_grad(lambda v: np.sum(np.log1p(v * v)))(m)[1]
    np.testing.assert_allclose(g, 2 * m / (1 + m * m), rtol=1e-15)


def _unavailable(*a, **k):
    raise RuntimeError('autograd / paragami are not installed: this entry point is a placeholder')


def install():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    def primitive(f):
        f.defvjp = lambda *a, **k: None
        return f

    ag = mod('autograd', value_and_grad=value_and_grad, vector_jacobian_product=vector_jacobian_product)
    ag.extend = mod('autograd.extend', primitive=primitive, defvjp=lambda *a, **k: None)
    agnp = mod('autograd.numpy')                                     # numpy, except for the tracked RandomState
    agnp.__getattr__ = lambda name: getattr(np, name)
    npr = mod('autograd.numpy.random', RandomState=TrackedRandomState)
    npr.__getattr__ = lambda name: getattr(np.random, name)          # randint, seed, ... : the global numpy stream
    agnp.random = npr
    agnp.linalg = np.linalg
    ag.numpy = agnp
    sc = mod('autograd.scipy')
    st = mod('autograd.scipy.stats')
    st.multivariate_normal = mod('autograd.scipy.stats.multivariate_normal', logpdf=mvn_logpdf)
    st.t = mod('autograd.scipy.stats.t', logpdf=t_logpdf)
    sc.stats = st
    sc.linalg = mod('autograd.scipy.linalg', sqrtm=scipy.linalg.sqrtm)
    sc.special = scipy.special
    sys.modules['autograd.scipy.special'] = scipy.special
    ag.scipy = sc
    mod('paragami', PatternDict=_unavailable, NumericVectorPattern=_unavailable,
        PSDSymmetricMatrixPattern=_unavailable, FlattenFunctionInput=_unavailable)


def load_reference():
    """-> (viabel.vb, viabel.functions) of the reference checkout, executed where they lie."""
    sys.dont_write_bytecode = True
    install()
    check_stand_ins()
    if REF not in sys.path:
        sys.path.insert(0, REF)
    return importlib.import_module('viabel.vb'), importlib.import_module('viabel.functions')
