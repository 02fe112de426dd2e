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
  * full-rank family only (`set_mode('fd')`): `eigh` / `sqrtm` as the reference calls them are not
    complex-analytic, so differentiation falls back to a 4-point central difference (h = 2e-3, truncation
    O(h^4): about 1e-10 relative on these smooth objectives) -- a weaker pin than the complex step, stated as such
    in the tests' tolerances.  `paragami` is stood in by a minimal restatement of the three patterns the
    reference uses, following paragami's documented free parameterisation (vector: identity; PSD matrix: row-major
    lower triangle of the Cholesky factor with the diagonal logged; PatternDict: concatenation in insertion
    order).  That layout is the one thing of the full-rank path this cannot pin.
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
_MODE = 'complex'
_FD_H = 2e-3


def set_mode(mode):
    global _MODE
    assert mode in ('complex', 'fd')
    _MODE = mode


def _partial(f, x, j, args, before, weight=None):
    """d/dx_j of f (scalar, or weight . f for a vector f) on the replayed random streams."""
    def scal(v):
        _restore(before)
        out = f(v, *args)
        return out if weight is None else np.sum(weight * out)
    if _MODE == 'complex':
        xc = x.astype(complex)
        xc[j] += 1j * _H
        return np.imag(scal(xc)) / _H
    e = np.zeros_like(x)
    e[j] = _FD_H
    return (8.0 * (scal(x + e) - scal(x - e)) - (scal(x + 2 * e) - scal(x - 2 * e))) / (12.0 * _FD_H)


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
        grad = np.array([_partial(f, x, j, args, before) for j in range(x.size)])
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
        out = np.array([_partial(f, x, j, args, before, weight=np.real(g)) for j in range(x.size)])
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
    set_mode('fd')
    g = value_and_grad(lambda v: np.sum(np.log1p(v * v)))(m)[1]
    set_mode('complex')
    np.testing.assert_allclose(g, 2 * m / (1 + m * m), rtol=1e-9)


class NumericVectorPattern:
    def __init__(self, length):
        self.length = length

    def flat_length(self, free=None):
        return self.length

    def fold(self, flat, free=None):
        return np.asarray(flat)


class PSDSymmetricMatrixPattern:
    def __init__(self, size):
        self.size = size

    def flat_length(self, free=None):
        return self.size * (self.size + 1) // 2

    def fold(self, flat, free=None):
        L = np.zeros((self.size, self.size), dtype=np.asarray(flat).dtype)
        L[np.tril_indices(self.size)] = flat
        L[np.diag_indices(self.size)] = np.exp(np.diag(L))
        return L @ L.T


class PatternDict:
    def __init__(self, free_default=None):
        self.free_default = free_default
        self.patterns = {}

    def __setitem__(self, name, pattern):
        self.patterns[name] = pattern

    def flat_length(self, free=None):
        return sum(p.flat_length(free) for p in self.patterns.values())

    def fold(self, flat, free=None):
        out, at = {}, 0
        for name, p in self.patterns.items():
            k = p.flat_length(free)
            out[name] = p.fold(flat[at:at + k], free)
            at += k
        return out


def FlattenFunctionInput(original_fun, patterns, free, argnums):
    def wrapped(*args):
        args = list(args)
        args[argnums] = patterns.fold(args[argnums], free)
        return original_fun(*args)
    return wrapped


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
    mod('paragami', PatternDict=PatternDict, NumericVectorPattern=NumericVectorPattern,
        PSDSymmetricMatrixPattern=PSDSymmetricMatrixPattern, FlattenFunctionInput=FlattenFunctionInput)


def load_reference():
    """-> (viabel.vb, viabel.functions) of the reference checkout, executed where they lie."""
    sys.dont_write_bytecode = True
    install()
    check_stand_ins()
    if REF not in sys.path:
        sys.path.insert(0, REF)
    return importlib.import_module('viabel.vb'), importlib.import_module('viabel.functions')
