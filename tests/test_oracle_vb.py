"""The closed-form VI oracle (oracle/vb_np.py) against (a) scipy.stats for the family log
densities (the arithmetic autograd.scipy wraps), (b) the literal torch-autograd transcription
of viabel/vb.py (oracle/vb_autograd.py).  The pin against the real reference (golden vectors made by running its
own vb.py here) is tests/test_oracle_vb_golden.py; these tests cover sizes and cases beyond the golden ones."""
import numpy as np
import pytest
from scipy import stats

from oracle import models_np as mm
from oracle import vb_autograd as va
from oracle import vb_np


def rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


def test_family_logdensity_vs_scipy():
    rs = np.random.RandomState(0)
    d = 5
    lam = np.concatenate([rs.randn(d), 0.3 * rs.randn(d)])
    x = rs.randn(64, d) * 2
    fam = vb_np.mean_field_gaussian_variational_family(d)
    ref = stats.multivariate_normal.logpdf(x, lam[:d], np.diag(np.exp(2 * lam[d:])))     # vb.py:65
    np.testing.assert_allclose(fam.logdensity(x, lam), ref, rtol=1e-12)
    fam = vb_np.mean_field_t_variational_family(d, 7)
    ref = np.sum(stats.t.logpdf(x, 7, lam[:d], np.exp(lam[d:])), axis=-1)               # vb.py:162
    np.testing.assert_allclose(fam.logdensity(x, lam), ref, rtol=1e-12)
    fam = vb_np.t_variational_family(d, 9)
    lam = 0.2 * rs.randn(fam.var_param_dim)
    mu, Sigma = fam.mean_and_cov(lam)
    Sigma = Sigma * (9 - 2) / 9
    ref = stats.multivariate_t.logpdf(x, mu, Sigma, df=9)                               # _distributions.py
    np.testing.assert_allclose(fam.logdensity(x, lam), ref, rtol=1e-10)
    # entropy = 1/2 log det Sigma (vb.py:213)
    np.testing.assert_allclose(fam.entropy(lam), 0.5 * np.linalg.slogdet(Sigma)[1], rtol=1e-12)


def test_family_errors_and_moments():
    with pytest.raises(ValueError, match='df must be greater than 2'):
        vb_np.mean_field_t_variational_family(3, 2)
    with pytest.raises(ValueError, match='df must be greater than 2'):
        vb_np.t_variational_family(3, 1.5)
    fam = vb_np.mean_field_t_variational_family(3, 3.5)
    with pytest.raises(ValueError, match='only p = 2 or 4'):
        fam.pth_moment(3, np.zeros(6))
    with pytest.raises(ValueError, match='df must be greater than p'):
        fam.pth_moment(4, np.zeros(6))
    g = vb_np.mean_field_gaussian_variational_family(2)
    lam = np.array([0., 0., np.log(2.), np.log(3.)])
    assert np.isclose(g.pth_moment(2, lam), 13.0)
    assert np.isclose(g.pth_moment(4, lam), 2 * (16 + 81) + 13 ** 2)
    # sample() follows the reference's stream discipline: seed=None advances a RandomState(0)
    s1 = g.sample(lam, 4)
    s2 = g.sample(lam, 4)
    rs = np.random.RandomState(0)
    np.testing.assert_array_equal(s1, rs.randn(4, 2) * np.exp(lam[2:]) + lam[:2])
    np.testing.assert_array_equal(s2, rs.randn(4, 2) * np.exp(lam[2:]) + lam[:2])
    np.testing.assert_array_equal(g.sample(lam, 4, seed=7),
                                  np.random.RandomState(7).randn(4, 2) * np.exp(lam[2:]) + lam[:2])


CASES = [
    ('mf_gauss', 2, None, lambda: mm.Funnel(), lambda x: va.funnel_logp(x)),
    ('mf_t', 2, 40, lambda: mm.Funnel(), lambda x: va.funnel_logp(x)),
    ('mf_t', 1, 7, lambda: mm.NormalMixture(), lambda x: va.mixture_logp(x)),
    ('mf_gauss', 10, None, lambda: mm.EightSchoolsNCP(),
     lambda x: va.eight_schools_logp(x, mm.EIGHT_SCHOOLS_Y, mm.EIGHT_SCHOOLS_SIGMA)),
    ('mf_t', 10, 40, lambda: mm.EightSchoolsNCP(),
     lambda x: va.eight_schools_logp(x, mm.EIGHT_SCHOOLS_Y, mm.EIGHT_SCHOOLS_SIGMA)),
]


def _family(kind, d, df):
    if kind == 'mf_gauss':
        return vb_np.mean_field_gaussian_variational_family(d)
    if kind == 'mf_t':
        return vb_np.mean_field_t_variational_family(d, df)
    return vb_np.t_variational_family(d, df)


@pytest.mark.parametrize('kind,d,df,mk,tl', CASES)
def test_closed_form_vs_autograd_mean_field(kind, d, df, mk, tl):
    rs = np.random.RandomState(3)
    fam, model = _family(kind, d, df), mk()
    lam = rs.randn(2 * d) * 0.3
    base = fam.impl.draw_base(np.random.RandomState(11), 400)
    v, g = vb_np.klvi_value_and_grad(fam.impl, model, lam, base)
    v2, g2 = va.klvi(kind, tl, lam, base, d, df)
    assert abs(v - v2) <= 1e-12 * abs(v2) and rel(g, g2) < 1e-12
    for alpha in (2.0, 1.5):
        v, g, lw = vb_np.chivi_value_and_grad(fam.impl, model, alpha, lam, base, True)
        v2, g2, lw2 = va.chivi(kind, tl, alpha, lam, base, d, df)
        assert abs(v - v2) <= 1e-12 * abs(v2) and rel(g, g2) < 1e-11 and rel(lw, lw2) < 1e-12


def test_closed_form_vs_autograd_regression():
    rs = np.random.RandomState(4)
    N, d = 300, 6
    X = rs.randn(N, d) / np.sqrt(d)
    bt = rs.randn(d)
    y = X @ bt + rs.standard_t(40, N)
    y -= y.mean()
    model = mm.StudentTRegression(X, y, 40.0, row_block=128)
    fam = vb_np.mean_field_t_variational_family(d, 40)
    lam = np.concatenate([bt + 0.01 * rs.randn(d), np.log(1e-1) * np.ones(d)])
    base = fam.impl.draw_base(np.random.RandomState(5), 200)
    tl = lambda x: va.regression_logp(x, X, y, 40.0)   # noqa: E731
    v, g = vb_np.klvi_value_and_grad(fam.impl, model, lam, base)
    v2, g2 = va.klvi('mf_t', tl, lam, base, d, 40)
    assert abs(v - v2) <= 1e-12 * abs(v2) and rel(g, g2) < 1e-11
    v, g = vb_np.chivi_value_and_grad(fam.impl, model, 2.0, lam, base)
    v2, g2, _ = va.chivi('mf_t', tl, 2.0, lam, base, d, 40)
    assert abs(v - v2) <= 1e-12 * abs(v2) and rel(g, g2) < 1e-10


def test_closed_form_vs_autograd_full_rank():
    """Includes the sqrtm / Sylvester / paragami-fold chain and the eigh-based log q of the reference."""
    rs = np.random.RandomState(6)
    d = 16
    A = rs.randn(d, d)
    Sp = A @ A.T / d + np.eye(d)
    mp, Pp = rs.randn(d), np.linalg.inv(Sp)
    model = mm.CorrelatedGaussian(mp, Pp)
    fam = vb_np.t_variational_family(d, 40)
    assert fam.var_param_dim == d + d * (d + 1) // 2
    lam = 0.1 * rs.randn(fam.var_param_dim)          # generic Sigma: avoids the eigh degeneracy at Sigma = I
    base = fam.impl.draw_base(np.random.RandomState(5), 400)
    tl = lambda x: va.gauss_logp(x, mp, Pp)          # noqa: E731
    v, g = vb_np.klvi_value_and_grad(fam.impl, model, lam, base)
    v2, g2 = va.klvi('fr_t', tl, lam, base, d, 40)
    assert abs(v - v2) <= 1e-12 * abs(v2) and rel(g, g2) < 1e-11
    v, g, lw = vb_np.chivi_value_and_grad(fam.impl, model, 2.0, lam, base, True)
    v2, g2, lw2 = va.chivi('fr_t', tl, 2.0, lam, base, d, 40)
    assert abs(v - v2) <= 1e-12 * abs(v2) and rel(g, g2) < 1e-10 and rel(lw, lw2) < 1e-11
    x = fam.impl.transform(lam, base)
    assert rel(fam.impl.logdensity(x, lam), fam.impl.logq_of_base(lam, base)) < 1e-12
    # var_param = 0 <=> mu = 0, Sigma = I
    mu, cov = fam.mean_and_cov(np.zeros(fam.var_param_dim))
    np.testing.assert_allclose(mu, 0)
    np.testing.assert_allclose(cov, 40 / 38 * np.eye(d))


def test_adagrad_and_schedule():
    sched = list(vb_np.learning_rate_schedule(8, 0.1, 0.01))
    assert sched[:2] == [0.1, 0.1] and sched[6:] == [0.01, 0.01]
    assert all(sched[i] >= sched[i + 1] for i in range(7))
    with pytest.raises(ValueError, match='learning rate must be positive'):
        list(vb_np.learning_rate_schedule(4, 0.0, None))
    with pytest.raises(ValueError, match='greater than final'):
        list(vb_np.learning_rate_schedule(4, 0.1, 0.2))
    # quadratic objective: iterate average approaches the minimiser
    target = np.array([1.0, -2.0])
    obj = lambda p: (0.5 * np.sum((p - target) ** 2), p - target)   # noqa: E731
    opt, hist, vals, norms = vb_np.adagrad_optimize(2000, obj, np.zeros(2), learning_rate=0.1)
    assert hist.shape == (500, 2) and vals.shape == (2000,) and norms.shape == (2000,)
    np.testing.assert_allclose(opt, target, atol=0.05)
