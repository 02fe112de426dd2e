"""The oracle's mean-field families and KLVI / CHIVI objectives against golden vectors produced by the REFERENCE's
own viabel/vb.py (scripts/make_golden_vb.py: the reference code run in the build container, autograd's
value_and_grad / vector_jacobian_product stood in by complex-step differentiation of the reference's own objective
functions, exact to rounding).  This pins `oracle/vb_np.py` for the mean-field families: random streams, entropy,
log density, moments, objective values and gradients."""
import os

import numpy as np
import pytest

import golden_inputs as gi
from oracle import models_np as mm
from oracle import vb_np

GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'vb_reference.npz'))
RTOL = 1e-12


def oracle_family(kind, dim, df):
    return vb_np.mean_field_gaussian_variational_family(dim) if kind == 'mf_gauss' \
        else vb_np.mean_field_t_variational_family(dim, df)


def oracle_model(name):
    return mm.Funnel(1.35) if name == 'funnel' else mm.EightSchoolsNCP()


@pytest.mark.parametrize('name', gi.VB_CASES)
def test_log_density_inputs_equal_the_oracle_models(name):
    kind, dim, df, logp_name, lam, n = gi.vb_case(name)
    x = np.random.RandomState(3).randn(20, dim)
    np.testing.assert_allclose(getattr(gi, logp_name + '_logp')(x), oracle_model(logp_name).logp_and_grad(x)[0],
                               rtol=1e-14)


@pytest.mark.parametrize('name', gi.VB_CASES)
def test_family_closures(name):
    kind, dim, df, logp_name, lam, n = gi.vb_case(name)
    g = lambda k: GOLD['vb/%s/%s' % (name, k)]      # noqa: E731
    fam = oracle_family(kind, dim, df)
    assert fam.var_param_dim == int(g('var_param_dim'))
    np.testing.assert_allclose(fam.sample(lam, 5), g('sample_1'), rtol=RTOL)
    np.testing.assert_allclose(fam.sample(lam, 5), g('sample_2'), rtol=RTOL)
    x = fam.sample(lam, 6, seed=11)
    np.testing.assert_allclose(x, g('sample_seed11'), rtol=RTOL)
    np.testing.assert_allclose(fam.entropy(lam), g('entropy'), rtol=RTOL)
    np.testing.assert_allclose(fam.logdensity(x, lam), g('logdensity'), rtol=RTOL)
    m, c = fam.mean_and_cov(lam)
    np.testing.assert_allclose(m, g('mean'), rtol=RTOL)
    np.testing.assert_allclose(c, g('cov'), rtol=RTOL)
    for p in (2, 4):
        np.testing.assert_allclose(fam.pth_moment(p, lam), g('pth_moment_%d' % p), rtol=RTOL)


@pytest.mark.parametrize('name', gi.VB_CASES)
def test_klvi_objective_and_grad(name):
    kind, dim, df, logp_name, lam, n = gi.vb_case(name)
    obj = vb_np.black_box_klvi(oracle_family(kind, dim, df), oracle_model(logp_name), n)
    for call in (1, 2):
        v, grad = obj(lam)
        np.testing.assert_allclose(v, GOLD['vb/%s/klvi_value_%d' % (name, call)], rtol=RTOL)
        np.testing.assert_allclose(grad, GOLD['vb/%s/klvi_grad_%d' % (name, call)], rtol=1e-11, atol=1e-13)


@pytest.mark.parametrize('alpha', [2.0, 1.5])
@pytest.mark.parametrize('name', gi.VB_CASES)
def test_chivi_objective_and_grad(name, alpha):
    kind, dim, df, logp_name, lam, n = gi.vb_case(name)
    np.random.seed(123)                                   # the per-call seed comes from the global stream, vb.py:258
    obj = vb_np.black_box_chivi(alpha, oracle_family(kind, dim, df), oracle_model(logp_name), n)
    for call in (1, 2):
        v, grad = obj(lam)[:2]
        np.testing.assert_allclose(v, GOLD['vb/%s/chivi%g_value_%d' % (name, alpha, call)], rtol=RTOL)
        np.testing.assert_allclose(grad, GOLD['vb/%s/chivi%g_grad_%d' % (name, alpha, call)], rtol=1e-11, atol=1e-13)


# ------------------------------------------------------------------------------------------ full-rank Student-t
# Values, samples, entropy, log density and moments come from the reference's own code (sqrtm, eigh, its
# multivariate_t_logpdf) and are compared at 1e-11; its gradients were taken by a 4-point central difference of the
# reference's objective (truncation ~1e-10 relative), hence the looser gradient tolerance.  paragami's flattening
# layout is a stand-in written from its documentation (scripts/ref_stand_ins.py): that layout stays unpinned.
FR_GRAD_RTOL, FR_GRAD_ATOL = 1e-7, 1e-8


def fr_model(name):
    assert name == 'gauss3'
    return mm.CorrelatedGaussian(gi._FR_MEAN, gi._FR_PREC)


@pytest.mark.parametrize('name', gi.VB_FR_CASES)
def test_full_rank_family_closures(name):
    dim, df, logp_name, lam, n = gi.vb_fr_case(name)
    g = lambda k: GOLD['vb/%s/%s' % (name, k)]      # noqa: E731
    x = np.random.RandomState(3).randn(20, dim)
    np.testing.assert_allclose(getattr(gi, logp_name + '_logp')(x), fr_model(logp_name).logp_and_grad(x)[0], rtol=1e-13)
    fam = vb_np.t_variational_family(dim, df)
    assert fam.var_param_dim == int(g('var_param_dim'))
    np.testing.assert_allclose(fam.sample(lam, 5), g('sample_1'), rtol=1e-11)
    np.testing.assert_allclose(fam.sample(lam, 5), g('sample_2'), rtol=1e-11)
    xs = fam.sample(lam, 6, seed=11)
    np.testing.assert_allclose(xs, g('sample_seed11'), rtol=1e-11)
    np.testing.assert_allclose(fam.entropy(lam), g('entropy'), rtol=1e-11)
    np.testing.assert_allclose(fam.logdensity(xs, lam), g('logdensity'), rtol=1e-11)
    m, c = fam.mean_and_cov(lam)
    np.testing.assert_allclose(m, g('mean'), rtol=1e-11)
    np.testing.assert_allclose(c, g('cov'), rtol=1e-11)
    for p in (2, 4):
        np.testing.assert_allclose(fam.pth_moment(p, lam), g('pth_moment_%d' % p), rtol=1e-11)


@pytest.mark.parametrize('name', gi.VB_FR_CASES)
def test_full_rank_objectives(name):
    dim, df, logp_name, lam, n = gi.vb_fr_case(name)
    obj = vb_np.black_box_klvi(vb_np.t_variational_family(dim, df), fr_model(logp_name), n)
    for call in (1, 2):
        v, grad = obj(lam)
        np.testing.assert_allclose(v, GOLD['vb/%s/klvi_value_%d' % (name, call)], rtol=1e-11)
        np.testing.assert_allclose(grad, GOLD['vb/%s/klvi_grad_%d' % (name, call)], rtol=FR_GRAD_RTOL, atol=FR_GRAD_ATOL)
    np.random.seed(123)
    obj = vb_np.black_box_chivi(2.0, vb_np.t_variational_family(dim, df), fr_model(logp_name), n)
    for call in (1, 2):
        v, grad = obj(lam)[:2]
        np.testing.assert_allclose(v, GOLD['vb/%s/chivi2_value_%d' % (name, call)], rtol=1e-11)
        np.testing.assert_allclose(grad, GOLD['vb/%s/chivi2_grad_%d' % (name, call)], rtol=FR_GRAD_RTOL, atol=FR_GRAD_ATOL)


# ------------------------------------------------------------------------------------------ path-derivative KLVI
@pytest.mark.parametrize('name', gi.VB_CASES + gi.VB_FR_CASES)
def test_klvi_pd_objective_and_grad(name):
    """black_box_klvi_pd / _pd2 (vb.py:268-296).  The goldens themselves show that the reference's gradient equals
    KLVI's and that _pd2 is the same function as _pd."""
    fr = name in gi.VB_FR_CASES
    if fr:
        dim, df, logp_name, lam, n = gi.vb_fr_case(name)
        fam, model = (lambda: vb_np.t_variational_family(dim, df)), fr_model(logp_name)
    else:
        kind, dim, df, logp_name, lam, n = gi.vb_case(name)
        fam, model = (lambda: oracle_family(kind, dim, df)), oracle_model(logp_name)
    gtol = dict(rtol=FR_GRAD_RTOL, atol=FR_GRAD_ATOL) if fr else dict(rtol=1e-11, atol=1e-13)
    for tag, make in (('klvipd', vb_np.black_box_klvi_pd), ('klvipd2', vb_np.black_box_klvi_pd2)):
        if fr and tag == 'klvipd2':
            continue                       # not callable with the full-rank family in the reference either
        obj = make(fam(), model, n)
        for call in (1, 2):
            v, grad = obj(lam)
            np.testing.assert_allclose(v, GOLD['vb/%s/%s_value_%d' % (name, tag, call)], rtol=1e-11)
            np.testing.assert_allclose(grad, GOLD['vb/%s/%s_grad_%d' % (name, tag, call)], **gtol)
    if not fr:
        np.testing.assert_array_equal(GOLD['vb/%s/klvipd_grad_1' % name], GOLD['vb/%s/klvi_grad_1' % name])
        np.testing.assert_array_equal(GOLD['vb/%s/klvipd_grad_1' % name], GOLD['vb/%s/klvipd2_grad_1' % name])


def test_c1_pipeline_end_to_end():
    """Config C1 at a small size, entirely in the reference's code (family stream -> KLVI -> AdaGrad with a decaying
    rate): the oracle follows all 240 iterates."""
    obj = vb_np.black_box_klvi(vb_np.mean_field_gaussian_variational_family(2), mm.Funnel(1.35), gi.C1_GOLDEN_N)
    res = vb_np.adagrad_optimize(gi.C1_GOLDEN_ITERS, obj, np.array([0., -1., 1., 1.]), learning_rate=.01,
                                 learning_rate_end=.001)
    np.testing.assert_allclose(res[2], GOLD['vb/c1_adagrad/value_history'], rtol=1e-9)
    np.testing.assert_allclose(res[1], GOLD['vb/c1_adagrad/param_history'], rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(res[0], GOLD['vb/c1_adagrad/param'], rtol=1e-9, atol=1e-11)
