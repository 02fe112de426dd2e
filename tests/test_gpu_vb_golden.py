"""The CUDA mean-field path against golden vectors of the REFERENCE's own viabel/vb.py
(tests/golden/vb_reference.npz, scripts/make_golden_vb.py).  The product draws from Philox, the reference from
numpy's RandomState: the reference's base draws are regenerated here from the same seeds and handed to the CUDA
objective through `base_draws=`, so values and gradients are compared on identical samples (fp64, rtol 1e-10).
The deterministic closures (entropy, logdensity, mean_and_cov, pth_moment) are compared directly."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import golden_inputs as gi   # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'vb_reference.npz'))
RTOL = 1e-10


def family(kind, dim, df):
    from viabel_b200 import vb
    return vb.mean_field_gaussian_variational_family(dim) if kind == 'mf_gauss' \
        else vb.mean_field_t_variational_family(dim, df)


def model(name):
    from viabel_b200 import models
    return models.Funnel(1.35) if name == 'funnel' else models.EightSchoolsNCP()


def reference_draws(rs, kind, df, n, dim):
    """The base draw of the reference's `sample` (vb.py:57 / :151)."""
    return rs.randn(n, dim) if kind == 'mf_gauss' else rs.standard_t(df, size=(n, dim))


@pytest.mark.parametrize('name', gi.VB_CASES)
def test_family_closures_golden(name):
    kind, dim, df, logp_name, lam, n = gi.vb_case(name)
    g = lambda k: GOLD['vb/%s/%s' % (name, k)]      # noqa: E731
    fam = family(kind, dim, df)
    assert fam.var_param_dim == int(g('var_param_dim'))
    np.testing.assert_allclose(fam.entropy(lam), g('entropy'), rtol=RTOL)
    np.testing.assert_allclose(fam.logdensity(g('sample_seed11'), lam), g('logdensity'), rtol=RTOL)
    m, c = fam.mean_and_cov(lam)
    np.testing.assert_allclose(m, g('mean'), rtol=RTOL)
    np.testing.assert_allclose(c, g('cov'), rtol=RTOL)
    for p in (2, 4):
        np.testing.assert_allclose(fam.pth_moment(p, lam), g('pth_moment_%d' % p), rtol=RTOL)
    # the model plugin on the reference's samples equals the log density the reference was given
    x = g('sample_seed11')
    np.testing.assert_allclose(model(logp_name)(x), getattr(gi, logp_name + '_logp')(x), rtol=RTOL)


@pytest.mark.parametrize('name', gi.VB_CASES)
def test_klvi_golden(name):
    from viabel_b200 import vb
    kind, dim, df, logp_name, lam, n = gi.vb_case(name)
    obj = vb.black_box_klvi(family(kind, dim, df), model(logp_name), n, group=False)
    rs = np.random.RandomState(0)                          # the family's persistent stream, vb.py:49 / :143
    for call in (1, 2):
        eps = reference_draws(rs, kind, df, n, dim)
        v, grad = obj(lam, base_draws={'eps': eps})
        np.testing.assert_allclose(v, GOLD['vb/%s/klvi_value_%d' % (name, call)], rtol=RTOL)
        np.testing.assert_allclose(grad, GOLD['vb/%s/klvi_grad_%d' % (name, call)], rtol=RTOL, atol=1e-12)


@pytest.mark.parametrize('alpha', [2.0, 1.5])
@pytest.mark.parametrize('name', gi.VB_CASES)
def test_chivi_golden(name, alpha):
    from viabel_b200 import vb
    kind, dim, df, logp_name, lam, n = gi.vb_case(name)
    obj = vb.black_box_chivi(alpha, family(kind, dim, df), model(logp_name), n, group=False)
    seeds = np.random.RandomState(123)                     # np.random.seed(123); npr.randint(2**32) per call, vb.py:258
    for call in (1, 2):
        eps = reference_draws(np.random.RandomState(seeds.randint(2 ** 32)), kind, df, n, dim)
        v, grad = obj(lam, base_draws={'eps': eps})[:2]
        np.testing.assert_allclose(v, GOLD['vb/%s/chivi%g_value_%d' % (name, alpha, call)], rtol=RTOL)
        np.testing.assert_allclose(grad, GOLD['vb/%s/chivi%g_grad_%d' % (name, alpha, call)], rtol=RTOL, atol=1e-12)


# ------------------------------------------------------------------------------------------ full-rank Student-t
# fp64 CUDA chain (eigh-based symmetric square root, GEMMs, weighted Gram matrix) against the reference's own
# sqrtm / eigh / multivariate_t_logpdf code.  The golden gradients are 4-point central differences of the
# reference's objective (good to ~1e-10), so gradients are compared at 1e-7; everything else at 1e-10.
def fr_base(rs, df, n, dim):
    """Draw order of the reference's `sample` (vb.py:204-206): the chi-square scales first, then the normals."""
    chi2 = rs.chisquare(df, n)
    return {'chi2': chi2, 'z': rs.randn(n, dim)}


@pytest.mark.parametrize('name', gi.VB_FR_CASES)
def test_full_rank_golden(name):
    from viabel_b200 import models, vb
    dim, df, logp_name, lam, n = gi.vb_fr_case(name)
    g = lambda k: GOLD['vb/%s/%s' % (name, k)]      # noqa: E731
    fam = vb.t_variational_family(dim, df)
    mdl = models.CorrelatedGaussian(gi._FR_MEAN, gi._FR_PREC)
    assert fam.var_param_dim == int(g('var_param_dim'))
    np.testing.assert_allclose(fam.entropy(lam), g('entropy'), rtol=RTOL)
    np.testing.assert_allclose(fam.logdensity(g('sample_seed11'), lam), g('logdensity'), rtol=RTOL)
    m, c = fam.mean_and_cov(lam)
    np.testing.assert_allclose(m, g('mean'), rtol=RTOL)
    np.testing.assert_allclose(c, g('cov'), rtol=RTOL)
    for p in (2, 4):
        np.testing.assert_allclose(fam.pth_moment(p, lam), g('pth_moment_%d' % p), rtol=RTOL)
    obj = vb.black_box_klvi(vb.t_variational_family(dim, df), mdl, n, group=False)
    rs = np.random.RandomState(0)
    for call in (1, 2):
        v, grad = obj(lam, base_draws=fr_base(rs, df, n, dim))
        np.testing.assert_allclose(v, g('klvi_value_%d' % call), rtol=RTOL)
        np.testing.assert_allclose(grad, g('klvi_grad_%d' % call), rtol=1e-7, atol=1e-8)
    obj = vb.black_box_chivi(2.0, vb.t_variational_family(dim, df), mdl, n, group=False)
    seeds = np.random.RandomState(123)
    for call in (1, 2):
        base = fr_base(np.random.RandomState(seeds.randint(2 ** 32)), df, n, dim)
        v, grad = obj(lam, base_draws=base)[:2]
        np.testing.assert_allclose(v, g('chivi2_value_%d' % call), rtol=RTOL)
        np.testing.assert_allclose(grad, g('chivi2_grad_%d' % call), rtol=1e-7, atol=1e-8)


# ------------------------------------------------------------------------------------------ path-derivative KLVI
@pytest.mark.parametrize('name', gi.VB_CASES)
def test_klvi_pd_golden(name):
    from viabel_b200 import vb
    kind, dim, df, logp_name, lam, n = gi.vb_case(name)
    for tag, make in (('klvipd', vb.black_box_klvi_pd), ('klvipd2', vb.black_box_klvi_pd2)):
        obj = make(family(kind, dim, df), model(logp_name), n, group=False)
        rs = np.random.RandomState(0)
        for call in (1, 2):
            v, grad = obj(lam, base_draws={'eps': reference_draws(rs, kind, df, n, dim)})
            np.testing.assert_allclose(v, GOLD['vb/%s/%s_value_%d' % (name, tag, call)], rtol=RTOL)
            np.testing.assert_allclose(grad, GOLD['vb/%s/%s_grad_%d' % (name, tag, call)], rtol=RTOL, atol=1e-12)


@pytest.mark.parametrize('name', gi.VB_FR_CASES)
def test_klvi_pd_full_rank_golden(name):
    from viabel_b200 import models, vb
    dim, df, logp_name, lam, n = gi.vb_fr_case(name)
    obj = vb.black_box_klvi_pd(vb.t_variational_family(dim, df), models.CorrelatedGaussian(gi._FR_MEAN, gi._FR_PREC),
                               n, group=False)
    rs = np.random.RandomState(0)
    for call in (1, 2):
        v, grad = obj(lam, base_draws=fr_base(rs, df, n, dim))
        np.testing.assert_allclose(v, GOLD['vb/%s/klvipd_value_%d' % (name, call)], rtol=RTOL)
        np.testing.assert_allclose(grad, GOLD['vb/%s/klvipd_grad_%d' % (name, call)], rtol=1e-7, atol=1e-8)


@pytest.mark.parametrize('n', [1, 257, 5000, 70001])
def test_klvi_pd_matches_oracle_fused_philox_and_dense(n):
    """Beyond the golden sizes: fused plugins (external draws and in-kernel Philox, which must equal the KLVI launch
    on the same stream in its gradient), and the dense regression plugin (vb_model_logp_grad + vb_mf_reduce)."""
    from oracle import models_np as mm
    from oracle import vb_np
    from viabel_b200 import models, vb
    d = 10
    lam = np.concatenate([0.3 * np.random.RandomState(1).randn(d), np.full(d, -0.2)])
    ofam = vb_np.mean_field_t_variational_family(d, 40)
    base = ofam.impl.draw_base(np.random.RandomState(n), n)
    v, g = vb.black_box_klvi_pd(vb.mean_field_t_variational_family(d, 40), models.EightSchoolsNCP(), n,
                                group=False)(lam, base_draws=base)
    ov, og = vb_np.klvi_pd_value_and_grad(ofam.impl, mm.EightSchoolsNCP(), lam, base)
    np.testing.assert_allclose(v, ov, rtol=RTOL)
    np.testing.assert_allclose(g, og, rtol=RTOL, atol=1e-12)
    # Philox: same stream position for both objectives -> identical draws -> identical gradient
    v1, g1 = vb.black_box_klvi_pd(vb.mean_field_t_variational_family(d, 40), models.EightSchoolsNCP(), n, group=False)(lam)
    v2, g2 = vb.black_box_klvi(vb.mean_field_t_variational_family(d, 40), models.EightSchoolsNCP(), n, group=False)(lam)
    np.testing.assert_allclose(g1, g2, rtol=1e-12, atol=1e-14)
    assert np.isfinite(v1) and v1 != v2
    if n <= 5000:
        rs = np.random.RandomState(5)
        N, dd = 300, 8
        X, y = rs.randn(N, dd) / np.sqrt(dd), rs.randn(N)
        lam8 = np.concatenate([0.1 * rs.randn(dd), np.full(dd, -1.0)])
        ofam8 = vb_np.mean_field_gaussian_variational_family(dd)
        base8 = ofam8.impl.draw_base(np.random.RandomState(n), n)
        v, g = vb.black_box_klvi_pd(vb.mean_field_gaussian_variational_family(dd), models.StudentTRegression(X, y, df=40.0),
                                    n, group=False)(lam8, base_draws=base8)
        ov, og = vb_np.klvi_pd_value_and_grad(ofam8.impl, mm.StudentTRegression(X, y, 40.0), lam8, base8)
        np.testing.assert_allclose(v, ov, rtol=RTOL)
        np.testing.assert_allclose(g, og, rtol=RTOL, atol=1e-12)


# ------------------------------------------------------------------------------------------ notebooks/experiments.py
EXP_GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'experiments_reference.npz'))


@pytest.mark.parametrize('name', gi.EXPERIMENT_CASES)
def test_experiments_helpers_golden(name):
    """get_samples_and_log_weights / psis_correction / improve_with_psis / check_approx_accuracy on the reference's
    draws (first call on a fresh family's RandomState(0)) against outputs of the reference's own
    notebooks/experiments.py."""
    from viabel_b200 import experiments as ex
    kind, dim, df, logp_name, lam, _ = gi.vb_case(name)
    n = gi.EXPERIMENT_N
    g = lambda k: EXP_GOLD['exp/%s/%s' % (name, k)]      # noqa: E731
    base = {'eps': reference_draws(np.random.RandomState(0), kind, df, n, dim)}
    x, lw = ex.get_samples_and_log_weights(model(logp_name), family(kind, dim, df), lam, n, base_draws=base)
    np.testing.assert_allclose(x[::97], g('samples_probe'), rtol=1e-12)
    np.testing.assert_allclose(lw, g('log_weights'), rtol=RTOL, atol=1e-11)
    xt, slw, khat = ex.psis_correction(model(logp_name), family(kind, dim, df), lam, n, base_draws=base)
    assert xt.shape == (dim, n)
    np.testing.assert_allclose(slw, g('smoothed_log_weights'), rtol=1e-9, atol=1e-10)
    np.testing.assert_allclose(khat, g('khat'), rtol=1e-9)
    tm, tc = gi.experiment_truth(name)
    res, am, ac = ex.improve_with_psis(model(logp_name), family(kind, dim, df), lam, n, tm, tc, base_draws=base)
    np.testing.assert_allclose(am, g('approx_mean'), rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(ac, g('approx_cov'), rtol=1e-9, atol=1e-11)
    for k in ('mean_error', 'cov_error_2', 'cov_norm_2', 'cov_error_nuc', 'cov_norm_nuc', 'std_error', 'rel_std_error',
              'khat'):
        np.testing.assert_allclose(res[k], g('res/' + k), rtol=1e-8, err_msg=k)
    res2 = ex.check_approx_accuracy(family(kind, dim, df), lam, tm, tc)
    for k in ('mean_error', 'cov_error_2', 'cov_error_nuc', 'std_error', 'rel_std_error'):
        np.testing.assert_allclose(res2[k], g('approx_res/' + k), rtol=1e-10, err_msg=k)


@pytest.mark.parametrize('device_loop', [False, True])
def test_c1_pipeline_end_to_end_golden(device_loop):
    """Config C1 at a small size against the reference's own run (its family stream, black_box_klvi and
    adagrad_optimize): the CUDA objective is fed the reference's draws iteration by iteration, through the host loop
    and through the device-resident loop."""
    from viabel_b200 import models, vb
    n, iters = gi.C1_GOLDEN_N, gi.C1_GOLDEN_ITERS
    rs = np.random.RandomState(0)
    draws = np.stack([rs.randn(n, 2) for _ in range(iters)])
    obj = vb.black_box_klvi(vb.mean_field_gaussian_variational_family(2), models.Funnel(1.35), n, group=False)
    res = vb.adagrad_optimize(iters, obj, np.array([0., -1., 1., 1.]), learning_rate=.01, learning_rate_end=.001,
                              device_loop=device_loop, base_draws=draws)
    np.testing.assert_allclose(res[2], GOLD['vb/c1_adagrad/value_history'], rtol=1e-8)
    np.testing.assert_allclose(res[1], GOLD['vb/c1_adagrad/param_history'], rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(res[0], GOLD['vb/c1_adagrad/param'], rtol=1e-8, atol=1e-10)
