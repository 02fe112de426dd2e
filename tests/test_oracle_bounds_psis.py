"""The oracle restatements of bounds.py / psis.py against (a) the committed outputs of the
reference run in the build container, (b) the reference's own known-answer tests
(tests/test_bounds.py:22-70), (c) the notebook golden numbers, (d) the live reference module
when it is present."""
import warnings

import numpy as np
import pytest
from scipy.special import factorial2
from scipy.stats import norm

import golden_inputs as gi
from oracle import bounds_np, psis_np

MC_SAMPLES = 2000000
MC_TOL = 5 / np.sqrt(MC_SAMPLES)


def _quiet(fn, *a, **k):
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        return fn(*a, **k)


@pytest.mark.parametrize('name', gi.BOUNDS_CASES)
def test_bounds_match_reference_outputs(golden, name):
    lw, samples, kw = gi.bounds_case(name)
    res = _quiet(bounds_np.all_bounds, lw, samples, **kw)
    for k, v in res.items():
        assert v == golden['bounds/%s/%s' % (name, k)], k          # bit-identical arithmetic
    for alpha in (1.5, 2.0, 3.0):
        assert _quiet(bounds_np.divergence_bound, lw, alpha) == golden['bounds/%s/d_alpha_%g' % (name, alpha)]


def test_normal_mixture_notebook_numbers():
    """notebooks/normal-mixture.ipynb:138-143,172-177,202-207 (printed to 3 significant digits)."""
    lw, samples, _ = gi.bounds_case('mixture_nb')
    q_var = 5

    def fmt(res):
        return ['{:.3g}'.format(res[k]) for k in ('W2', 'd2', 'mean_error', 'std_error', 'cov_error')]
    assert fmt(_quiet(bounds_np.all_bounds, lw, samples)) == ['6.08', '0.768', '4.79', '6.08', '101']
    assert fmt(_quiet(bounds_np.all_bounds, lw, samples, q_var=q_var, log_norm_bound=0)) == \
        ['4.41', '0.277', '2.52', '4.41', '58.6']
    mfn = lambda order: factorial2(order - 1) ** (1 / order) * np.sqrt(q_var)   # noqa: E731 (notebook :97)
    assert fmt(_quiet(bounds_np.all_bounds, lw, moment_bound_fn=mfn, q_var=q_var)) == \
        ['2.72', '0.768', '2.72', '2.72', '26.9']


def _gauss_alpha_div(alpha, var1, var2):
    tmp = alpha * var2 - (alpha - 1) * var1
    if tmp < 0:
        return np.inf
    return -0.5 / (alpha - 1) * np.log(tmp) + .5 * alpha / (alpha - 1) * np.log(var2) - .5 * np.log(var1)


def _gauss_kl(var1, var2):
    return .5 * (var1 / var2 + np.log(var2 / var1) - 1)


def test_reference_known_answers_divergence():
    """tests/test_bounds.py:22-38 (fewer draws, tolerance scaled the same way)."""
    np.random.seed(846)
    var1, var2 = 4, 16
    samples = norm(scale=np.sqrt(var2)).rvs(MC_SAMPLES)
    lw = norm(scale=np.sqrt(var1)).logpdf(samples) - norm(scale=np.sqrt(var2)).logpdf(samples)
    for alpha in [1.5, 2, 3]:
        for elbo in [None, 0]:
            expected = _gauss_alpha_div(alpha, var1, var2)
            if elbo is None:
                expected += alpha / (alpha - 1) * _gauss_kl(var2, var1)
            np.testing.assert_allclose(_quiet(bounds_np.divergence_bound, lw, alpha, elbo), expected,
                                       atol=MC_TOL, rtol=MC_TOL)


def test_reference_known_answers_wasserstein_and_all():
    """tests/test_bounds.py:41-70."""
    np.random.seed(341)
    d2, stdev = 5.0, 3.5
    samples = norm.rvs(scale=stdev, size=MC_SAMPLES)
    res = bounds_np.wasserstein_bounds(d2, samples)
    np.testing.assert_allclose(res['W1'], 2 * stdev * np.sqrt(np.expm1(d2)), rtol=MC_TOL)
    np.testing.assert_allclose(res['W2'], 2 * stdev * (3 * np.expm1(d2)) ** 0.25, rtol=2 * MC_TOL)
    np.random.seed(1639)
    var1, var2 = 2.5, 9.3
    samples = norm(scale=np.sqrt(var2)).rvs(MC_SAMPLES)
    lw = norm(scale=np.sqrt(var1)).logpdf(samples) - norm(scale=np.sqrt(var2)).logpdf(samples)
    res = _quiet(bounds_np.all_bounds, lw, samples, q_var=var2, log_norm_bound=None)
    expected_d2 = _gauss_alpha_div(2, var1, var2) + 2 * _gauss_kl(var2, var1)
    np.testing.assert_allclose(res['d2'], expected_d2, rtol=MC_TOL)


def test_bounds_errors_and_warning_text():
    with pytest.raises(ValueError, match='alpha must be greater than 1'):
        bounds_np.divergence_bound(np.zeros(4), alpha=1.0)
    with pytest.raises(ValueError, match='must provides samples'):
        bounds_np.wasserstein_bounds(1.0)
    with pytest.warns(UserWarning, match='significant Monte Carlo error when computing ELBO'):
        bounds_np.divergence_bound(np.array([0.0, 10.0, -10.0, 5.0]))


@pytest.mark.parametrize('name', gi.PSIS_CASES)
def test_psis_matches_reference_outputs(golden, name):
    lw = gi.psis_case(name)
    out, k, det = _quiet(psis_np.psislw_detail, lw)
    pre = 'psis/%s/' % name
    assert k == golden[pre + 'k'] or (np.isinf(k) and np.isinf(golden[pre + 'k']))
    assert det['xcutoff'] == golden[pre + 'xcutoff']
    tail_sorted = det['tailinds'][det['tail_order']] if det['n2'] > 4 else None
    if tail_sorted is not None:
        assert np.array_equal(tail_sorted, golden[pre + 'tail_idx_sorted'])          # bit-exact set + order
        assert np.array_equal(out[tail_sorted], golden[pre + 'tail_values'])
    assert np.array_equal(out[gi.probe_indices(len(lw))], golden[pre + 'probe_values'])
    assert psis_np.sumlogs(lw) == golden[pre + 'sumlogs']
    # the stable tie-break variant agrees on tie-free input
    out2, k2, _ = _quiet(psis_np.psislw_detail, lw, sort_kind='stable')
    assert np.array_equal(out, out2) and (k == k2 or np.isinf(k))


@pytest.mark.parametrize('name', gi.GPD_CASES)
def test_gpdfit_matches_reference_outputs(golden, name):
    xs = gi.gpd_case(name)
    k, sigma = _quiet(psis_np.gpdfitnew, xs, sort=False)
    assert k == golden['gpd/%s/k' % name] and sigma == golden['gpd/%s/sigma' % name]


def test_gpinv_matches_reference_outputs(golden):
    p = gi.gpinv_probs()
    for (k, s) in gi.GPINV_PARAMS:
        np.testing.assert_array_equal(psis_np.gpinv(p, k, s), golden['gpinv/%g_%g' % (k, s)])


def test_psis_edge_cases():
    with pytest.raises(ValueError, match='1 or 2 dimensional'):
        psis_np.psislw(np.zeros((2, 2, 2)))
    with pytest.raises(ValueError, match='More than one'):
        psis_np.psislw(np.zeros(1))
    # two columns are smoothed independently
    lw2 = np.stack([gi.psis_case('gpd05_n5000'), gi.psis_case('gpd05_n5000')[::-1]], axis=1)
    out, ks = _quiet(psis_np.psislw, lw2)
    assert out.shape == lw2.shape and ks.shape == (2,)
    np.testing.assert_allclose(ks[0], ks[1], rtol=1e-10)


@pytest.mark.parametrize('name', gi.LOO_CASES)
def test_psisloo_matches_reference_outputs(golden, name):
    """psisloo / 2-D psislw (psis.py:68-109,136-137) against outputs of the real reference (scripts/make_golden.py)."""
    ll = gi.loo_case(name)
    keep = ll.copy()
    loo, loos, ks = _quiet(psis_np.psisloo, ll)
    assert np.array_equal(ll, keep)                                   # the input is not modified (lw = -log_lik copy)
    pre = 'loo/%s/' % name
    assert loo == golden[pre + 'loo'] and np.array_equal(loos, golden[pre + 'loos'])
    assert np.array_equal(ks, golden[pre + 'ks'])
    lw2, ks2 = _quiet(psis_np.psislw, -ll)
    assert lw2.flags.f_contiguous and np.array_equal(ks2, golden[pre + 'ks'])
    assert np.array_equal(lw2[gi.probe_indices(ll.shape[0])], golden[pre + 'lw_probe'])


def test_live_reference_agrees(reference_modules):
    """A/B against the imported reference on fresh random inputs (build container only)."""
    rb, rp = reference_modules
    rs = np.random.RandomState(99)
    lw = rs.randn(40000) * 1.3 - 7.0
    smp = rs.randn(40000, 4)
    a, b = _quiet(rb.all_bounds, lw, smp), _quiet(bounds_np.all_bounds, lw, smp)
    assert all(a[k] == b[k] for k in a)
    for n, kk in [(50, 0.5), (3000, 0.4), (60000, 0.6)]:
        w = gi.gpd_log_weights(n, kk, 1000 + n)
        o1, k1 = _quiet(rp.psislw, w)
        o2, k2 = _quiet(psis_np.psislw, w)
        assert np.array_equal(o1, o2) and k1 == k2
    ll = -gi.gpd_log_weights(1500 * 6, 0.55, 4242).reshape(1500, 6)
    a, b = _quiet(rp.psisloo, ll.copy()), _quiet(psis_np.psisloo, ll.copy())
    assert a[0] == b[0] and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
