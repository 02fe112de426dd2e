"""GPU parity of the callers either side of the hot path: the fused
get_samples_and_log_weights pass and the importance-weighted moments behind improve_with_psis, against the
oracle restatement of notebooks/experiments.py:60-89 (oracle/experiments_np.py) on identical base draws."""
import os
import sys
import warnings

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from oracle import experiments_np as ex_np      # noqa: E402
from oracle import models_np as mm              # noqa: E402
from oracle import psis_np, vb_np               # noqa: E402

pytestmark = pytest.mark.gpu


def _families(kind, d):
    from viabel_b200 import vb
    if kind == 'gauss':
        return vb.mean_field_gaussian_variational_family(d), vb_np.mean_field_gaussian_variational_family(d)
    return vb.mean_field_t_variational_family(d, 40), vb_np.mean_field_t_variational_family(d, 40)


@pytest.mark.parametrize('kind,model', [('gauss', 'funnel'), ('t', 'funnel'), ('t', 'eight'), ('gauss', 'mixture')])
def test_fused_samples_and_log_weights_match_oracle(kind, model):
    from viabel_b200 import experiments, models
    m, om = {'funnel': (models.Funnel(), mm.Funnel()), 'eight': (models.EightSchoolsNCP(), mm.EightSchoolsNCP()),
             'mixture': (models.NormalMixture(), mm.NormalMixture())}[model]
    d = om.dim
    fam, ofam = _families(kind, d)
    rs = np.random.RandomState(3)
    lam = np.concatenate([0.3 * rs.randn(d), np.log(np.full(d, 0.8))])
    base = ofam.impl.draw_base(np.random.RandomState(4), 20000)
    x, lw = experiments.get_samples_and_log_weights(m, fam, lam, 20000, base_draws=base)
    ox, olw = ex_np.get_samples_and_log_weights(om, ofam, lam, base)
    assert isinstance(x, np.ndarray) and x.shape == (20000, d) and lw.shape == (20000,)
    np.testing.assert_allclose(x, ox, rtol=1e-14, atol=1e-14)
    np.testing.assert_allclose(lw, olw, rtol=1e-10, atol=1e-10)
    # Philox mode: same law, device-resident outputs feed all_bounds without a host copy
    import torch
    import viabel_b200
    xd, lwd = experiments.get_samples_and_log_weights(m, fam, lam, 200000, device=True)
    assert xd.is_cuda and lwd.is_cuda and torch.isfinite(lwd).all()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        res = viabel_b200.all_bounds(lwd, samples=xd)
    assert np.isfinite(res['d2'])
    np.testing.assert_allclose(lwd.mean().item(), olw.mean(), atol=6 * olw.std() / np.sqrt(20000) + 0.05)


def test_composed_path_full_rank_and_gauss_plugin():
    from viabel_b200 import experiments, models, vb
    d = 6
    rs = np.random.RandomState(0)
    A = rs.randn(d, d)
    P = np.linalg.inv(A @ A.T / d + np.eye(d))
    model = models.CorrelatedGaussian(rs.randn(d), P)
    fam = vb.t_variational_family(d, 40)
    lam = 0.1 * rs.randn(fam.var_param_dim)
    x, lw = experiments.get_samples_and_log_weights(model, fam, lam, 5000)
    assert x.shape == (5000, d) and np.all(np.isfinite(lw))
    np.testing.assert_allclose(lw, model(x) - fam.logdensity(x, lam), rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize('n,d,dtype', [(1000, 1, 'f64'), (50000, 10, 'f64'), (200000, 33, 'f64'), (30000, 16, 'f32')])
def test_weighted_mean_and_cov_matches_numpy(n, d, dtype):
    from viabel_b200 import experiments
    rs = np.random.RandomState(n % 13)
    x = rs.randn(n, d) * (1 + np.arange(d)) + np.arange(d)
    if dtype == 'f32':
        x = x.astype(np.float32)
    lw = 2.0 * rs.randn(n) - 50.0
    import torch
    xin = torch.from_numpy(x).cuda() if dtype == 'f32' else x
    mean, cov = experiments.weighted_mean_and_cov(xin, lw)
    omean, ocov = ex_np.weighted_mean_and_cov(x.astype(np.float64).T, lw)
    np.testing.assert_allclose(mean, omean, rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(cov, ocov, rtol=1e-10, atol=1e-11)


def test_improve_with_psis_end_to_end():
    from viabel_b200 import experiments, models
    m, om = models.Funnel(), mm.Funnel()
    fam, ofam = _families('t', 2)
    lam = np.array([0.0, -0.5, 1.2, 0.3])
    true_mean = np.zeros(2)
    true_cov = np.diag([np.exp(2 * 1.35 ** 2), 1.35 ** 2])
    # oracle on the SAME samples: run the device pass once, then restate the rest on the host
    x, lw = experiments.get_samples_and_log_weights(m, fam, lam, 100000)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        slw, k = psis_np.psislw(lw, sort_kind='stable')
    omean, ocov = ex_np.weighted_mean_and_cov(x.T, slw)
    mean, cov = experiments.weighted_mean_and_cov(x, experiments.psislw(lw)[0])
    np.testing.assert_allclose(mean, omean, rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(cov, ocov, rtol=1e-8, atol=1e-10)
    res, am, ac = experiments.improve_with_psis(m, fam, lam, 100000, true_mean, true_cov)
    assert set(res) >= {'mean_error', 'cov_error_2', 'cov_norm_2', 'cov_error_nuc', 'std_error', 'rel_std_error', 'khat'}
    assert am.shape == (2,) and ac.shape == (2, 2) and np.isfinite(res['khat'])
    s, slw_d, kh = experiments.psis_correction(m, fam, lam, 5000)
    assert s.shape == (2, 5000) and slw_d.shape == (5000,)
    np.testing.assert_allclose(np.sum(np.exp(slw_d)), 1.0, rtol=1e-10)
