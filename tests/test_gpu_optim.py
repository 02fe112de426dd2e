"""R-hat diagnostics (csrc/rhat.cu behind viabel_b200.functions) and the R-hat-monitored optimisers against the
golden vectors of the REFERENCE's own functions (tests/golden/optimizers.npz, scripts/make_golden_optim.py)
and against the oracle on larger seeded histories.  Tolerance: the device sums the iterations in 8 stripes, the
reference row by row -> rtol 1e-10 on var_hat / R-hat; the iterate averages use numpy's order -> exact."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import golden_inputs as gi                                    # noqa: E402
from test_oracle_optim import GOLD, check_optim_case, run_optim_case   # noqa: E402

pytestmark = pytest.mark.gpu
RTOL = 1e-10


@pytest.fixture(scope='module')
def fn():
    from viabel_b200 import functions
    return functions


@pytest.mark.parametrize('name', gi.RHAT_CASES)
def test_compute_r_hat_golden(fn, name):
    chains, kw = gi.rhat_case(name)
    var_hat, rhat = fn.compute_R_hat(chains, **kw)
    assert isinstance(rhat, np.ndarray)
    np.testing.assert_allclose(var_hat, GOLD['rhat/%s/var_hat' % name], rtol=RTOL)
    np.testing.assert_allclose(rhat, GOLD['rhat/%s/rhat' % name], rtol=RTOL)
    dv, dr = fn.compute_R_hat(torch.from_numpy(chains).cuda(), **kw)       # device in -> device out
    assert dr.is_cuda
    np.testing.assert_array_equal(dr.cpu().numpy(), rhat)


@pytest.mark.parametrize('name', gi.RHAT_WINDOW_CASES)
def test_r_hat_windows_golden(fn, name):
    chains, window, interval, start = gi.rhat_window_case(name)
    np.testing.assert_allclose(fn.compute_R_hat_adaptive_numpy(chains, window_size=window),
                               GOLD['rhat_adaptive/%s' % name], rtol=RTOL)
    np.testing.assert_allclose(fn.compute_R_hat_halfway(chains, interval=interval, start=start),
                               GOLD['rhat_halfway/%s' % name], rtol=RTOL)


@pytest.mark.parametrize('name', gi.AVERAGING_CASES)
def test_iterate_averaging_golden(fn, name):
    est, start = gi.averaging_case(name)
    iters, mean = fn.stochastic_iterate_averaging(est, start)
    np.testing.assert_array_equal(iters, GOLD['averaging/%s/iters' % name])
    np.testing.assert_array_equal(mean, GOLD['averaging/%s/mean' % name])
    # a column block of a device-resident history is read in place
    wide = torch.from_numpy(np.concatenate([est, est[:, ::-1] * 2.0], axis=1).copy()).cuda()
    it2, _ = fn.stochastic_iterate_averaging(wide[:, :est.shape[1]], start)
    np.testing.assert_array_equal(it2.cpu().numpy(), iters)


def test_error_behaviour(fn):
    chains, _ = gi.rhat_case('c4_t400_p6')
    with pytest.raises(ValueError):
        fn.compute_R_hat(chains[:, :101], warmup=0)           # odd length: the reference's reshape error
    with pytest.raises(ValueError):
        fn.compute_R_hat_adaptive_numpy(chains, window_size=150)
    with pytest.raises(TypeError):
        fn.stochastic_iterate_averaging(chains[0], 400)


@pytest.mark.parametrize('shape', [(2, 5000, 20), (1, 2000, 1000), (3, 400, 33), (4, 64, 131840 // 64)])
def test_r_hat_against_oracle_larger(fn, shape):
    from oracle import functions_np as ofn
    chains = gi._ar_chains(*shape, seed=77, drift=2.0)
    var_hat, rhat = fn.compute_R_hat(chains, warmup=0)
    ov, orh = ofn.compute_R_hat(chains, warmup=0)
    np.testing.assert_allclose(var_hat, ov, rtol=RTOL)
    np.testing.assert_allclose(rhat, orh, rtol=RTOL)
    w = shape[1] // 4
    if w % 2 == 0:
        np.testing.assert_allclose(fn.compute_R_hat_adaptive_numpy(chains, window_size=w),
                                   ofn.compute_R_hat_adaptive_numpy(chains, window_size=w), rtol=RTOL)
    it, m = fn.stochastic_iterate_averaging(chains[0], shape[1] // 3)
    oi, om = ofn.stochastic_iterate_averaging(chains[0], shape[1] // 3)
    np.testing.assert_array_equal(it, oi)


@pytest.mark.parametrize('name', gi.OPTIM_CASES)
def test_optimisers_golden(name):
    from viabel_b200 import vb
    got = run_optim_case(name, vb.adagrad_optimize, vb.rmsprop_IA_optimize_with_rhat, vb.adam_IA_optimize_with_rhat)
    check_optim_case(name, got, RTOL)


@pytest.mark.parametrize('rule', ['rmsprop', 'adam'])
def test_optimisers_on_the_cuda_objective(rule):
    """The same fused KLVI objective (fresh family, same Philox stream) driven by this package's optimiser and by
    the oracle's: same iterates, R-hats and averages."""
    from oracle import functions_np as ofn
    from viabel_b200 import models, vb

    def make():
        fam = vb.mean_field_gaussian_variational_family(2)
        return vb.black_box_klvi(fam, models.Funnel(1.35), 500)
    init = np.array([0., -1., 1., 1.])
    kw = dict(rhat_window=100, tail_avg_iters=100, learning_rate=.02, n_optimisers=2, window=10)
    mine = getattr(vb, rule + '_IA_optimize_with_rhat')(300, make(), init, 2, **kw)
    theirs = getattr(ofn, rule + '_IA_optimize_with_rhat')(300, make(), init, 2, **kw)
    np.testing.assert_allclose(mine[0], theirs[0], rtol=1e-12)
    np.testing.assert_allclose(mine[1], theirs[1], rtol=1e-12)
    np.testing.assert_allclose(mine[4], theirs[4], rtol=1e-12)
    for k in theirs[6]:
        np.testing.assert_allclose(mine[6][k], theirs[6][k], rtol=1e-9, err_msg=k)
    for a, b in zip(mine[2] + mine[3], theirs[2] + theirs[3]):
        np.testing.assert_allclose(a, b, rtol=1e-12)


@pytest.mark.parametrize('objective', ['klvi', 'chivi', 'klvi_pd'])
@pytest.mark.parametrize('n', [500, 5000])
@pytest.mark.parametrize('rule', ['rmsprop', 'adam'])
def test_device_loop_is_bit_identical_to_the_host_loop(rule, n, objective):
    """vb_mf_optimizer_create: every restart on the device (n = 500: many iterations inside one single-CTA launch;
    n = 5000: one fused launch per iteration replayed from a CUDA graph) against the reference-shaped host loop over
    the same CUDA objective -- same Philox streams, same rounding: array_equal on every iterate and value."""
    from viabel_b200 import models, vb

    def make():
        fam = vb.mean_field_t_variational_family(10, 40)
        if objective == 'klvi':
            return vb.black_box_klvi(fam, models.EightSchoolsNCP(), n)
        if objective == 'klvi_pd':
            return vb.black_box_klvi_pd(fam, models.EightSchoolsNCP(), n)
        return vb.black_box_chivi(2.0, fam, models.EightSchoolsNCP(), n)
    init = np.concatenate([np.zeros(10), np.full(10, -1.0)])
    kw = dict(rhat_window=50, tail_avg_iters=100, learning_rate=.03, learning_rate_end=.005, n_optimisers=2, window=2)
    f = getattr(vb, rule + '_IA_optimize_with_rhat')
    dev = f(260, make(), init, 10, device_loop=True, **kw)
    host = f(260, make(), init, 10, device_loop=False, **kw)
    assert dev[1].shape == host[1].shape == (2, 200, 20)            # history capped at 100 * window
    np.testing.assert_array_equal(dev[1], host[1])
    np.testing.assert_array_equal(dev[0], host[0])
    np.testing.assert_array_equal(dev[4], host[4])
    for a, b in zip(dev[2] + dev[3], host[2] + host[3]):
        np.testing.assert_array_equal(a, b)
    for k in host[6]:
        np.testing.assert_array_equal(dev[6][k], host[6][k])


def test_device_loop_validation():
    from viabel_b200 import models, vb
    fam = vb.mean_field_gaussian_variational_family(2)
    obj = vb.black_box_klvi(fam, models.Funnel(1.35), 100)
    init = np.array([0., -1., 1., 1.])
    with pytest.raises(TypeError):
        vb.rmsprop_IA_optimize_with_rhat(100, obj, init, 2, avg_grad_norm=True, device_loop=True)
    with pytest.raises(TypeError):
        vb.adam_IA_optimize_with_rhat(100, gi.NoisyQuadratic(4, 1), init, 2, device_loop=True)
    with pytest.raises(ValueError):
        vb.adam_IA_optimize_with_rhat(100, obj, init, 2, learning_rate=-1.0)


def test_custom_ops_forward_to_the_same_entry_points(fn):
    chains, window, interval, start = gi.rhat_window_case('c2_t600_p8')
    dev = torch.from_numpy(chains).cuda()
    win = torch.tensor([[i * window, window] for i in range(chains.shape[1] // window)], dtype=torch.int64, device='cuda')
    vh, rh = torch.ops.viabel_b200.rhat_windows(dev, win)
    np.testing.assert_array_equal(rh.cpu().numpy(), fn.compute_R_hat_adaptive_numpy(chains, window_size=window))
    avg = torch.ops.viabel_b200.iterate_average(dev[0], 100)
    np.testing.assert_array_equal(avg.cpu().numpy(), fn.stochastic_iterate_averaging(chains[0], 100)[0])
