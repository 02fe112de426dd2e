"""The oracle's optimisers and R-hat diagnostics against golden vectors produced by the REFERENCE's own
viabel/vb.py and viabel/functions.py (scripts/make_golden_optim.py; autograd.numpy stood in by numpy).  This is
what pins `adagrad_optimize`, `learning_rate_schedule`, `rmsprop_IA_optimize_with_rhat`,
`adam_IA_optimize_with_rhat`, `compute_R_hat*` and `stochastic_iterate_averaging`."""
import os

import numpy as np
import pytest

import golden_inputs as gi
from oracle import functions_np as fn
from oracle import vb_np

GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'optimizers.npz'))
RTOL = 1e-12          # same numpy operations in the same order: agreement is in fact exact on this machine


@pytest.mark.parametrize('name', gi.RHAT_CASES)
def test_compute_r_hat(name):
    chains, kw = gi.rhat_case(name)
    var_hat, rhat = fn.compute_R_hat(chains, **kw)
    np.testing.assert_allclose(var_hat, GOLD['rhat/%s/var_hat' % name], rtol=RTOL)
    np.testing.assert_allclose(rhat, GOLD['rhat/%s/rhat' % name], rtol=RTOL)


def test_compute_r_hat_odd_length_is_the_references_reshape_error():
    chains, _ = gi.rhat_case('c4_t400_p6')
    with pytest.raises(ValueError):
        fn.compute_R_hat(chains[:, :101], warmup=0)


@pytest.mark.parametrize('name', gi.RHAT_WINDOW_CASES)
def test_r_hat_windows(name):
    chains, window, interval, start = gi.rhat_window_case(name)
    np.testing.assert_allclose(fn.compute_R_hat_adaptive_numpy(chains, window_size=window),
                               GOLD['rhat_adaptive/%s' % name], rtol=RTOL)
    np.testing.assert_allclose(fn.compute_R_hat_halfway(chains, interval=interval, start=start),
                               GOLD['rhat_halfway/%s' % name], rtol=RTOL)


@pytest.mark.parametrize('name', gi.AVERAGING_CASES)
def test_iterate_averaging(name):
    est, start = gi.averaging_case(name)
    iters, mean = fn.stochastic_iterate_averaging(est, start)
    np.testing.assert_array_equal(iters, GOLD['averaging/%s/iters' % name])
    np.testing.assert_array_equal(mean, GOLD['averaging/%s/mean' % name])
    with pytest.raises(TypeError):
        fn.stochastic_iterate_averaging(est, len(est))


@pytest.mark.parametrize('name', ['const', 'decay', 'short'])
def test_learning_rate_schedule(name):
    args = {'const': (40, .05, None), 'decay': (40, .05, .005), 'short': (7, .1, .01)}[name]
    np.testing.assert_array_equal(np.array(list(vb_np.learning_rate_schedule(*args))), GOLD['schedule/%s' % name])


def run_optim_case(name, adagrad, rmsprop, adam):
    kind, n_iters, init, kw, seed = gi.optim_case(name)
    obj = gi.NoisyQuadratic(len(init), seed, has_log_norm=kw.get('has_log_norm', False))
    if kind == 'adagrad':
        res = adagrad(n_iters, obj, init, **kw)
        return dict(zip(['param', 'param_history', 'value_history', 'log_norm_history'], res))
    res = (rmsprop if kind == 'rmsprop' else adam)(n_iters, obj, init, len(init) // 2, **kw)
    out = {'param': res[0], 'chains': res[1], 'value_history': res[4], 'log_norm_history': res[5]}
    for o, (m, s) in enumerate(zip(res[2], res[3])):
        out['avg_mean_%d' % o], out['avg_sigma_%d' % o] = m, s
    for k, v in res[6].items():
        out['log/' + k] = v
    return out


def check_optim_case(name, got, rtol):
    keys = [k for k in GOLD.files if k.startswith('optim/%s/' % name)]
    assert len(keys) >= 4
    for k in keys:
        np.testing.assert_allclose(np.asarray(got[k[len('optim/%s/' % name):]], dtype=float), GOLD[k], rtol=rtol,
                                   atol=1e-14, err_msg=k)


@pytest.mark.parametrize('name', gi.OPTIM_CASES)
def test_oracle_optimisers_match_reference(name):
    got = run_optim_case(name, vb_np.adagrad_optimize, fn.rmsprop_IA_optimize_with_rhat,
                         fn.adam_IA_optimize_with_rhat)
    check_optim_case(name, got, RTOL)
