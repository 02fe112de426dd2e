"""The oracle's notebook helpers (oracle/experiments_np.py) against golden vectors produced by the REFERENCE's own
notebooks/experiments.py (scripts/make_golden_experiments.py): samples, log weights, PSIS-smoothed weights, k-hat,
weighted mean / covariance.  Pins SURVEY 8(f) N1 / N2."""
import os
import warnings

import numpy as np
import pytest

import golden_inputs as gi
from oracle import experiments_np as ex
from oracle import psis_np
from test_oracle_vb_golden import oracle_family, oracle_model

GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'experiments_reference.npz'))


@pytest.mark.parametrize('name', gi.EXPERIMENT_CASES)
def test_samples_log_weights_psis_and_moments(name):
    kind, dim, df, logp_name, lam, _ = gi.vb_case(name)
    n = gi.EXPERIMENT_N
    g = lambda k: GOLD['exp/%s/%s' % (name, k)]      # noqa: E731
    fam, model = oracle_family(kind, dim, df), oracle_model(logp_name)
    base = fam.impl.draw_base(np.random.RandomState(0), n)          # first call on a fresh family's stream
    x, lw = ex.get_samples_and_log_weights(model, fam, lam, base)
    np.testing.assert_allclose(x[::97], g('samples_probe'), rtol=1e-12)
    np.testing.assert_allclose(lw, g('log_weights'), rtol=1e-11, atol=1e-12)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        slw, khat = psis_np.psislw(lw)
        mean, cov, khat2 = ex.improve_with_psis(model, fam, lam, base, sort_kind=None)
    np.testing.assert_allclose(slw, g('smoothed_log_weights'), rtol=1e-10, atol=1e-11)
    np.testing.assert_allclose(khat, g('khat'), rtol=1e-10)
    np.testing.assert_allclose(mean, g('approx_mean'), rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(cov, g('approx_cov'), rtol=1e-9, atol=1e-11)
