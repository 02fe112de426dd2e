"""CPU restatement of the notebook helpers around the hot path (test infrastructure only):
``get_samples_and_log_weights`` / ``psis_correction`` / ``improve_with_psis`` of
``notebooks/experiments.py:60-89``, on the oracle families of ``oracle/vb_np.py`` with externally
supplied base draws (the reference draws from the family's RandomState; parity tests hand both
sides the same draws).  PINNED: golden outputs of the reference's own ``notebooks/experiments.py`` run in the build
container (``tests/golden/experiments_reference.npz``, generator ``scripts/make_golden_experiments.py``; matplotlib /
seaborn and autograd stood in, see ``scripts/ref_stand_ins.py``), ``tests/test_oracle_experiments_golden.py``."""
import numpy as np

from . import psis_np


def get_samples_and_log_weights(logdensity, var_family, var_param, base):
    """experiments.py:60-63 with ``samples = transform(var_param, base)`` standing in for ``sample``."""
    samples = var_family.impl.transform(var_param, base)
    log_weights = logdensity(samples) - var_family.logdensity(samples, var_param)
    return samples, log_weights


def weighted_mean_and_cov(samples_t, slw):
    """experiments.py:81-86: ``samples_t`` is (d, n) like the reference's ``samples.T``."""
    slw = slw - np.max(slw)
    wts = np.exp(slw)
    wts /= np.sum(wts)
    approx_mean = np.sum(wts[np.newaxis, :] * samples_t, axis=1)
    approx_cov = np.cov(samples_t, aweights=wts, ddof=0)
    return approx_mean, np.atleast_2d(approx_cov)


def improve_with_psis(logdensity, var_family, var_param, base, sort_kind='stable'):
    samples, lw = get_samples_and_log_weights(logdensity, var_family, var_param, base)
    slw, khat = psis_np.psislw(lw, sort_kind=sort_kind)
    mean, cov = weighted_mean_and_cov(samples.T, slw)
    return mean, cov, khat
