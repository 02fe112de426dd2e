"""Golden vectors for the notebook helpers around the hot path: `get_samples_and_log_weights`,
`psis_correction`, `improve_with_psis`, `check_accuracy` of notebooks/experiments.py, run from the REFERENCE's own
file in the build container.  matplotlib / seaborn (absent, only used by its plotting helpers) are empty stand-ins;
viabel.vb comes through scripts/ref_stand_ins.py (nothing here differentiates).

Run:  python scripts/make_golden_experiments.py      ->  tests/golden/experiments_reference.npz
"""
import importlib
import os
import sys
import types
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
sys.path.insert(0, os.path.join(ROOT, 'scripts'))
sys.dont_write_bytecode = True
import golden_inputs as gi   # noqa: E402
from ref_stand_ins import REF, load_reference   # noqa: E402


def main():
    vb, _ = load_reference()
    for name in ('matplotlib', 'matplotlib.pyplot', 'seaborn'):
        sys.modules[name] = types.ModuleType(name)
    sys.modules['matplotlib'].pyplot = sys.modules['matplotlib.pyplot']
    sys.path.insert(0, os.path.join(REF, 'notebooks'))
    ex = importlib.import_module('experiments')
    out = {}
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        for name in gi.EXPERIMENT_CASES:
            kind, dim, df, logp_name, lam, _ = gi.vb_case(name)
            logp = getattr(gi, logp_name + '_logp')
            n = gi.EXPERIMENT_N

            def family():
                return vb.mean_field_gaussian_variational_family(dim) if kind == 'mf_gauss' \
                    else vb.mean_field_t_variational_family(dim, df)
            pre = 'exp/%s/' % name
            x, lw = ex.get_samples_and_log_weights(logp, family(), lam, n)
            out[pre + 'samples_probe'], out[pre + 'log_weights'] = x[::97].copy(), lw
            st, slw, khat = ex.psis_correction(logp, family(), lam, n)
            assert st.shape == (dim, n)
            out[pre + 'smoothed_log_weights'], out[pre + 'khat'] = slw, np.float64(khat)
            tm, tc = gi.experiment_truth(name)
            res, am, ac = ex.improve_with_psis(logp, family(), lam, n, tm, tc)
            out[pre + 'approx_mean'], out[pre + 'approx_cov'] = am, ac
            for k, v in res.items():
                out[pre + 'res/' + k] = np.float64(v)
            fam = family()
            res2 = ex.check_approx_accuracy(fam, lam, tm, tc)
            for k, v in res2.items():
                out[pre + 'approx_res/' + k] = np.float64(v)
    path = os.path.join(ROOT, 'tests', 'golden', 'experiments_reference.npz')
    np.savez_compressed(path, **out)
    print('wrote %s: %d arrays, %.1f kB' % (path, len(out), os.path.getsize(path) / 1e3))


if __name__ == '__main__':
    main()
