"""Generate tests/golden/*.npz by running the REFERENCE (kunalghosh/viabel, mounted read-only at
/root/reference) in the build container.  Inputs are defined by seeds (tests/golden_inputs.py)
so only the reference's outputs are stored.  Run:  python scripts/make_golden.py
"""
import importlib.util
import os
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
sys.dont_write_bytecode = True
import golden_inputs as gi   # noqa: E402

REF = os.environ.get('VIABEL_REFERENCE', '/root/reference')


def load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    rb = load('ref_bounds', os.path.join(REF, 'viabel', 'bounds.py'))
    rp = load('ref_psis', os.path.join(REF, 'notebooks', 'psis.py'))
    out = {}
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        # ---- bounds
        for name in gi.BOUNDS_CASES:
            lw, samples, kw = gi.bounds_case(name)
            res = rb.all_bounds(lw, samples, **kw)
            for k, v in res.items():
                out['bounds/%s/%s' % (name, k)] = np.float64(v)
            for alpha in (1.5, 2.0, 3.0):
                out['bounds/%s/d_alpha_%g' % (name, alpha)] = np.float64(rb.divergence_bound(lw, alpha))
        # ---- psis
        for name in gi.PSIS_CASES:
            lw = gi.psis_case(name)
            n = len(lw)
            # replay psislw's own bookkeeping (psis.py:157-196) to pin the tail set and permutation
            x = lw - np.max(lw)
            cutoff_ind = -int(np.ceil(min(0.2 * n, 3 * np.sqrt(n)))) - 1
            srt = np.argsort(x)
            xcutoff = max(x[srt[cutoff_ind]], np.log(np.finfo(float).tiny))
            tailinds, = np.where(x > xcutoff)
            x2si = np.argsort(x[tailinds])
            smoothed, k = rp.psislw(lw)
            out['psis/%s/k' % name] = np.float64(k)
            out['psis/%s/xcutoff' % name] = np.float64(xcutoff)
            out['psis/%s/tail_idx_sorted' % name] = tailinds[x2si].astype(np.int64)
            out['psis/%s/tail_values' % name] = smoothed[tailinds[x2si]]
            probe = gi.probe_indices(n)
            out['psis/%s/probe_values' % name] = smoothed[probe]
            out['psis/%s/sumlogs' % name] = np.float64(rp.sumlogs(lw))
        # ---- gpdfitnew / gpinv
        for name in gi.GPD_CASES:
            xs = gi.gpd_case(name)
            k, sigma = rp.gpdfitnew(xs, sort=False)
            out['gpd/%s/k' % name] = np.float64(k)
            out['gpd/%s/sigma' % name] = np.float64(sigma)
        # ---- psisloo and 2-D psislw (psis.py:68-109, 136-137)
        for name in gi.LOO_CASES:
            ll = gi.loo_case(name)
            loo, loos, ks = rp.psisloo(ll.copy())
            out['loo/%s/loo' % name] = np.float64(loo)
            out['loo/%s/loos' % name] = loos
            out['loo/%s/ks' % name] = ks
            lw2, ks2 = rp.psislw(-ll)
            out['loo/%s/lw_probe' % name] = np.ascontiguousarray(lw2[gi.probe_indices(ll.shape[0])])
            assert np.array_equal(ks, ks2)
        p = gi.gpinv_probs()
        for (k, s) in gi.GPINV_PARAMS:
            out['gpinv/%g_%g' % (k, s)] = rp.gpinv(p, k, s)
    dst = os.path.join(ROOT, 'tests', 'golden', 'reference_outputs.npz')
    np.savez_compressed(dst, **out)
    print('wrote', dst, os.path.getsize(dst), 'bytes,', len(out), 'arrays')


if __name__ == '__main__':
    main()
