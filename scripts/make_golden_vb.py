"""Golden vectors for the mean-field part of viabel/vb.py: the families' closures and the KLVI / CHIVI
`objective_and_grad`, produced by running the REFERENCE's own code in the build container with the stand-ins of
scripts/ref_stand_ins.py (numpy for autograd.numpy; complex-step differentiation of the reference's own
objective functions for autograd's value_and_grad / vector_jacobian_product, exact to rounding).

Run:  python scripts/make_golden_vb.py      ->  tests/golden/vb_reference.npz
"""
import os
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
sys.path.insert(0, os.path.join(ROOT, 'scripts'))
sys.dont_write_bytecode = True
import golden_inputs as gi   # noqa: E402
from ref_stand_ins import load_reference, set_mode   # noqa: E402


def main():
    vb, _ = load_reference()
    out = {}
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        for name in gi.VB_CASES:
            kind, dim, df, logp_name, lam, n = gi.vb_case(name)
            logp = getattr(gi, logp_name + '_logp')

            def family():
                return vb.mean_field_gaussian_variational_family(dim) if kind == 'mf_gauss' \
                    else vb.mean_field_t_variational_family(dim, df)
            fam = family()
            pre = 'vb/%s/' % name
            out[pre + 'var_param_dim'] = np.int64(fam.var_param_dim)
            out[pre + 'sample_1'] = fam.sample(lam, 5)                  # persistent stream, first call
            out[pre + 'sample_2'] = fam.sample(lam, 5)                  # ... advanced by the first
            x = fam.sample(lam, 6, seed=11)
            out[pre + 'sample_seed11'] = x
            out[pre + 'entropy'] = np.float64(fam.entropy(lam))
            out[pre + 'logdensity'] = np.asarray(fam.logdensity(x, lam))
            m, c = fam.mean_and_cov(lam)
            out[pre + 'mean'], out[pre + 'cov'] = np.asarray(m), np.asarray(c)
            for p in (2, 4):
                if df is None or df > p:
                    out[pre + 'pth_moment_%d' % p] = np.float64(fam.pth_moment(p, lam))
            # KLVI: two consecutive calls on a fresh family (the second sees the advanced stream)
            obj = vb.black_box_klvi(family(), logp, n)
            for call in (1, 2):
                v, g = obj(lam)
                out[pre + 'klvi_value_%d' % call], out[pre + 'klvi_grad_%d' % call] = np.float64(v), g
            for tag, make in (('klvipd', vb.black_box_klvi_pd), ('klvipd2', vb.black_box_klvi_pd2)):
                obj = make(family(), logp, n)
                for call in (1, 2):
                    v, g = obj(lam)
                    out[pre + '%s_value_%d' % (tag, call)], out[pre + '%s_grad_%d' % (tag, call)] = np.float64(v), g
            # CHIVI: the seed of every call comes from the GLOBAL numpy stream (vb.py:258)
            for alpha in (2.0, 1.5):
                np.random.seed(123)
                obj = vb.black_box_chivi(alpha, family(), logp, n)
                for call in (1, 2):
                    v, g = obj(lam)
                    out[pre + 'chivi%g_value_%d' % (alpha, call)] = np.float64(v)
                    out[pre + 'chivi%g_grad_%d' % (alpha, call)] = g
        # ---- config C1 end to end on the reference's own streams: family + black_box_klvi + adagrad_optimize
        init = np.array([0., -1., 1., 1.])                                   # funnel-distribution.ipynb:241-243
        obj = vb.black_box_klvi(vb.mean_field_gaussian_variational_family(2), gi.funnel_logp, gi.C1_GOLDEN_N)
        res = vb.adagrad_optimize(gi.C1_GOLDEN_ITERS, obj, init, learning_rate=.01, learning_rate_end=.001)
        for k, v in zip(['param', 'param_history', 'value_history'], res[:3]):
            out['vb/c1_adagrad/' + k] = np.asarray(v)
        # ---- full-rank Student-t family: finite-difference gradients (see ref_stand_ins.py)
        set_mode('fd')
        for name in gi.VB_FR_CASES:
            dim, df, logp_name, lam, n = gi.vb_fr_case(name)
            logp = getattr(gi, logp_name + '_logp')
            fam = vb.t_variational_family(dim, df)
            pre = 'vb/%s/' % name
            out[pre + 'var_param_dim'] = np.int64(fam.var_param_dim)
            out[pre + 'sample_1'] = fam.sample(lam, 5)
            out[pre + 'sample_2'] = fam.sample(lam, 5)
            x = fam.sample(lam, 6, seed=11)
            out[pre + 'sample_seed11'] = x
            out[pre + 'entropy'] = np.float64(fam.entropy(lam))
            out[pre + 'logdensity'] = np.asarray(fam.logdensity(x, lam))
            m, c = fam.mean_and_cov(lam)
            out[pre + 'mean'], out[pre + 'cov'] = np.asarray(m), np.asarray(c)
            for p in (2, 4):
                out[pre + 'pth_moment_%d' % p] = np.float64(fam.pth_moment(p, lam))
            obj = vb.black_box_klvi(vb.t_variational_family(dim, df), logp, n)
            for call in (1, 2):
                v, g = obj(lam)
                out[pre + 'klvi_value_%d' % call], out[pre + 'klvi_grad_%d' % call] = np.float64(v), g
            # (black_box_klvi_pd2 passes var_param by keyword, which the full-rank family's flattened logdensity
            # does not accept in the reference either: only the _pd form exists for this family)
            for tag, make in (('klvipd', vb.black_box_klvi_pd),):
                obj = make(vb.t_variational_family(dim, df), logp, n)
                for call in (1, 2):
                    v, g = obj(lam)
                    out[pre + '%s_value_%d' % (tag, call)], out[pre + '%s_grad_%d' % (tag, call)] = np.float64(v), g
            for alpha in (2.0,):
                np.random.seed(123)
                obj = vb.black_box_chivi(alpha, vb.t_variational_family(dim, df), logp, n)
                for call in (1, 2):
                    v, g = obj(lam)
                    out[pre + 'chivi%g_value_%d' % (alpha, call)] = np.float64(v)
                    out[pre + 'chivi%g_grad_%d' % (alpha, call)] = g
        set_mode('complex')
    path = os.path.join(ROOT, 'tests', 'golden', 'vb_reference.npz')
    np.savez_compressed(path, **out)
    print('wrote %s: %d arrays, %.1f kB' % (path, len(out), os.path.getsize(path) / 1e3))


if __name__ == '__main__':
    main()
