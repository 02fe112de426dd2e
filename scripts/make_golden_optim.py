"""Golden vectors for the optimisers and the R-hat diagnostics, produced by running the
REFERENCE's own viabel/vb.py and viabel/functions.py in the build container.

`autograd` and `paragami` are not installed here.  The optimisers (`adagrad_optimize`,
`rmsprop_IA_optimize_with_rhat`, `adam_IA_optimize_with_rhat`, `learning_rate_schedule`) and the diagnostics
(`compute_R_hat*`, `stochastic_iterate_averaging`) only use the numpy namespace those packages re-export, so the
modules are imported with the stand-ins of scripts/ref_stand_ins.py (`autograd.numpy` forwards to numpy; nothing
here differentiates).  The reference sources are executed where they lie (read-only); only outputs are stored.

Run:  python scripts/make_golden_optim.py      ->  tests/golden/optimizers.npz
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

from ref_stand_ins import load_reference   # noqa: E402


def main():
    vb, fn = load_reference()
    out = {}
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        # ---- R-hat family on seeded chains
        for name in gi.RHAT_CASES:
            chains, kw = gi.rhat_case(name)
            var_hat, rhat = fn.compute_R_hat(chains, **kw)
            out['rhat/%s/var_hat' % name] = var_hat
            out['rhat/%s/rhat' % name] = rhat
        for name in gi.RHAT_WINDOW_CASES:
            chains, window, interval, start = gi.rhat_window_case(name)
            out['rhat_adaptive/%s' % name] = fn.compute_R_hat_adaptive_numpy(chains, window_size=window)
            out['rhat_halfway/%s' % name] = fn.compute_R_hat_halfway(chains, interval=interval, start=start)
        for name in gi.AVERAGING_CASES:
            est, start = gi.averaging_case(name)
            iters, mean = fn.stochastic_iterate_averaging(est, start)
            out['averaging/%s/iters' % name] = iters
            out['averaging/%s/mean' % name] = mean
        # ---- optimisers on a deterministic noisy objective
        for name in gi.OPTIM_CASES:
            kind, n_iters, init, kw, obj_seed = gi.optim_case(name)
            obj = gi.NoisyQuadratic(len(init), obj_seed, has_log_norm=kw.get('has_log_norm', False))
            if kind == 'adagrad':
                res = vb.adagrad_optimize(n_iters, obj, init, **kw)
                keys = ['param', 'param_history', 'value_history', 'log_norm_history']
                for k, v in zip(keys, res):
                    out['optim/%s/%s' % (name, k)] = np.asarray(v)
            else:
                f = vb.rmsprop_IA_optimize_with_rhat if kind == 'rmsprop' else vb.adam_IA_optimize_with_rhat
                K = len(init) // 2
                res = f(n_iters, obj, init, K, **kw)
                out['optim/%s/param' % name] = np.asarray(res[0])
                out['optim/%s/chains' % name] = np.asarray(res[1])
                for o, (m, s) in enumerate(zip(res[2], res[3])):
                    out['optim/%s/avg_mean_%d' % (name, o)] = np.asarray(m)
                    out['optim/%s/avg_sigma_%d' % (name, o)] = np.asarray(s)
                out['optim/%s/value_history' % name] = np.asarray(res[4])
                out['optim/%s/log_norm_history' % name] = np.asarray(res[5])
                for k, v in res[6].items():
                    out['optim/%s/log/%s' % (name, k)] = np.asarray(v)
        sched = {'const': (40, .05, None), 'decay': (40, .05, .005), 'short': (7, .1, .01)}
        for name, (n, lr, lre) in sched.items():
            out['schedule/%s' % name] = np.array(list(vb.learning_rate_schedule(n, lr, lre)))
    path = os.path.join(ROOT, 'tests', 'golden', 'optimizers.npz')
    np.savez_compressed(path, **out)
    print('wrote %s: %d arrays, %.1f kB' % (path, len(out), os.path.getsize(path) / 1e3))


if __name__ == '__main__':
    main()
