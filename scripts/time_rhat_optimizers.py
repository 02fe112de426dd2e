"""Wall-clock of the R-hat-monitored optimisers on the C1 objective (funnel, MF-Gaussian, 1000 samples):
device-resident loop against the reference-shaped host loop over the same CUDA objective.  Run on the GPU box."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, '.')
from viabel_b200 import models, vb   # noqa: E402

n_iters = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
init = np.array([0., -1., 1., 1.])
for rule in ('rmsprop', 'adam'):
    f = getattr(vb, rule + '_IA_optimize_with_rhat')
    for mode in (True, False):
        obj = vb.black_box_klvi(vb.mean_field_gaussian_variational_family(2), models.Funnel(1.35), 1000)
        f(200, obj, init, 2, rhat_window=100, tail_avg_iters=100, device_loop=mode)     # warm-up
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = f(n_iters, obj, init, 2, rhat_window=500, device_loop=mode)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print('%-8s %-6s %d iterations: %.1f ms total, %.2f us / iteration, final %s' % (
            rule, 'device' if mode else 'host', n_iters, dt * 1e3, dt / n_iters * 1e6, np.round(res[0], 3)), flush=True)
