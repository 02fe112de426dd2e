import sys, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from oracle import models_np as mm
from viabel_b200 import models
import os
def reg(N, d, seed=0):
    rs = np.random.RandomState(seed)
    X = rs.randn(N, d) / np.sqrt(d); bt = rs.randn(d)
    y = X @ bt + rs.standard_t(40, N)
    return X, y - y.mean(), bt
def rel(a, b): return np.max(np.abs(a - b)) / np.max(np.abs(b))
for path in ('0', '1'):
    os.environ['VB_REG_PATH'] = path
    for (N, d, n, sd) in [(300, 8, 40, .1), (5000, 33 if path == '0' else 32, 129, .1), (40000, 128, 512, .1), (70001, 64, 300, .1), (1 << 19, 128, 128, .02)]:
        X, y, bt = reg(N, d)
        X32, y32 = X.astype(np.float32), y.astype(np.float32)
        model = models.StudentTRegression(X32, y32, df=40.0, dtype='f32')
        om = mm.StudentTRegression(X32.astype(np.float64), y32.astype(np.float64), 40.0)
        b = (bt + sd * np.random.RandomState(3).randn(n, d)).astype(np.float32)
        lp, g = model.logp_and_grad(b, dtype='f32')
        olp, og = om.logp_and_grad(b.astype(np.float64))
        print('path', path, (N, d, n), 'logp rel %.2e grad rel %.2e' % (rel(lp, olp), rel(g, og)), flush=True)
