import sys, numpy as np, torch
sys.path.insert(0,'.')
from viabel_b200 import vb
for d in [32, 64, 128, 256, 384, 512, 640, 768, 1024]:
    fam = vb.t_variational_family(d, 40)
    rs = np.random.RandomState(0)
    lam = torch.from_numpy(0.02 * rs.randn(fam.var_param_dim)).cuda()
    impl = fam.impl
    for _ in range(3): impl.chain_kernel(lam)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): impl.chain_kernel(lam)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    info = impl.chain_info()
    print('d %4d forward %.3f ms  GEMMs %d  -> %.2f us/GEMM  (%.1f TFLOP/s)  info %s' % (d, ms, info[4], 1e3 * ms / info[4], info[4] * 2.0 * d**3 / (ms * 1e-3) / 1e12, info))
