import sys, ctypes as C, numpy as np, torch
sys.path.insert(0,'.')
lib = C.CDLL('scratch/libchain_trace.so')   # nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --expt-relaxed-constexpr -DVB_CHAIN_TRACE -shared -o scratch/libchain_trace.so viabel_b200/csrc/fr_chain.cu -lcuda
lib.vb_fr_chain_workspace_bytes.restype = C.c_size_t
lib.vb_fr_chain_forward.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
lib.vb_fr_chain_trace_dump.argtypes = [C.c_void_p, C.c_int, C.c_int]
import bench_workloads as W
d = int(sys.argv[1]) if len(sys.argv) > 1 else 512
if d == 512:
    mp, Pp, lam = W.c3_problem(512)
else:
    lam = 0.02 * np.random.RandomState(0).randn(d + d * (d + 1) // 2)
vp = torch.from_numpy(lam).cuda()
L = torch.empty(d, d, dtype=torch.float64, device='cuda'); S = torch.empty_like(L)
scal = torch.empty(4, dtype=torch.float64, device='cuda')
ws = torch.zeros(lib.vb_fr_chain_workspace_bytes(d), dtype=torch.uint8, device='cuda')
def run():
    rc = lib.vb_fr_chain_forward(vp.data_ptr(), d, 0, L.data_ptr(), S.data_ptr(), None, scal.data_ptr(), ws.data_ptr(), ws.numel(), None)
    assert rc == 0, rc
for _ in range(3): run()
buf = (C.c_ulonglong * 8192)()
lib.vb_fr_chain_trace_dump(buf, 8192, 1)
run()
n = lib.vb_fr_chain_trace_dump(buf, 8192, 1)
ev = [(buf[i] >> 8, buf[i] & 255) for i in range(n)]
print('events', n)
names = {1: 'gemm start', 2: 'first slab landed', 3: 'k loop done', 4: 'exchange done', 5: 'epilogue done', 6: 'after grid barrier'}
from collections import defaultdict
acc = defaultdict(list)
for (t0, a), (t1, b) in zip(ev[:-1], ev[1:]):
    acc[(a, b)].append(t1 - t0)
tot = ev[-1][0] - ev[0][0]
print('total cycles', tot, '= %.1f us at 1.92 GHz' % (tot / 1920.0))
for k, v in sorted(acc.items()):
    print('%-22s -> %-22s n=%3d mean %7.0f cyc  sum %8d (%.1f%%)' % (names[k[0]], names[k[1]], len(v), np.mean(v), sum(v), 100.0 * sum(v) / tot))
