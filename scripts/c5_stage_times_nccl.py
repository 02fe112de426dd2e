import os, sys, time, warnings
sys.path.insert(0, '.')
import torch, torch.distributed as dist
warnings.simplefilter('ignore')
rank = int(os.environ['RANK']); world = int(os.environ['WORLD_SIZE']); lr = int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(lr)
dist.init_process_group('nccl')
import bench_workloads as W
from viabel_b200 import psis_dist
w = W.C5(rank, world, world); w.setup()
for _ in range(5): w.step_device()
torch.cuda.synchronize(); dist.barrier()
# wrap stages and comm
ev = []
def wrap(obj, name):
    f = getattr(obj, name)
    def g(*a, **k):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); r = f(*a, **k); e1.record(); ev.append((name, e0, e1)); return r
    setattr(obj, name, g)
for n in ['sample', 'pivot', 'filter', 'prepare_hist_dev', 'refine_select_dev', 'gather_compact', 'tail_dev', 'write', 'merge_div']:
    wrap(psis_dist.CudaStages, n)
for n in ['all_gather', 'all_reduce_max', 'all_reduce_sum']:
    wrap(psis_dist.DistComm, n)
N = 20
tot = {}
t0 = time.perf_counter()
e_all0 = torch.cuda.Event(enable_timing=True); e_all1 = torch.cuda.Event(enable_timing=True)
e_all0.record()
for _ in range(N):
    w.step_device()
e_all1.record()
torch.cuda.synchronize()
if rank == 0:
    print('step %.3f ms (events) %.3f ms (wall)' % (e_all0.elapsed_time(e_all1) / N, (time.perf_counter() - t0) / N * 1e3))
    order = []
    for name, e0, e1 in ev:
        if name not in tot: tot[name] = 0.0; order.append(name)
        tot[name] += e0.elapsed_time(e1)
    s = 0
    for n in order:
        print('%-20s %.1f us' % (n, tot[n] / N * 1e3)); s += tot[n] / N * 1e3
    print('sum of stages %.1f us' % s)
dist.destroy_process_group()
