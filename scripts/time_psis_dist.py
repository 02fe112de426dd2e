"""Per-stage wall time of the sharded psislw protocol (viabel_b200/psis_dist.py) and of the sharded all_bounds,
under torchrun:  torchrun --nproc-per-node 2 scripts/time_psis_dist.py.  Every stage / collective is bracketed by
device synchronisation, so the numbers add up to MORE than the pipelined step; they show where the time goes."""
import collections
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
    if os.environ.get('NCCL_DEBUG', '').upper() == 'VERSION':      # (WARN prints the banner too)
        os.environ.pop('NCCL_DEBUG', None)
    dist.init_process_group('nccl', device_id=torch.device('cuda', int(os.environ['LOCAL_RANK'])))
    import viabel_b200
    from viabel_b200 import psis_dist, psis
    n = 125_000_000
    g = torch.Generator(device='cuda')
    g.manual_seed(7 + rank)
    u = torch.rand(n, dtype=torch.float64, device='cuda', generator=g)
    lw = torch.log(torch.expm1(-0.5 * torch.log1p(-u)) / 0.5)
    del u
    acc = collections.OrderedDict()

    def wrap(obj, name):
        fn = getattr(obj, name)

        def timed(*a, **k):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = fn(*a, **k)
            torch.cuda.synchronize()
            acc[name] = acc.get(name, 0.0) + time.perf_counter() - t0
            return r
        setattr(obj, name, timed)

    st, comm = psis_dist.CudaStages(), psis_dist.DistComm(None)
    for name in ('sample', 'pivot', 'filter', 'prepare', 'refine', 'refine_hist', 'refine_select', 'tail', 'write'):
        wrap(st, name)
    for name in ('all_gather', 'all_reduce_sum', 'all_reduce_max'):
        wrap(comm, name)
    reps = 10
    for it in range(3 + reps):
        if it == 3:
            acc.clear()
            torch.cuda.synchronize()
            dist.barrier()
            t_all = time.perf_counter()
        psis_dist.psislw_sharded(lw, 1.0, stages=st, comm=comm)
    torch.cuda.synchronize()
    total = (time.perf_counter() - t_all) / reps
    # un-instrumented, pipelined
    for _ in range(3):
        psis.psislw(lw, group=True)
    torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
    for _ in range(reps):
        psis.psislw(lw, group=True)
    torch.cuda.synchronize(); t_psis = (time.perf_counter() - t0) / reps
    fn = lambda p: 3.0 if p == 2 else 27.0          # noqa: E731
    for _ in range(3):
        viabel_b200.all_bounds(lw, moment_bound_fn=fn, q_var=1.0, group=True)
    torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
    for _ in range(reps):
        viabel_b200.all_bounds(lw, moment_bound_fn=fn, q_var=1.0, group=True)
    torch.cuda.synchronize(); t_bounds = (time.perf_counter() - t0) / reps
    if rank == 0:
        print('world %d: instrumented psislw %.3f ms; pipelined psislw %.3f ms; all_bounds %.3f ms'
              % (world, total * 1e3, t_psis * 1e3, t_bounds * 1e3))
        for k, v in acc.items():
            print('  %-16s %8.3f ms' % (k, v / reps * 1e3))
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
