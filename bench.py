#!/usr/bin/env python
"""Benchmark of the Monte Carlo variational hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1|c2|c3|c4|c5] [--impl reference]

Prints ONE JSON line.  A "step" is one pass of the hot path over one batch of synthetic input
(one objective_and_grad call, or one psislw + all_bounds pass for c5).  `value` is whole-job
throughput with inputs resident in HBM, timed with CUDA events (max over ranks); `e2e` is the
same metric through the public Python API with host buffers (H2D/D2H inside the timed region);
`roofline` is for the dominant kernel timed alone; `cpu_baseline` is the oracle port timed on
this box's host cores on a bounded sample.  Under torchrun the sample axis is sharded (weak
scaling: per-GPU work fixed).  See bench_workloads.py for the workload definitions.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=None)
    ap.add_argument('--warmup', type=int, default=None)
    ap.add_argument('--workload', default=os.environ.get('VB_BENCH_WORKLOAD', 'c2'),
                    help='c2 (default: BASELINE.json configs[1]), c1, c3, c3chol, c4, c5, c2f64, n3 (psisloo), n4r (R-hat diagnostics)')
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--profile-kernel-only', action='store_true',
                    help='run only the dominant-kernel loop (for ncu captures)')
    ap.add_argument('--profile-step', action='store_true', help='run only the device-resident step loop (for ncu)')
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
         'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.gpu), '--query-gpu=' + self.Q, '--format=csv,noheader,nounits',
                 '-lms', '100'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(names, r[5:9]):
                if val.lower().startswith('active'):
                    reasons.add(name)
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as fh:
            d = json.load(fh)
        return d.get('hbm_gbs', 6650.0), d.get('bf16_tflops', 1590.0), 'measured'
    return 6650.0, 1590.0, 'fallback'


def main():
    args = parse()
    # the image exports NCCL_DEBUG=VERSION, which makes NCCL print its banner on stdout next to the ONE JSON line;
    # set before torch / NCCL are loaded
    if os.environ.get('NCCL_DEBUG', '').upper() == 'VERSION':      # (WARN prints the banner too)
        os.environ.pop('NCCL_DEBUG', None)
    import bench_workloads as W
    wl_cls = W.WORKLOADS[args.workload]
    steps = args.steps if args.steps is not None else wl_cls.default_steps
    warmup = args.warmup if args.warmup is not None else wl_cls.default_warmup
    warmup = max(warmup, 3)
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))

    if args.impl == 'reference':
        # the reference is a single-process CPU program: rank 0 alone runs it
        if rank != 0:
            return
        wl = wl_cls(rank=0, world=1, n_gpus=args.gpus)
        line = wl.reference_arm(steps, warmup)
        line.update({'impl': 'reference', 'n_gpus': args.gpus, 'steps': steps, 'warmup': warmup,
                     'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'data': 'synthetic'})
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    import viabel_b200  # noqa: F401  (fails loudly when the CUDA library is missing)

    wl = wl_cls(rank=rank, world=world, n_gpus=args.gpus)
    wl.setup()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k, w):
        """w warm-up calls, then k calls bracketed by barrier + synchronize; device time, max over ranks."""
        for _ in range(w):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device='cuda')
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    if args.profile_kernel_only or args.profile_step:
        timed(wl.kernel_only if args.profile_kernel_only else wl.step_device, steps, warmup)
        return

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_dev = timed(wl.step_device, steps, warmup)
    clocks = sampler.stop() if rank == 0 else None
    # end to end through the public API with host buffers: wall clock around synchronous calls
    for _ in range(warmup):
        wl.step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        wl.step_e2e()
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    # dominant kernel alone, for the roofline
    ms_kernel = timed(wl.kernel_only, max(steps, 10), warmup) / max(steps, 10)

    if rank == 0:
        hbm, tf, src = peaks()
        units_total = wl.units_per_step * world * steps
        roof = wl.roofline(ms_kernel, hbm, tf)
        roof['peak_source'] = src
        tpath = os.path.join(ROOT, 'profiles', 'traffic.json')      # measured DRAM bytes of the same kernel(s) (ncu)
        if os.path.exists(tpath):
            with open(tpath) as fh:
                t = json.load(fh).get(args.workload)
            if t:
                roof['traffic'], roof['traffic_unit'], roof['traffic_source'] = t['bytes'], 'bytes/launch', t['source']
        line = {
            'metric': wl.metric, 'value': units_total / (ms_dev / 1e3), 'unit': wl.unit,
            'n_gpus': world, 'steps': steps, 'warmup': warmup, 'ms_per_step': ms_dev / steps,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': wl.dtype,
            'data': 'synthetic', 'config': wl.config(),
            'e2e': {'value': units_total / e2e_s, 'unit': wl.unit,
                    'h2d_bytes_per_step': wl.h2d_bytes, 'd2h_bytes_per_step': wl.d2h_bytes},
            'gpu_launches': wl.launches_per_step * steps,
            'roofline': roof, 'clocks': clocks,
        }
        if not args.no_cpu_baseline:
            line['cpu_baseline'] = wl.cpu_baseline()
        line.update(wl.extra())
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
