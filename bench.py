#!/usr/bin/env python
"""Benchmark of the Monte Carlo variational hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1|c2|c3|c4|c5] [--impl reference]

With no --workload the line's headline is c2 (BASELINE.json configs[1]) and its "workloads" object carries the same
keys (value, ms_per_step, e2e, roofline, cpu_baseline, clocks) for c3, c4 and c5, so the driver's BENCH / SCALE
records hold the north-star target configs too.  Prints ONE JSON line.  A "step" is one pass of the hot path over one batch of synthetic input
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
    ap.add_argument('--workload', default=os.environ.get('VB_BENCH_WORKLOAD') or None,
                    help='default: headline c2 (BASELINE.json configs[1]) plus a "workloads" object with c3, c4, c5; '
                         'a name (c1, c2, c3, c3chol, c4, c5, c2f64, n3, n4r) runs that workload alone')
    ap.add_argument('--headline-only', action='store_true', help='c2 without the c3 / c4 / c5 entries')
    ap.add_argument('--stage-times', action='store_true', help='per-stage CUDA-event times of the step (max / min over ranks)')
    ap.add_argument('--cpu-budget', type=float, default=10.0, help='seconds of CPU work per cpu_baseline')
    ap.add_argument('--ref-budget', type=float, default=75.0, help='seconds of CPU work of the --impl reference headline')
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


_TF32_PEAK = None


def measure_tf32_peak():
    """Dense TF32 tensor peak of THIS GPU, measured live with MEASURED_PEAKS.json's protocol for bf16:
    torch.matmul 8192^3 on fp32 operands with allow_tf32 (cuBLAS tf32 tensor-core GEMM), 2 N^3 flop, best of 10
    (burst) and back to back for about 1 s (sustained), CUDA events.  Library call used as the yardstick only."""
    global _TF32_PEAK
    if _TF32_PEAK is not None:
        return _TF32_PEAK
    import torch
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn(n, n, device='cuda', dtype=torch.float32)
        b = torch.randn(n, n, device='cuda', dtype=torch.float32)
        c = torch.empty(n, n, device='cuda', dtype=torch.float32)
        for _ in range(3):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize()
        best = float('inf')
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b, out=c)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        reps = max(10, int(1000.0 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize()
        sustained = e0.elapsed_time(e1) / reps
        flops = 2.0 * n ** 3
        _TF32_PEAK = {'burst': flops / (best * 1e-3) / 1e12, 'sustained': flops / (sustained * 1e-3) / 1e12,
                      'how': 'torch.matmul fp32 8192^3 with allow_tf32, best of 10 / back to back for 1 s, measured in this run'}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old
        del a, b, c
        torch.cuda.empty_cache()
    return _TF32_PEAK


def run_workload(wl_cls, args, steps, warmup, rank, world, local_rank, want_cpu, stage_times=False):
    """Time one workload; returns the dict of its JSON keys (rank 0) or None."""
    import torch
    import torch.distributed as dist
    wl = wl_cls(rank=rank, world=world, n_gpus=args.gpus)
    wl.setup()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world > 1:
            t = torch.tensor([x], dtype=torch.float64, device='cuda')
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

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
        return max_over_ranks(e0.elapsed_time(e1))

    if args.profile_kernel_only or args.profile_step:
        timed(wl.kernel_only if args.profile_kernel_only else wl.step_device, steps, warmup)
        return None

    # let the board's power limiter recover from whatever ran before (the previous workload's tensor-heavy steps or a
    # peak measurement): without the pause the 0.3 ms steps of c4 were timed at 1.1 - 1.5 GHz under sw_power_cap
    barrier()
    time.sleep(1.5)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_dev = timed(wl.step_device, steps, warmup)
    clocks = sampler.stop() if rank == 0 else None
    # dominant kernel alone, for the roofline: right after the step loop (same thermal / power state), before e2e
    kk = max(steps, 10)
    ms_kernel = timed(wl.kernel_only, kk, warmup) / kk
    # end to end through the public API with host buffers: wall clock around synchronous calls
    for _ in range(warmup):
        wl.step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        wl.step_e2e()
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    stages = wl.stage_times(steps) if (stage_times and hasattr(wl, 'stage_times')) else None
    if stages is not None and world > 1:
        keys = sorted(stages)
        t = torch.tensor([stages[k] for k in keys], dtype=torch.float64, device='cuda')
        tmax, tmin = t.clone(), t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        stages = {k: {'max_ms': float(a), 'min_ms': float(b)} for k, a, b in zip(keys, tmax.tolist(), tmin.tolist())}

    out = None
    if rank == 0:
        hbm, tf, src = peaks()
        units_total = wl.units_per_step * world * steps
        roof = wl.roofline(ms_kernel, hbm, tf)
        if roof.get('bound') == 'tensor' and roof.get('peak_kind') == 'f16':
            roof.setdefault('peak_source', 'MEASURED_PEAKS.json bf16_tflops' if src == 'measured' else 'fallback')
        elif roof.get('bound') == 'tensor' and roof.get('peak_kind') == 'f64':
            pass                                         # the workload measured its own fp64 MMA peak (vb_measure_dmma_peak)
        elif roof.get('bound') == 'tensor':
            pk = measure_tf32_peak()
            roof['peak'], roof['frac'] = pk['burst'], roof['achieved'] / pk['burst']
            roof['peak_sustained'], roof['peak_source'] = pk['sustained'], pk['how']
        else:
            roof['peak_source'] = 'MEASURED_PEAKS.json (driver-written copy bandwidth)' if src == 'measured' else 'fallback'
        roof['share_of_step'] = ms_kernel * wl.kernel_launches_per_step / (ms_dev / steps) if ms_dev > 0 else None
        tpath = os.path.join(ROOT, 'profiles', 'traffic.json')      # measured DRAM bytes of the same kernel(s) (ncu)
        if os.path.exists(tpath):
            with open(tpath) as fh:
                t = json.load(fh).get(wl.name)
            if t:
                roof['traffic'], roof['traffic_unit'], roof['traffic_source'] = t['bytes'], 'bytes/launch', t['source']
        out = {
            'value': units_total / (ms_dev / 1e3), 'unit': wl.unit, 'ms_per_step': ms_dev / steps, 'steps': steps,
            'warmup': warmup, 'dtype': wl.dtype, 'config': wl.config(),
            'e2e': {'value': units_total / e2e_s, 'unit': wl.unit,
                    'h2d_bytes_per_step': wl.h2d_bytes, 'd2h_bytes_per_step': wl.d2h_bytes},
            'gpu_launches': wl.launches_per_step * steps,
            'roofline': roof, 'clocks': clocks,
        }
        if stages is not None:
            out['stage_times_ms_per_step'] = stages
        out.update(wl.extra())
    del wl
    torch.cuda.empty_cache()
    return out


def main():
    args = parse()
    # the image exports NCCL_DEBUG=VERSION, which makes NCCL print its banner on stdout next to the ONE JSON line;
    # set before torch / NCCL are loaded
    if os.environ.get('NCCL_DEBUG', '').upper() == 'VERSION':      # (WARN prints the banner too)
        os.environ.pop('NCCL_DEBUG', None)
    import bench_workloads as W
    single = args.workload is not None
    head_name = args.workload if single else 'c2'
    extra_names = [] if single or args.headline_only else ['c3', 'c4', 'c5']
    wl_cls = W.WORKLOADS[head_name]
    steps = args.steps if args.steps is not None else wl_cls.default_steps
    warmup = args.warmup if args.warmup is not None else wl_cls.default_warmup
    warmup = max(warmup, 3)
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))

    def sub_steps(cls):
        # the headline times EXACTLY --steps; the other configs use their own defaults (sized so that the timed region
        # is tens of milliseconds: shorter ones sit inside clock ramps)
        return cls.default_steps, max(3, cls.default_warmup)

    if args.impl == 'reference':
        # the reference is a single-process CPU program: rank 0 alone runs it
        if rank != 0:
            return
        wl = wl_cls(rank=0, world=1, n_gpus=args.gpus)
        line = wl.reference_arm(steps, warmup, budget_s=args.ref_budget)
        line.update({'impl': 'reference', 'n_gpus': args.gpus, 'steps': steps, 'warmup': warmup,
                     'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'data': 'synthetic'})
        subs = {}
        for name in extra_names:
            cls = W.WORKLOADS[name]
            k, w = sub_steps(cls)
            k, w = min(k, 3), 1
            sub = cls(rank=0, world=1, n_gpus=args.gpus).reference_arm(k, w, budget_s=args.ref_budget / 4)
            sub.update({'steps': k, 'warmup': w})
            subs[name] = sub
        if subs:
            line['workloads'] = subs
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    import viabel_b200  # noqa: F401  (fails loudly when the CUDA library is missing)

    want_cpu = not args.no_cpu_baseline and world == 1       # CPU baseline: rank 0 at N = 1 only
    head = run_workload(wl_cls, args, steps, warmup, rank, world, local_rank, want_cpu, args.stage_times)
    subs = {}
    for name in extra_names:
        cls = W.WORKLOADS[name]
        k, w = sub_steps(cls)
        sub = run_workload(cls, args, k, w, rank, world, local_rank, want_cpu, args.stage_times)
        if sub is not None:
            subs[name] = sub
    if rank == 0 and head is not None and want_cpu:
        # CPU baselines AFTER every GPU timing: their thread pools / fork pools would otherwise perturb the
        # host-synchronising steps (c5) of the workloads timed after them
        head['cpu_baseline'] = wl_cls(rank=0, world=1, n_gpus=args.gpus).cpu_baseline(budget_s=args.cpu_budget)
        for name in subs:
            subs[name]['cpu_baseline'] = W.WORKLOADS[name](rank=0, world=1, n_gpus=args.gpus).cpu_baseline(
                budget_s=args.cpu_budget / 2)
    if rank == 0 and head is not None:
        line = {'metric': W.METRIC, 'value': head.pop('value'), 'unit': head.pop('unit'), 'n_gpus': world,
                'steps': steps, 'warmup': warmup, 'ms_per_step': head.pop('ms_per_step'), 'higher_is_better': True,
                'scaling': 'weak', 'vs_baseline': None, 'dtype': head.pop('dtype'), 'data': 'synthetic',
                'config': head.pop('config')}
        line.update(head)
        if subs:
            line['workloads'] = subs
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
