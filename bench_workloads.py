"""Workload definitions for bench.py — the five BASELINE.json configs."""
import math
import os
import time

import numpy as np

METRIC = 'MC grad samples/s (KLVI/CHIVI) & PSIS+bounds weights/s; %HBM/TC roofline'


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def _pool_sum(fn, jobs, procs):
    """Run fn(job) for every job in a fork pool and return the list of results."""
    import multiprocessing as mp
    if procs <= 1 or len(jobs) <= 1:
        return [fn(j) for j in jobs]
    ctx = mp.get_context('fork')
    with ctx.Pool(procs) as pool:
        return pool.map(fn, jobs)


def blas_threads(n):
    """Context manager pinning the BLAS / OpenMP pools to n threads (torchrun exports OMP_NUM_THREADS=1, which
    would silently change the CPU denominator)."""
    from threadpoolctl import threadpool_limits
    return threadpool_limits(limits=int(n))


class Workload:
    default_steps, default_warmup = 20, 3
    metric, unit, dtype = METRIC, 'samples/s', 'f64'
    h2d_bytes = d2h_bytes = 0
    launches_per_step = 1
    kernel_launches_per_step = 1   # how often the roofline kernel (the unit kernel_only times) runs in one step
    # ---- CPU legs (oracle port on the host cores).  A workload states how many of its units one CPU step
    # processes; the rate is MEASURED on that bounded sample (units / seconds), never scaled from a different shape.
    cpu_probe_units = 0            # tiny step used to calibrate the sample size
    cpu_max_units = 0              # the full config (per GPU)
    cpu_round = 1                  # sample sizes are multiples of this
    cpu_cores = 1                  # threads the port actually uses

    def __init__(self, rank, world, n_gpus):
        self.rank, self.world, self.n_gpus = rank, world, n_gpus

    def extra(self):
        return {}

    # subclasses: cpu_prepare() builds host inputs once; cpu_step(units) runs the oracle on `units` units;
    # cpu_literal_step(units) (optional) the literal torch-autograd transcription; cpu_sample_desc(units) says what ran
    def cpu_prepare(self):
        pass

    def cpu_literal_step(self, units):
        return None

    def _cpu_pick_units(self, n_steps, budget_s):
        """Sample size so that n_steps CPU steps take about budget_s: calibrated on a small probe step."""
        self.cpu_prepare()
        self.cpu_step(self.cpu_probe_units)                      # warm caches / thread pools
        t0 = time.perf_counter()
        self.cpu_step(self.cpu_probe_units)
        per_unit = (time.perf_counter() - t0) / self.cpu_probe_units
        units = int(budget_s / max(n_steps, 1) / max(per_unit, 1e-12))
        units = max(self.cpu_probe_units, min(self.cpu_max_units, units // self.cpu_round * self.cpu_round))
        return units

    def _cpu_time(self, fn, units, k):
        t0 = time.perf_counter()
        for _ in range(k):
            fn(units)
        return (time.perf_counter() - t0) / k

    def cpu_baseline(self, budget_s=12.0, k=2):
        """The oracle port timed on this box's host cores on a bounded sample of the same workload."""
        with blas_threads(self.cpu_cores):
            units = self._cpu_pick_units(k, budget_s)
            dt = self._cpu_time(self.cpu_step, units, k)
        return {'value': units / dt, 'unit': self.unit, 'cores': self.cpu_cores, 'kind': 'port',
                'sample': self.cpu_sample_desc(units) + '; %d steps of %.2f s' % (k, dt)}

    def reference_arm(self, steps, warmup, budget_s=75.0):
        """bench.py --impl reference: the oracle port (the reference itself is pure Python and viabel.vb cannot be
        imported here) on all host threads it can use.  Runs EXACTLY `steps` timed steps after `warmup` untimed ones;
        each step is a bounded sample of the workload (stated in cpu_baseline.sample) sized by a probe so that the
        whole run ends within a few minutes.  `value` = units processed / seconds, measured, not extrapolated."""
        with blas_threads(self.cpu_cores):
            units = self._cpu_pick_units(steps + warmup, budget_s)
            for _ in range(warmup):
                self.cpu_step(units)
            dt = self._cpu_time(self.cpu_step, units, steps)
            rate = units / dt
            arms = {'closed_form': rate}
            # the literal transcription (forward under a tape + backward, CHIVI forward twice) is what the
            # reference's autograd does; the FASTER of the two is the denominator
            lit_units = max(self.cpu_probe_units, min(units, 4 * self.cpu_probe_units))
            t0 = time.perf_counter()
            if self.cpu_literal_step(lit_units) is not None:
                arms['literal_autograd'] = lit_units / (time.perf_counter() - t0)
        best = max(arms, key=arms.get)
        cb = {'value': arms[best], 'unit': self.unit, 'cores': self.cpu_cores, 'kind': 'port', 'arm': best,
              'arms': arms, 'sample': self.cpu_sample_desc(units) + '; %d timed steps of %.2f s after %d warm-up steps'
              % (steps, dt, warmup)}
        ms = dt * 1e3 if best == 'closed_form' else units / arms[best] * 1e3
        return {'metric': self.metric, 'value': arms[best], 'unit': self.unit, 'ms_per_step': ms, 'dtype': 'f64',
                'config': self.config(), 'units_per_step': units, 'cpu_baseline': cb,
                'e2e': {'value': arms[best], 'unit': self.unit, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}


class MeanFieldStages:
    """--stage-times for the sample-sharded mean-field objectives: CUDA-event time of (i) the local step (draws, model,
    reductions -> this rank's partial vector), (ii) the one all-gather of the partial vectors, (iii) the finalise
    kernel, each over `steps` back-to-back repetitions; bench.py reports max and min over ranks, which separates the
    collective's cost from rank skew."""

    def stage_times(self, steps):
        import torch
        import torch.distributed as dist
        from viabel_b200 import _lib
        from viabel_b200._device import all_gather_rows
        obj = self.obj
        vp = self.lam_dev
        impl = obj.impl
        base = getattr(self, '_stage_base', None)
        n_local = self.n
        index0 = self.rank * self.n

        def timed(fn):
            torch.cuda.synchronize()
            if self.world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / steps

        state = {}

        def local():
            state['out'], state['partials'] = obj._mf(vp, base, 0, 0, index0, n_local, finalize=(self.world == 1))
        t_local = timed(local)
        out = {'local_step': t_local}
        if self.world > 1:
            def gather():
                state['g'] = all_gather_rows(state['partials'], self.world, None)
            out['all_gather'] = timed(gather)

            def fin():
                obj._mf_merge(vp, state['partials'], self.world, None)
            out['all_gather_plus_finalize'] = timed(fin)
        return out


# ======================================================================================= C4
EIGHT_SCHOOLS_LAM = np.array(
    [4.4, 0.8, 0.3, 0.1, -0.1, 0.1, -0.2, -0.1, 0.35, 0.06] +
    list(np.log([3.2, 0.73, 0.96, 0.94, 0.97, 0.95, 0.92, 0.94, 0.93, 0.98])))


def _c4_cpu_job(job):
    """One shard of the oracle KLVI step: draws its own standard-t base draws like vb.py:151."""
    seed, n = job
    from oracle import models_np, vb_np
    fam = vb_np.mean_field_t_variational_family(10, 40)
    base = fam.impl.draw_base(np.random.RandomState(seed), n)
    v, g = vb_np.klvi_value_and_grad(fam.impl, models_np.EightSchoolsNCP(), EIGHT_SCHOOLS_LAM, base)
    return v, g


class C4(MeanFieldStages, Workload):
    """eight-schools non-centred, mean-field t (df 40), KLVI, 2^22 samples per step per GPU."""
    name = 'c4'
    n = 1 << 22
    d = 10
    default_steps, default_warmup = 200, 20

    def setup(self):
        import torch
        from viabel_b200 import models, vb, _lib
        self.torch, self.lib = torch, _lib
        self.fam = vb.mean_field_t_variational_family(self.d, 40)
        self.model = models.EightSchoolsNCP()
        self.obj = vb.black_box_klvi(self.fam, self.model, self.n * self.world, group=self.world > 1)
        # resident base draws for the streaming (HBM-roofline) mode: this rank's slice of the stream
        start = self.rank * self.n
        self.eps = self.fam.impl.base_draws(self.n, 0, 0, index0=start)['eps']
        self.lam_host = EIGHT_SCHOOLS_LAM.copy()
        self.lam_dev = torch.from_numpy(self.lam_host).cuda()
        self.units_per_step = self.n
        self.h2d_bytes = self.lam_host.nbytes
        self.d2h_bytes = 8 * (1 + 2 * self.d)
        # raw C-ABI call for the kernel-only timing
        from viabel_b200._device import ptr
        NP = 2 * self.d + 4
        self._out = torch.empty(1 + 2 * self.d, dtype=torch.float64, device='cuda')
        self._part = torch.empty(NP, dtype=torch.float64, device='cuda')
        self._ws = torch.zeros(_lib.lib.vb_mf_step_workspace_bytes(self.d), dtype=torch.uint8, device='cuda')
        self._args = (_lib.FAMILY_MF_T, 40.0, self.model.handle, _lib.OBJ_KLVI, 1.0, _lib.F64, ptr(self.lam_dev),
                      self.n, ptr(self.eps), 0, 0, 0, ptr(self._out), ptr(self._out) + 8, ptr(self._part),
                      ptr(self._ws), self._ws.numel())

    def step_device(self):
        # the draw is part of the step (vb.py:239 -> :151): Philox + Marsaglia-Tsang inside the fused kernel
        return self.obj(self.lam_dev)

    def step_e2e(self):
        return self.obj(self.lam_host)            # same, host var_param in / host (value, grad) out

    def kernel_only(self):
        # roofline instrument: the same kernel in streaming mode (fp64 base draws resident in HBM, read once)
        from viabel_b200._device import stream_ptr
        self.lib.check(self.lib.lib.vb_mf_step(*self._args, stream_ptr()))

    def roofline(self, ms_kernel, hbm_gbs, bf16_tflops):
        bytes_per_launch = self.n * self.d * 8            # base draws read once, nothing written
        ach = bytes_per_launch / (ms_kernel * 1e-3) / 1e9
        return {'bound': 'hbm', 'achieved': ach, 'peak': hbm_gbs, 'unit': 'GB/s', 'frac': ach / hbm_gbs,
                'traffic': None, 'kernel': 'mf_step_kernel<double, MF_T, EightSchools<8>, KLVI, streaming>',
                'ms_per_launch': ms_kernel, 'algorithmic_bytes_per_launch': bytes_per_launch}

    def config(self):
        return {'workload': 'c4: eight-schools ncp (d=10), mean-field Student-t(df=40) KLVI, 2^22 samples/step/GPU',
                'samples_per_step_per_gpu': self.n,
                'value_mode': 'Philox + Marsaglia-Tsang draws fused in-kernel (no HBM input: issue-bound, HBM fraction '
                              'undefined), device var_param',
                'e2e_mode': 'same kernel, host var_param in, host (value, grad) out',
                'roofline_mode': 'streaming instrument: the same kernel fed 335 MB of resident fp64 base draws',
                'l2': 'streaming inputs (335 MB/launch) larger than L2 (126 MB); Philox mode reads no input', 'parallelism': 'sample-sharded x%d' % self.world}

    # ---- CPU legs: fork pool, one oracle KLVI step shard per process (draws its own standard-t base draws)
    cpu_probe_units, cpu_max_units, cpu_round = 1 << 15, 1 << 22, 1 << 12

    @property
    def cpu_cores(self):
        return host_cores()

    def cpu_step(self, units):
        procs = host_cores()
        per = max(units // procs, 1)
        with blas_threads(1):
            _pool_sum(_c4_cpu_job, [(1000 + i, per) for i in range(procs)], procs)

    def cpu_literal_step(self, units):
        import torch
        from oracle import vb_autograd as va, vb_np
        fam = vb_np.mean_field_t_variational_family(10, 40)
        base = fam.impl.draw_base(np.random.RandomState(1), units)
        y = [28., 8., -3., 7., -1., 1., 18., 12.]
        sg = [15., 10., 16., 11., 9., 11., 10., 18.]
        torch.set_num_threads(host_cores())
        return va.klvi('mf_t', lambda x: va.eight_schools_logp(x, y, sg), EIGHT_SCHOOLS_LAM, base, 10, 40)

    def cpu_sample_desc(self, units):
        return ('oracle vb_np KLVI step (standard_t draws + log p + closed-form grad) on %d of the 2^22 samples, '
                'fork pool of %d processes' % (units, host_cores()))

    def config_name(self):
        return 'c4: eight-schools ncp (d=10), mean-field Student-t(df=40) KLVI, 2^22 samples/step/GPU'


# ======================================================================================= C1
C1_ITERS = 500          # AdaGrad iterations per bench step (10 CUDA-graph replays of 50 iterations)


def _c1_cpu_job(job):
    seed, iters = job
    from oracle import models_np, vb_np
    fam = vb_np.mean_field_gaussian_variational_family(2)
    obj = vb_np.black_box_klvi(fam, models_np.Funnel(), 1000)
    vb_np.adagrad_optimize(iters, obj, np.array([0., -1., 1., 1.]), learning_rate=.01, learning_rate_end=.001)
    return iters


class C1(Workload):
    """funnel d=2, mean-field Gaussian KLVI, 1000 samples per AdaGrad iteration (latency-bound; roofline n/a).
    One bench step = 500 iterations of adagrad_optimize on the device-resident loop (vb_mf_adagrad_*)."""
    name = 'c1'
    n = 1000
    default_steps, default_warmup = 10, 3

    def setup(self):
        import ctypes
        import torch
        from viabel_b200 import models, vb, _lib
        from viabel_b200._device import ptr, stream_ptr
        self.lib, self.vb, self.torch = _lib, vb, torch
        self.fam = vb.mean_field_gaussian_variational_family(2)
        self.model = models.Funnel()
        self.obj = vb.black_box_klvi(self.fam, self.model, self.n, group=False)
        self.init = np.array([0., -1., 1., 1.])                 # notebooks/funnel-distribution.ipynb:241-243
        self.units_per_step = self.n * C1_ITERS
        self.h2d_bytes = 32 + C1_ITERS * (8 + 8 + 4)            # init_param + schedule + Philox (seed, offset) arrays
        self.d2h_bytes = 8 * C1_ITERS + 8 * 4 * (C1_ITERS - 3 * C1_ITERS // 4)   # value history + kept iterates
        self.launches_per_step = C1_ITERS                       # one fused (step + update) launch per iteration
        # device-resident plan for `value` / the kernel-only timing: a long run the timed steps are slices of
        total = 1 << 18
        dev = 'cuda'
        self._param = torch.from_numpy(self.init.copy()).to(dev)
        self._rates = torch.full((total,), 0.001, dtype=torch.float64, device=dev)
        self._seeds = torch.zeros(total, dtype=torch.int64, device=dev)
        self._offsets = torch.arange(total, dtype=torch.int32, device=dev)
        self._values = torch.zeros(total, dtype=torch.float64, device=dev)
        self._kept = torch.zeros((total - 3 * total // 4, 4), dtype=torch.float64, device=dev)
        self._state = torch.zeros(_lib.lib.vb_mf_adagrad_state_bytes(2, 10), dtype=torch.uint8, device=dev)
        self._ws = torch.zeros(_lib.lib.vb_mf_step_workspace_bytes(2), dtype=torch.uint8, device=dev)
        self._plan = ctypes.c_void_p()
        _lib.check(_lib.lib.vb_mf_adagrad_create(
            _lib.FAMILY_MF_GAUSS, 0.0, self.model.handle, _lib.OBJ_KLVI, 1.0, _lib.F64, ptr(self._param), self.n, 0,
            total, 10, 0.1, ptr(self._rates), ptr(self._seeds), ptr(self._offsets), ptr(self._values),
            ptr(self._kept), ptr(self._state), self._state.numel(), ptr(self._ws), self._ws.numel(), 50, 1,
            stream_ptr(), ctypes.byref(self._plan)), 'vb_mf_adagrad_create')
        # single fused step (no update) for the kernel-only line
        self._out = torch.empty(5, dtype=torch.float64, device=dev)
        self._part = torch.empty(8, dtype=torch.float64, device=dev)
        self._args = [_lib.FAMILY_MF_GAUSS, 0.0, self.model.handle, _lib.OBJ_KLVI, 1.0, _lib.F64, ptr(self._param),
                      self.n, 0, 0, 0, 0, ptr(self._out), ptr(self._out) + 8, ptr(self._part),
                      ptr(self._ws), self._ws.numel()]
        self._off = 0

    def step_device(self):
        from viabel_b200._device import stream_ptr
        self.lib.check(self.lib.lib.vb_mf_adagrad_run(self._plan, C1_ITERS, stream_ptr()), 'vb_mf_adagrad_run')

    def step_e2e(self):
        # the public call: host init_param in, (averaged iterate, kept iterates, value history) out
        return self.vb.adagrad_optimize(C1_ITERS, self.obj, self.init, learning_rate=.01, learning_rate_end=.001)

    def kernel_only(self):
        from viabel_b200._device import stream_ptr
        self._args[10] = self._off
        self._off += 1
        self.lib.check(self.lib.lib.vb_mf_step(*self._args, stream_ptr()))

    def roofline(self, ms_kernel, hbm_gbs, bf16_tflops):
        return {'bound': 'hbm', 'achieved': 0.0, 'peak': hbm_gbs, 'unit': 'GB/s', 'frac': 0.0, 'traffic': None,
                'kernel': 'mf_step_kernel<double, MF_GAUSS, Funnel, KLVI, philox>', 'ms_per_launch': ms_kernel,
                'note': 'latency-bound (one CTA-sized problem, ~16 kB touched): roofline fraction not meaningful'}

    def config(self):
        return {'workload': self.config_name(), 'iterations_per_step': C1_ITERS,
                'value_mode': 'device-resident AdaGrad loop: CUDA-graph replays of 50 iterations, one launch per iteration '
                              '(the step kernel\'s last CTA applies the update)',
                'e2e_mode': 'adagrad_optimize(500, objective, host init_param) -> host histories, plan built per call',
                'l2': 'no resident input (Philox in-kernel)', 'parallelism': 'replicas only'}

    def config_name(self):
        return ('c1: funnel d=2, mean-field Gaussian KLVI, 1000 samples/iteration, adagrad_optimize '
                '(%d iterations per bench step)' % C1_ITERS)

    def cpu_baseline(self, budget_s=12.0, k=2):
        t0 = time.perf_counter()
        _c1_cpu_job((0, 3000))
        rate = 3000 * self.n / (time.perf_counter() - t0)
        return {'value': rate, 'unit': self.unit, 'cores': 1, 'kind': 'port',
                'sample': '3000 oracle adagrad_optimize iterations of 1000 samples (randn + funnel log p + '
                          'closed-form grad + windowed update)'}

    def reference_arm(self, steps, warmup, budget_s=75.0):
        k = steps * C1_ITERS                                    # exactly `steps` bench steps of 500 iterations
        _c1_cpu_job((0, max(warmup, 1) * 50))
        t0 = time.perf_counter()
        _c1_cpu_job((0, k))
        dt = time.perf_counter() - t0
        rate = k * self.n / dt
        cb = {'value': rate, 'unit': self.unit, 'cores': 1, 'kind': 'port',
              'sample': '%d sequential oracle adagrad_optimize iterations (the optimiser loop is sequential: 1 process)' % k}
        return {'metric': self.metric, 'value': rate, 'unit': self.unit, 'ms_per_step': dt / k * C1_ITERS * 1e3,
                'dtype': 'f64', 'config': self.config(), 'cpu_baseline': cb,
                'e2e': {'value': rate, 'unit': self.unit, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}


# ======================================================================================= C5
def c5_log_weights_host(n, seed=7):
    """lw = log of GPD(k=0.5, sigma=1) variates."""
    u = np.random.RandomState(seed).rand(n)
    return np.log(np.expm1(-0.5 * np.log1p(-u)) / 0.5)


class C5(Workload):
    """psislw + all_bounds on 1.25e8 fp64 log weights per GPU (10^9 over 8 GPUs), heavy Pareto tail."""
    name = 'c5'
    n = 125_000_000
    unit = 'weights/s'
    default_steps, default_warmup = 20, 3

    def setup(self):
        import torch
        import viabel_b200
        from viabel_b200 import psis, _lib
        from viabel_b200._device import ptr
        self.torch, self.psis, self.vb, self.lib = torch, psis, viabel_b200, _lib
        g = torch.Generator(device='cuda')
        g.manual_seed(7 + self.rank)
        u = torch.rand(self.n, dtype=torch.float64, device='cuda', generator=g)
        self.lw = torch.log(torch.expm1(-0.5 * torch.log1p(-u)) / 0.5)      # synthetic input, built once
        del u
        self.lw_host = None
        self.units_per_step = self.n
        self.h2d_bytes = 2 * self.n * 8        # psislw(lw_host) and all_bounds(lw_host) each upload the weights
        self.d2h_bytes = self.n * 8
        self.launches_per_step = 30       # profiles/r3d_launches_c5.txt (psislw; all_bounds launches nothing on a cache hit)
        self._out = torch.empty_like(self.lw)
        self._diag = torch.empty(8, dtype=torch.float64, device='cuda')
        nb = _lib.lib.vb_psis_smooth_workspace_bytes(self.n, 1.0)
        self._ws = torch.zeros(nb, dtype=torch.uint8, device='cuda')
        self._ptr = ptr

    @property
    def _group(self):
        return True if self.world > 1 else None

    def _bounds(self, lw):
        return self.vb.all_bounds(lw, moment_bound_fn=lambda p: 3.0 if p == 2 else 27.0, q_var=1.0, group=self._group)

    def step_device(self):
        out, k = self.psis.psislw(self.lw, group=self._group)
        res = self._bounds(self.lw)
        return k, res

    def step_e2e(self):
        if self.lw_host is None:
            # the step's input lives in PINNED host memory (bench contract); the result comes back into a pinned
            # buffer through psislw(out=): both directions run at link speed
            pin_in = self.torch.empty(self.n, dtype=self.torch.float64, pin_memory=True)
            pin_in.copy_(self.lw)
            self._pins = (pin_in, self.torch.empty(self.n, dtype=self.torch.float64, pin_memory=True))
            self.lw_host, self.out_host = pin_in.numpy(), self._pins[1].numpy()
        out, k = self.psis.psislw(self.lw_host, group=self._group, out=self.out_host)
        res = self._bounds(self.lw_host)
        return k, res

    def kernel_only(self):
        from viabel_b200._device import stream_ptr
        p = self._ptr
        self.lib.check(self.lib.lib.vb_psis_smooth(p(self.lw), p(self._out), self.n, 1.0, p(self._diag), 0,
                                                   p(self._ws), self._ws.numel(), stream_ptr()))

    def roofline(self, ms_kernel, hbm_gbs, bf16_tflops):
        b = 24 * self.n          # read for max/select + read for normalise + write
        ach = b / (ms_kernel * 1e-3) / 1e9
        return {'bound': 'hbm', 'achieved': ach, 'peak': hbm_gbs, 'unit': 'GB/s', 'frac': ach / hbm_gbs,
                'traffic': None, 'kernel': 'vb_psis_smooth (all passes: max, prefilter, normalise + tail work)',
                'ms_per_launch': ms_kernel, 'algorithmic_bytes_per_launch': b,
                'note': 'two streaming passes (filter: 1 read; normalise: 1 read + 1 write) = the algorithmic 24 B/weight, '
                        'measured DRAM traffic 2.95 GB; the exp of the filter pass is instruction-bound (DESIGN 6) and '
                        '~0.28 ms of small dependent tail kernels sit between the passes'}

    def config(self):
        return {'workload': 'c5: psislw + all_bounds on 1.25e8 fp64 log weights per GPU (GPD k=0.5 tail)',
                'weights_per_gpu': self.n, 'l2': 'input (1 GB) larger than L2 (126 MB)',
                'parallelism': 'sample-sharded x%d' % self.world}

    config_name = lambda self: self.config()['workload']      # noqa: E731

    # ---- CPU legs: the reference algorithm is one argsort + numpy ufunc passes, single-threaded
    cpu_probe_units, cpu_max_units, cpu_round, cpu_cores = 1_000_000, 125_000_000, 1_000_000, 1

    def cpu_prepare(self):
        if getattr(self, '_cpu_lw', None) is None:
            self._cpu_lw = c5_log_weights_host(40_000_000)      # 0.3 GB; a step uses a prefix of it

    def cpu_step(self, units):
        import warnings
        from oracle import bounds_np, psis_np
        units = min(units, self._cpu_lw.size)
        lw = self._cpu_lw[:units]
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            psis_np.psislw(lw)
            bounds_np.all_bounds(lw, moment_bound_fn=lambda p: 3.0 if p == 2 else 27.0, q_var=1.0)

    def _cpu_pick_units(self, n_steps, budget_s):
        return min(Workload._cpu_pick_units(self, n_steps, budget_s), 40_000_000)

    def cpu_sample_desc(self, units):
        return ('oracle psislw + all_bounds (restatement of notebooks/psis.py + viabel/bounds.py: full argsort + numpy '
                'passes, single-threaded as numpy is) on %d of the 1.25e8 weights per GPU' % units)


# ======================================================================================= N3
def n3_log_lik(n, m, seed=31):
    """(n, m) log-likelihood draws whose importance ratios -log_lik have GPD(k = 0.5) right tails."""
    u = np.random.RandomState(seed).rand(n, m)
    return -np.log(np.expm1(-0.5 * np.log1p(-u)) / 0.5)


class N3(Workload):
    """psisloo (notebooks/psis.py:68-109) on 4000 posterior draws x 8192 observations: one CTA per column."""
    name = 'n3'
    n, m = 4000, 8192
    unit = 'weights/s'
    default_steps, default_warmup = 20, 3
    cpu_cols = 256

    def setup(self):
        import torch
        from viabel_b200 import psis, _lib
        from viabel_b200._device import ptr
        self.torch, self.psis, self.lib, self._ptr = torch, psis, _lib, ptr
        g = torch.Generator(device='cuda')
        g.manual_seed(31 + self.rank)
        u = torch.rand((self.n, self.m), dtype=torch.float64, device='cuda', generator=g)
        self.ll = -torch.log(torch.expm1(-0.5 * torch.log1p(-u)) / 0.5)          # (n, m), as the reference takes it
        del u
        self.units_per_step = self.n * self.m
        self.h2d_bytes, self.d2h_bytes = self.n * self.m * 8, 8 * (1 + 2 * self.m)
        self.launches_per_step = 3                                                # negate+transpose, smooth, sum
        self._cols = (-self.ll).t().contiguous()
        self._out = torch.empty_like(self._cols)
        self._ks = torch.empty(self.m, dtype=torch.float64, device='cuda')
        self._loo = torch.empty(self.m, dtype=torch.float64, device='cuda')
        self._status = torch.empty(self.m, dtype=torch.int32, device='cuda')
        self._pin = None

    def step_device(self):
        return self.psis.psisloo(self.ll)

    def step_e2e(self):
        if self._pin is None:
            self._pin = self.torch.empty((self.n, self.m), dtype=self.torch.float64, pin_memory=True)
            self._pin.copy_(self.ll)
        return self.psis.psisloo(self._pin.numpy())

    def kernel_only(self):
        from viabel_b200._device import stream_ptr
        p = self._ptr
        self.lib.check(self.lib.lib.vb_psis_smooth_columns(p(self._cols), p(self._out), self.n, self.m, 1.0,
                                                           p(self._ks), p(self._loo), p(self._status), stream_ptr()))

    def roofline(self, ms_kernel, hbm_gbs, bf16_tflops):
        b = 16 * self.n * self.m         # each weight read once and written once (the second read for loo hits L2)
        ach = b / (ms_kernel * 1e-3) / 1e9
        return {'bound': 'hbm', 'achieved': ach, 'peak': hbm_gbs, 'unit': 'GB/s', 'frac': ach / hbm_gbs,
                'traffic': None, 'kernel': 'psis_columns_kernel (one CTA per column: sort + GPD fit + smoothing + loo)',
                'ms_per_launch': ms_kernel, 'algorithmic_bytes_per_launch': b,
                'note': 'per column the CTA sorts 4096 (key, index) pairs in shared memory (78 compare-exchange '
                        'steps): the kernel is bound by shared-memory sort latency, not by HBM'}

    def config(self):
        return {'workload': self.config_name(), 'draws': self.n, 'columns': self.m,
                'l2': 'input + output (524 MB) larger than L2', 'parallelism': 'columns sharded x%d' % self.world}

    def config_name(self):
        return 'n3: psisloo on 4000 draws x 8192 observations (GPD k=0.5 ratios), one CTA per column'

    def _cpu(self, cols):
        import warnings
        from oracle import psis_np
        ll = n3_log_lik(self.n, cols)
        t0 = time.perf_counter()
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            psis_np.psisloo(ll)
        return self.n * cols / (time.perf_counter() - t0)

    def cpu_baseline(self, budget_s=None, k=None):       # fixed sample (a few seconds); the budget arguments are accepted
        rate = self._cpu(self.cpu_cols)
        return {'value': rate, 'unit': self.unit, 'cores': 1, 'kind': 'port',
                'sample': 'oracle psisloo on %d of the %d columns (4000 draws each), single-threaded numpy'
                          % (self.cpu_cols, self.m)}

    def reference_arm(self, steps, warmup, budget_s=None):
        k = max(1, min(steps, 3))
        t0 = time.perf_counter()
        rates = [self._cpu(self.cpu_cols) for _ in range(k)]
        dt = (time.perf_counter() - t0) / k
        rate = float(np.mean(rates))
        cb = {'value': rate, 'unit': self.unit, 'cores': 1, 'kind': 'port',
              'sample': 'oracle psisloo on %d columns x 4000 draws per step (argsort + GPD fit per column)' % self.cpu_cols}
        return {'metric': self.metric, 'value': rate, 'unit': self.unit, 'ms_per_step': dt * 1e3, 'dtype': 'f64',
                'config': {'workload': self.config_name()}, 'cpu_baseline': cb,
                'e2e': {'value': rate, 'unit': self.unit, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}


class N4R(Workload):
    """Convergence diagnostics of the R-hat optimisers (viabel/functions.py) over a C3-sized iterate history:
    2 restarts x 2000 iterates x 131 840 parameters (4.2 GB fp64): windowed split R-hat (8 windows of 250),
    halfway R-hat of the whole history, iterate averages of the last 500 iterates of both restarts."""
    name = 'n4r'
    chains, iters, params = 2, 2000, 131840
    window = 250
    unit = 'iterate values/s'
    default_steps, default_warmup = 5, 3
    cpu_params = 4096

    def setup(self):
        import torch
        from viabel_b200 import functions, _lib
        from viabel_b200._device import ptr
        self.torch, self.fn, self.lib, self._ptr = torch, functions, _lib, ptr
        g = torch.Generator(device='cuda')
        g.manual_seed(41 + self.rank)
        self.hist = torch.randn((self.chains, self.iters, self.params), dtype=torch.float64, device='cuda', generator=g)
        self.hist += torch.linspace(1.0, 0.0, self.iters, dtype=torch.float64, device='cuda')[None, :, None] ** 4
        self.units_per_step = self.chains * self.iters * self.params
        self.h2d_bytes = self.units_per_step * 8
        self.d2h_bytes = 8 * self.params * (self.iters // self.window + 1)
        self.launches_per_step = 2 + 2 + 2 * self.chains
        nw = self.iters // self.window
        self._win = torch.tensor([[i * self.window, self.window] for i in range(nw)], dtype=torch.int64, device='cuda')
        self._vh = torch.empty((nw, self.params), dtype=torch.float64, device='cuda')
        self._rh = torch.empty((nw, self.params), dtype=torch.float64, device='cuda')
        nb = _lib.lib.vb_rhat_workspace_bytes(self.chains, self.params, nw)
        self._ws = torch.zeros(nb, dtype=torch.uint8, device='cuda')
        self._pin = None

    def _diagnose(self, hist):
        r = self.fn.compute_R_hat_adaptive_numpy(hist, window_size=self.window)
        h = self.fn.compute_R_hat(hist, warmup=self.iters // 2)[1]
        dev = hist if hasattr(hist, 'is_cuda') else None
        avgs = []
        if dev is not None:
            for o in range(self.chains):
                avgs.append(self.fn.stochastic_iterate_averaging(dev[o], self.iters - 500)[1])
        return r, h, avgs

    def step_device(self):
        return self._diagnose(self.hist)

    def step_e2e(self):
        if self._pin is None:
            self._pin = self.torch.empty(self.hist.shape, dtype=self.torch.float64, pin_memory=True)
            self._pin.copy_(self.hist)
        dev = self._pin.to('cuda', non_blocking=True)                  # the history is uploaded once per step
        r, h, avgs = self._diagnose(dev)
        return r.cpu(), h.cpu(), [a.cpu() for a in avgs]

    def kernel_only(self):
        from viabel_b200._device import stream_ptr
        p = self._ptr
        self.lib.check(self.lib.lib.vb_rhat_windows(p(self.hist), self.chains, self.iters, self.params, p(self._win),
                                                    self._win.shape[0], p(self._vh), p(self._rh), p(self._ws),
                                                    self._ws.numel(), stream_ptr()))

    def roofline(self, ms_kernel, hbm_gbs, bf16_tflops):
        b = 8 * self.units_per_step      # every iterate value read once (the second pass about the mean hits L2)
        ach = b / (ms_kernel * 1e-3) / 1e9
        return {'bound': 'hbm', 'achieved': ach, 'peak': hbm_gbs, 'unit': 'GB/s', 'frac': ach / hbm_gbs,
                'traffic': None, 'kernel': 'rhat_halfchain_kernel + rhat_finish_kernel (vb_rhat_windows, 8 windows)',
                'ms_per_launch': ms_kernel, 'algorithmic_bytes_per_launch': b,
                'note': 'two-pass variance as the reference computes it: a CTA re-reads its 125 x 256 B slab from L2'}

    def config(self):
        return {'workload': self.config_name(), 'chains': self.chains, 'iterates': self.iters, 'params': self.params,
                'l2': 'history (4.2 GB) larger than L2', 'parallelism': 'replicas x%d' % self.world}

    def config_name(self):
        return ('n4r: windowed + halfway split R-hat and iterate averages of a 2 x 2000 x 131840 fp64 iterate '
                'history (C3 parameter count)')

    def _cpu(self, params):
        from oracle import functions_np as ofn
        rs = np.random.RandomState(5)
        hist = rs.randn(self.chains, self.iters, params) + np.linspace(1.0, 0.0, self.iters)[None, :, None] ** 4
        t0 = time.perf_counter()
        ofn.compute_R_hat_adaptive_numpy(hist, window_size=self.window)
        ofn.compute_R_hat(hist, warmup=self.iters // 2)
        for o in range(self.chains):
            ofn.stochastic_iterate_averaging(hist[o], self.iters - 500)
        return self.chains * self.iters * params / (time.perf_counter() - t0)

    def cpu_baseline(self, budget_s=None, k=None):
        rate = self._cpu(self.cpu_params)
        return {'value': rate, 'unit': self.unit, 'cores': 1, 'kind': 'port',
                'sample': 'oracle diagnostics on %d of the %d parameters, single-threaded numpy'
                          % (self.cpu_params, self.params)}

    def reference_arm(self, steps, warmup, budget_s=None):
        k = max(1, min(steps, 3))
        t0 = time.perf_counter()
        rates = [self._cpu(self.cpu_params) for _ in range(k)]
        dt = (time.perf_counter() - t0) / k
        rate = float(np.mean(rates))
        cb = {'value': rate, 'unit': self.unit, 'cores': 1, 'kind': 'port',
              'sample': 'oracle diagnostics on %d parameters x %d x %d iterates per step' % (self.cpu_params, self.chains, self.iters)}
        return {'metric': self.metric, 'value': rate, 'unit': self.unit, 'ms_per_step': dt * 1e3, 'dtype': 'f64',
                'config': {'workload': self.config_name()}, 'cpu_baseline': cb,
                'e2e': {'value': rate, 'unit': self.unit, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}


WORKLOADS = {'c1': C1, 'c4': C4, 'c5': C5, 'n3': N3, 'n4r': N4R}



# ======================================================================================= C3
def c3_problem(d=512):
    """Sigma_p = A A'/d + I (seed 4), mu_p (seed 5), var_param = 0.1 N(0,1) (seed 6): SURVEY 8(d) C3."""
    A = np.random.RandomState(4).randn(d, d)
    Sp = A @ A.T / d + np.eye(d)
    mp = np.random.RandomState(5).randn(d)
    lam = 0.1 * np.random.RandomState(6).randn(d + d * (d + 1) // 2)
    return mp, np.linalg.inv(Sp), lam


class C3(Workload):
    """CHIVI (alpha 2), full-rank Student-t (df 40), d = 512, correlated-Gaussian target, 2^16 samples/step."""
    name = 'c3'
    n, d = 1 << 16, 512
    dtype = 'f32 storage, 3xTF32 tensor-core products, fp64 accumulation of reductions'
    default_steps, default_warmup = 10, 3
    scale = 'sqrtm'

    def setup(self):
        import torch
        from viabel_b200 import models, vb, _lib
        from viabel_b200._device import ptr
        self.torch, self.lib = torch, _lib
        mp, Pp, lam = c3_problem(self.d)
        self.model = models.CorrelatedGaussian(mp, Pp)
        self.fam = vb.t_variational_family(self.d, 40, dtype='f32', scale=self.scale)
        self.obj = vb.black_box_chivi(2.0, self.fam, self.model, self.n * self.world, group=self.world > 1)
        self.base = self.fam.impl.base_draws(self.n, 0, 0, index0=self.rank * self.n)
        self.lam_host = lam
        self.lam_dev = torch.from_numpy(lam).cuda()
        self.units_per_step = self.n
        self.h2d_bytes = lam.nbytes
        self.d2h_bytes = 8 * (1 + lam.size)
        self.launches_per_step = 17       # profiles/r3g_launches_c3.txt: 2 draws, 2 chain, 3 GEMMs + 10 small kernels per step
        # pre-split operands for the kernel-only timing: X (n, d) = Z (n, d) S (d, d)
        lib = _lib.lib
        z = self.base['z']
        S = torch.randn(self.d, self.d, dtype=torch.float32, device='cuda')
        self._zs = torch.empty(lib.vb_split_bytes(self.n, self.d) // 4, dtype=torch.float32, device='cuda')
        self._ss = torch.empty(lib.vb_split_bytes(self.d, self.d) // 4, dtype=torch.float32, device='cuda')
        _lib.check(lib.vb_split_f32(ptr(z), self.n, self.d, self.d, ptr(self._zs), 0))
        _lib.check(lib.vb_split_f32(ptr(S), self.d, self.d, self.d, ptr(self._ss), 0))
        self._c = torch.empty((self.n, self.d), dtype=torch.float32, device='cuda')
        self._ptr = ptr
        g = torch.Generator(device='cuda')
        g.manual_seed(3)
        self._GS = torch.randn((self.d, self.d), dtype=torch.float64, device='cuda', generator=g)
        self._gmu = torch.randn(self.d, dtype=torch.float64, device='cuda', generator=g)
        self._rows = torch.tensor([0.0, float(self.n) / 8, -1.0, -2.0], dtype=torch.float64, device='cuda')
        self._out = torch.empty(1 + lam.size, dtype=torch.float64, device='cuda')

    def step_device(self):
        # the draw is part of the step (vb.py:258 -> :203-208): Philox chi-square + normals generated on the device
        return self.obj(self.lam_dev)

    def step_e2e(self):
        return self.obj(self.lam_host)

    def kernel_only(self):
        """The step's dominant kernels: the O(d^3) chain pair (fr_chain_forward_kernel + fr_chain_backward_kernel)."""
        from viabel_b200._device import stream_ptr
        impl, lib, p, d = self.fam.impl, self.lib.lib, self._ptr, self.d
        ch = impl.chain_kernel(self.lam_dev)
        ws = impl.ws.get('chain', lib.vb_fr_chain_workspace_bytes(d))
        self.lib.check(lib.vb_fr_chain_backward(d, 0 if self.scale == 'sqrtm' else 1, 1, 2.0, float(self.n), p(ch['L']),
                                                p(ch['S']), p(self._GS), p(self._gmu), p(self._rows), p(ch['scal']),
                                                p(self._out), p(ws), ws.numel(), stream_ptr()), 'vb_fr_chain_backward')

    def roofline(self, ms_kernel, hbm_gbs, bf16_tflops):
        info = self.fam.impl.chain_info()       # [polar, inverse, smith iterations, status, forward GEMMs, backward GEMMs]
        gemms, sym = info[4] + info[5], info[6] + info[7]      # counted by the kernels themselves
        nb = (self.d + 31) // 32
        sym_frac = (nb * (nb + 1) / 2) / float(nb * nb)          # a symmetric product computes the lower-triangular blocks
        flops = ((gemms - sym) + sym * sym_frac) * 2.0 * self.d ** 3
        ach = flops / (ms_kernel * 1e-3) / 1e12 if ms_kernel > 0 else 0.0
        # fp64 MMA peak of THIS GPU, measured live by the library's own microbenchmark (DMMA.8x8x4 on register operands)
        import ctypes
        import torch
        pk = ctypes.c_double(0.0)
        nb = self.lib.lib.vb_measure_dmma_peak_workspace_bytes()
        pws = torch.empty(nb, dtype=torch.uint8, device='cuda')
        self.lib.check(self.lib.lib.vb_measure_dmma_peak(ctypes.byref(pk), pws.data_ptr(), nb, None), 'vb_measure_dmma_peak')
        peak = pk.value
        del pws
        return {'bound': 'tensor', 'achieved': ach, 'peak': peak, 'unit': 'TFLOP/s', 'frac': ach / peak, 'traffic': None,
                'peak_kind': 'f64', 'peak_source': 'measured live: vb_measure_dmma_peak (back-to-back DMMA.8x8x4, all SMs, '
                                                   'best of 4; nominal 40 TFLOP/s); MEASURED_PEAKS.json has no fp64 figure',
                'kernel': 'fr_chain_forward_kernel + fr_chain_backward_kernel (O(d^3) chain: polar Newton-Schulz sqrtm, '
                          'squared-Smith Sylvester solve, packing; fp64 mma.sync m8n8k4)',
                'ms_per_launch': ms_kernel, 'executed_gemms': gemms, 'symmetric_gemms': sym, 'iterations': info[:4],
                'executed_flops_per_launch': flops,
                'note': 'the chain pair is the dominant kernel of the step (see share_of_step); `achieved` counts the '
                        'EXECUTED fp64 GEMM flops (%d GEMMs of 512^3), not an algorithmic minimum: the iteration trades '
                        'flops for having no eigendecomposition (cuSOLVER syevd took 4.7 ms here).  The per-sample products '
                        '(z S, (x - mu_p) P_p, Z\' diag(w) G: 6 d^2 flop per sample) run on tcgen05 3xTF32 at 0.14 / 0.19 / '
                        '0.14 ms per launch (profiles/r2u_launches_c3.txt); symmetric products (52 of 73 at C3) execute half of 2 d^3 and '
                        'are counted as executed' % gemms}

    def config(self):
        return {'workload': self.config_name(), 'samples_per_step_per_gpu': self.n, 'scale': self.scale,
                'value_mode': 'streaming: fp32 base draws (134 MB) resident in HBM',
                'e2e_mode': 'Philox draws, host var_param (131840 doubles) in, host (value, grad) out',
                'l2': 'inputs (134 MB z + intermediates) larger than L2', 'parallelism': 'sample-sharded x%d' % self.world}

    def config_name(self):
        return 'c3: CHIVI alpha=2, full-rank Student-t(df=40) d=512, correlated-Gaussian target, 2^16 samples/step/GPU'

    # ---- CPU legs: closed-form oracle CHIVI step (numpy + multithreaded BLAS), O(d^3) chain included once per step
    cpu_probe_units, cpu_max_units, cpu_round = 1 << 11, 1 << 16, 1 << 10

    @property
    def cpu_cores(self):
        return host_cores()

    def cpu_prepare(self):
        if getattr(self, '_cpu_in', None) is None:
            from oracle import models_np, vb_np
            mp, Pp, lam = c3_problem(self.d)
            fam = vb_np.t_variational_family(self.d, 40, scale=self.scale)
            self._cpu_in = (fam, models_np.CorrelatedGaussian(mp, Pp), lam, mp, Pp)

    def cpu_step(self, units):
        from oracle import vb_np
        fam, model, lam, _, _ = self._cpu_in
        base = fam.impl.draw_base(np.random.RandomState(1), units)      # the draw is part of the step (vb.py:204-206)
        vb_np.chivi_value_and_grad(fam.impl, model, 2.0, lam, base)

    def cpu_literal_step(self, units):
        if self.scale != 'sqrtm':
            return None
        import torch
        from oracle import vb_autograd as va
        fam, _, lam, mp, Pp = self._cpu_in
        base = fam.impl.draw_base(np.random.RandomState(1), units)
        torch.set_num_threads(host_cores())
        return va.chivi('fr_t', lambda x: va.gauss_logp(x, mp, Pp), 2.0, lam, base, self.d, 40)

    def cpu_sample_desc(self, units):
        return ('oracle closed-form CHIVI step (chi2 + normal draws, z S, log p + grad, weighted outer product, eigh / '
                'Sylvester chain) on %d of the 2^16 samples, numpy with %d BLAS threads' % (units, host_cores()))


class C3Chol(C3):
    """Same as c3 with the documented same-distribution fast mode x = mu + L ztilde (no eigh / Sylvester)."""
    name = 'c3chol'
    scale = 'chol'

    def config_name(self):
        return C3.config_name(self) + ' [scale=chol: deviation from the reference per-draw map]'


# ======================================================================================= C2
def c2_problem(N=1 << 20, d=128):
    """X ~ N(0,1)/sqrt(d) (seed 1), beta* (seed 2), y = X beta* + t_40 centred, var_param (seed 3)."""
    X = (np.random.RandomState(1).randn(N, d) / np.sqrt(d)).astype(np.float32)
    bt = np.random.RandomState(2).randn(d)
    y = X.astype(np.float64) @ bt + np.random.RandomState(12).standard_t(40, N)
    y = (y - y.mean()).astype(np.float32)
    lam = np.concatenate([bt + 0.01 * np.random.RandomState(3).randn(d), np.full(d, np.log(1e-3))])
    return X, y, lam


class C2(MeanFieldStages, Workload):
    """Student-t robust regression N = 2^20 rows, d = 128; KLVI, mean-field Student-t (df 40), 4096 samples/step."""
    name = 'c2'
    n, N, d = 4096, 1 << 20, 128
    dtype = 'f32 storage; tcgen05 fp16 (hi, lo) x 3 products with fp32 accumulate (fp32-accurate: 22 significant bits per operand); fp64 final sums'
    compute = 'f32'
    default_steps, default_warmup = 5, 3

    def setup(self):
        import torch
        from viabel_b200 import models, vb, _lib
        self.lib = _lib
        X, y, lam = c2_problem(self.N, self.d)
        self.model = models.StudentTRegression(X, y, df=40.0, dtype='f32')
        self.fam = vb.mean_field_t_variational_family(self.d, 40, dtype=self.compute)
        self.obj = vb.black_box_klvi(self.fam, self.model, self.n * self.world, group=self.world > 1)
        self.eps = self.fam.impl.base_draws(self.n, 0, 0, index0=self.rank * self.n)['eps']
        self.lam_host = lam
        self.lam_dev = torch.from_numpy(lam).cuda()
        self.units_per_step = self.n
        self.h2d_bytes, self.d2h_bytes = lam.nbytes, 8 * (1 + lam.size)
        self.launches_per_step = 8        # base draws, mf_sample, y prescale, fused kernel, final reduce, weights, colsum, finish
        self._x = torch.from_numpy(np.tile(lam[:self.d], (self.n, 1))).cuda()
        if self.compute == 'f32':
            self._x = self._x.float()

    def step_device(self):
        # the draw is part of the step (vb.py:239 -> :54-57): Philox normals generated on the device
        return self.obj(self.lam_dev)

    def step_e2e(self):
        return self.obj(self.lam_host)

    def kernel_only(self):
        code = self.lib.F64 if self.compute == 'f64' else self.lib.F32
        self.model.logp_and_grad_device(self._x, code, True)

    def roofline(self, ms_kernel, hbm_gbs, bf16_tflops):
        flops = 4.0 * self.N * self.d * self.n
        ach = flops / (ms_kernel * 1e-3) / 1e12
        if self.compute == 'f64':
            return {'bound': 'tensor', 'achieved': ach, 'peak': bf16_tflops, 'unit': 'TFLOP/s', 'frac': ach / bf16_tflops,
                    'traffic': None, 'kernel': 'regression fp64 SIMT path (DFMA pipe, parity instrument)',
                    'ms_per_launch': ms_kernel, 'algorithmic_flops_per_launch': flops, 'peak_kind': 'f16',
                    'note': 'fp64 path runs on the DFMA pipe, not the tensor pipe; peak shown for reference only'}
        fused = os.environ.get('VB_REG_PATH', '0') != '1'
        if fused:
            return {'bound': 'tensor', 'achieved': ach, 'peak': bf16_tflops, 'unit': 'TFLOP/s', 'frac': ach / bf16_tflops,
                    'traffic': None, 'kernel': 'reg_fused_kernel (one persistent launch: GEMM1 B X\' -> residual / log1p / psi '
                    'epilogue in registers -> GEMM2 Psi X, Psi hi/lo handed over in tensor memory)',
                    'ms_per_launch': ms_kernel, 'algorithmic_flops_per_launch': flops, 'peak_kind': 'f16',
                    'peak_source': 'MEASURED_PEAKS.json bf16_tflops (driver-written cuBLAS bf16 burst; fp16 MMAs run at the bf16 rate)',
                    'note': 'timed unit = the whole plugin call vb_model_logp_grad (y prescale + fused kernel + final '
                            'reduction); tcgen05 kind::f16, fp16 (hi, lo) operand pairs = 3 MMAs per algorithmic product, '
                            'so frac <= 1/3 by construction and the tensor-pipe rate is 3 x achieved'}
        return {'bound': 'tensor', 'achieved': ach, 'peak': bf16_tflops / 2.0, 'unit': 'TFLOP/s', 'frac': ach / (bf16_tflops / 2.0),
                'traffic': None, 'kernel': 'gemm_tf32x3_kernel<PsiEpi> + gemm_tf32x3_kernel<SplitKEpi> (two-GEMM 3xTF32 '
                'pipeline, VB_REG_PATH=1)', 'ms_per_launch': ms_kernel, 'algorithmic_flops_per_launch': flops,
                'note': 'A/B arm: Psi round-trips through HBM; tf32 peak replaced by the live measurement in bench.py'}

    def config(self):
        return {'workload': self.config_name(), 'samples_per_step_per_gpu': self.n, 'rows': self.N, 'compute': self.compute,
                'l2': 'design matrix (512 MB fp32) larger than L2', 'parallelism': 'sample-sharded x%d' % self.world}

    def config_name(self):
        return 'c2: Student-t regression N=2^20 rows d=128, KLVI mean-field Student-t(df=40), 4096 samples/step/GPU'

    # ---- CPU legs: closed-form oracle KLVI step on a bounded number of the 4096 MC samples over ALL 2^20 rows, so
    # that samples/s is measured directly (the per-sample cost 4 N d flop does not depend on the batch)
    cpu_probe_units, cpu_max_units, cpu_round = 16, 4096, 16

    @property
    def cpu_cores(self):
        return host_cores()

    def cpu_prepare(self):
        if getattr(self, '_cpu_in', None) is None:
            from oracle import models_np, vb_np
            X, y, lam = c2_problem(self.N, self.d)
            model = models_np.StudentTRegression(X.astype(np.float64), y.astype(np.float64), 40.0, row_block=1 << 13)
            self._cpu_in = (vb_np.mean_field_t_variational_family(self.d, 40), model, lam)

    def cpu_step(self, units):
        from oracle import vb_np
        fam, model, lam = self._cpu_in
        base = fam.impl.draw_base(np.random.RandomState(1), units)
        vb_np.klvi_value_and_grad(fam.impl, model, lam, base)

    def cpu_literal_step(self, units):
        import torch
        from oracle import vb_autograd as va
        fam, model, lam = self._cpu_in
        base = fam.impl.draw_base(np.random.RandomState(1), units)
        torch.set_num_threads(host_cores())
        Xt, yt = torch.from_numpy(model.X), torch.from_numpy(model.y)
        return va.klvi('mf_t', lambda b: va.regression_logp(b, Xt, yt, 40.0), lam, base, self.d, 40)

    def cpu_sample_desc(self, units):
        return ('oracle closed-form KLVI step on %d of the 4096 MC samples x all 2^20 rows (standard_t draws, X b, psi, '
                "X' psi in fp64), numpy with %d BLAS threads" % (units, host_cores()))


class C2F64(C2):
    """Same as c2 in full fp64 (SIMT DFMA path): the parity instrument, not the throughput path."""
    name = 'c2f64'
    dtype = 'f64'
    compute = 'f64'
    default_steps, default_warmup = 2, 3

    def config_name(self):
        return C2.config_name(self) + ' [fp64 SIMT parity path]'


WORKLOADS.update({'c2': C2, 'c2f64': C2F64, 'c3': C3, 'c3chol': C3Chol})
