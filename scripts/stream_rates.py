"""Achieved HBM rates of the streaming kernels on 1.25e8 fp64 weights (1 GB): a pure read (vb_max), the PSIS
filter pass, the divergence pass and a torch copy, timed alone with CUDA events.  Answers whether the exp-carrying
passes are limited by memory or by instruction issue."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from viabel_b200 import _lib                      # noqa: E402
from viabel_b200._device import ptr, stream_ptr   # noqa: E402


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    n = 125_000_000
    g = torch.Generator(device='cuda')
    g.manual_seed(7)
    u = torch.rand(n, dtype=torch.float64, device='cuda', generator=g)
    lw = torch.log(torch.expm1(-0.5 * torch.log1p(-u)) / 0.5)
    out = torch.empty_like(lw)
    lib = _lib.lib
    res = torch.empty(8, dtype=torch.float64, device='cuda')
    ws = torch.zeros(max(lib.vb_max_workspace_bytes(n), lib.vb_bounds_divergence_workspace_bytes(n),
                         lib.vb_logsumexp_workspace_bytes(n)) + 4096, dtype=torch.uint8, device='cuda')
    rows = []
    rows.append(('torch copy (read + write)', timed(lambda: out.copy_(lw)), 16))
    rows.append(('vb_max (read, fmax)', timed(lambda: lib.vb_max(ptr(lw), n, ptr(res), ptr(ws), ws.numel(), stream_ptr())), 8))
    rows.append(('vb_logsumexp (max pass + exp pass)', timed(lambda: lib.vb_logsumexp(ptr(lw), n, ptr(res), ptr(ws), ws.numel(), stream_ptr())), 16))
    rows.append(('vb_bounds_divergence (read, exp + 5 sums)', timed(lambda: lib.vb_bounds_divergence(ptr(lw), n, 2.0, ptr(res), ptr(ws), ws.numel(), stream_ptr())), 8))
    pivot = torch.tensor([float(torch.quantile(lw[:1 << 20], 0.99))], dtype=torch.float64, device='cuda')
    cap = 1 << 22
    cand = torch.empty((cap, 2), dtype=torch.int64, device='cuda')
    cnt = torch.zeros(1, dtype=torch.int64, device='cuda')
    bm = torch.empty(2, dtype=torch.float64, device='cuda')
    fws = torch.zeros(lib.vb_psis_filter_workspace_bytes(n), dtype=torch.uint8, device='cuda')
    rows.append(('vb_psis_filter (read, exp + compaction)', timed(lambda: lib.vb_psis_filter(
        ptr(lw), n, 0, ptr(pivot), ptr(cand), cap, ptr(cnt), ptr(bm), ptr(fws), fws.numel(), stream_ptr())), 8))
    rows.append(('torch sum (read)', timed(lambda: lw.sum()), 8))
    for name, ms, bpe in rows:
        print('%-45s %8.3f ms  %7.1f GB/s' % (name, ms, bpe * n / ms / 1e6))


if __name__ == '__main__':
    main()
