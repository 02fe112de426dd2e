"""Multi-rank parity check, run under torchrun (one process per GPU, NCCL):
    torchrun --nproc-per-node 2 scripts/check_multi_gpu.py
Every rank holds a contiguous shard of the SAME externally supplied base draws / log weights; the sharded results
(partials exchanged by all-gather, merged in rank order) must equal the oracle's unsharded ones.  Used by
tests/test_gpu_multi.py when at least two GPUs are visible."""
import os
import sys
import warnings

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
    if os.environ.get('NCCL_DEBUG', '').upper() == 'VERSION':
        os.environ.pop('NCCL_DEBUG', None)
    dist.init_process_group('nccl', device_id=torch.device('cuda', int(os.environ['LOCAL_RANK'])))
    from oracle import models_np as mm
    from oracle import bounds_np, psis_np, vb_np
    from viabel_b200 import bounds, models, psis, vb

    d, n = 10, 100_003
    lam = np.concatenate([0.3 * np.random.RandomState(0).randn(d), np.full(d, -0.1)])
    ofam = vb_np.mean_field_t_variational_family(d, 40)
    base = ofam.impl.draw_base(np.random.RandomState(1), n)
    start, count = vb._shard(n, rank, world)
    mine = {'eps': base['eps'][start:start + count]}
    fam, model, omodel = vb.mean_field_t_variational_family(d, 40), models.EightSchoolsNCP(), mm.EightSchoolsNCP()
    checks = []
    v, g = vb.black_box_klvi(fam, model, n, group=True)(lam, base_draws=mine)
    ov, og = vb_np.klvi_value_and_grad(ofam.impl, omodel, lam, base)
    checks.append(('klvi', v, ov, g, og))
    v, g = vb.black_box_chivi(2.0, fam, model, n, group=True)(lam, base_draws=mine)[:2]
    ov, og = vb_np.chivi_value_and_grad(ofam.impl, omodel, 2.0, lam, base)
    checks.append(('chivi', v, ov, g, og))
    v, g = vb.black_box_klvi_pd(fam, model, n, group=True)(lam, base_draws=mine)
    ov, og = vb_np.klvi_pd_value_and_grad(ofam.impl, omodel, lam, base)
    checks.append(('klvi_pd', v, ov, g, og))
    for name, v, ov, g, og in checks:
        np.testing.assert_allclose(v, ov, rtol=1e-10, err_msg=name)
        np.testing.assert_allclose(g, og, rtol=1e-10, atol=1e-12, err_msg=name)

    # sharded psislw and divergence bound on a heavy-tailed weight vector
    n2 = 300_000
    u = np.random.RandomState(9).rand(n2)
    lw = np.log(np.expm1(-0.5 * np.log1p(-u)) / 0.5)
    s2, c2 = vb._shard(n2, rank, world)
    shard = torch.from_numpy(lw[s2:s2 + c2].copy()).cuda()
    out, k = psis.psislw(shard, group=True)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        oout, ok = psis_np.psislw(lw, sort_kind='stable')
        od2 = bounds_np.divergence_bound(lw, alpha=2.0)
        d2 = bounds.divergence_bound(shard, alpha=2.0, group=True)
    np.testing.assert_allclose(k, ok, rtol=1e-9)
    np.testing.assert_allclose(out.cpu().numpy(), oout[s2:s2 + c2], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(d2, od2, rtol=1e-10)
    dist.barrier()
    if rank == 0:
        print('MULTI_GPU_OK world=%d' % world)
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
