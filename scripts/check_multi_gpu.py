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
    # larger weight vector: more than 8192 candidates in total, so the histogram-refinement exchange
    # (refine_hist / refine_select all-reduce) runs over NCCL; tail permutation must equal the oracle's
    n3 = 4_000_000
    u = np.random.RandomState(10).rand(n3)
    lw3 = np.log(np.expm1(-0.5 * np.log1p(-u)) / 0.5)
    s3, c3 = vb._shard(n3, rank, world)
    shard3 = torch.from_numpy(lw3[s3:s3 + c3].copy()).cuda()
    out3, k3 = psis.psislw(shard3, group=True)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        oout3, ok3 = psis_np.psislw(lw3, sort_kind='stable')
    np.testing.assert_allclose(k3, ok3, rtol=1e-9)
    np.testing.assert_allclose(out3.cpu().numpy(), oout3[s3:s3 + c3], rtol=1e-9, atol=1e-9)

    # sharded full-rank CHIVI / KLVI (all_reduce(max) + rescale + one all-gather of [sums, g_mu, G_S])
    dfr, nfr = 24, 20_011
    rs = np.random.RandomState(3)
    A = rs.randn(dfr, dfr)
    mp_, Pp = rs.randn(dfr), np.linalg.inv(A @ A.T / dfr + np.eye(dfr))
    for scale in ('sqrtm', 'chol'):
        ofr = vb_np.t_variational_family(dfr, 40, scale=scale)
        lam_fr = 0.1 * np.random.RandomState(6).randn(ofr.impl.n_param)
        full = ofr.impl.draw_base(np.random.RandomState(4), nfr)
        sf, cf = vb._shard(nfr, rank, world)
        mine_fr = {k_: v_[sf:sf + cf] for k_, v_ in full.items()}
        fr = vb.t_variational_family(dfr, 40, scale=scale)
        gm, ogm = models.CorrelatedGaussian(mp_, Pp), mm.CorrelatedGaussian(mp_, Pp)
        v, g = vb.black_box_chivi(2.0, fr, gm, nfr, group=True)(lam_fr, base_draws=mine_fr)
        ov, og = vb_np.chivi_value_and_grad(ofr.impl, ogm, 2.0, lam_fr, full)
        np.testing.assert_allclose(v, ov, rtol=1e-10, err_msg='fr chivi ' + scale)
        assert np.max(np.abs(g - og)) <= 1e-9 * np.max(np.abs(og)), 'fr chivi grad ' + scale
        v, g = vb.black_box_klvi(fr, gm, nfr, group=True)(lam_fr, base_draws=mine_fr)
        ov, og = vb_np.klvi_value_and_grad(ofr.impl, ogm, lam_fr, full)
        np.testing.assert_allclose(v, ov, rtol=1e-10, err_msg='fr klvi ' + scale)
        assert np.max(np.abs(g - og)) <= 1e-9 * np.max(np.abs(og)), 'fr klvi grad ' + scale

    # sharded all_bounds with samples (column means, central moments and covariance merged across ranks)
    ns, ds = 50_001, 7
    smp = np.random.RandomState(12).randn(ns, ds) * np.arange(1, ds + 1) + 0.5
    lws = np.random.RandomState(13).randn(ns) * 0.3 - 1.0
    ss, cs = vb._shard(ns, rank, world)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        res = bounds.all_bounds(torch.from_numpy(lws[ss:ss + cs].copy()).cuda(),
                                samples=torch.from_numpy(smp[ss:ss + cs].copy()).cuda(), group=True)
        ores = bounds_np.all_bounds(lws, samples=smp)
    for key in ores:
        np.testing.assert_allclose(res[key], ores[key], rtol=1e-9, err_msg='all_bounds ' + key)
    dist.barrier()
    if rank == 0:
        print('MULTI_GPU_OK world=%d' % world)
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
