"""world_size-2 gloo tests (CPU) of the sample-axis sharding logic: the shard arithmetic the
product uses (vb._shard / vb._dist_info), the exchange pattern (one all_gather of the
VB_MF_NPART partial vector per step; one all_gather of 6 doubles for divergence_bound) and the
fixed-order merge formulas, checked against the unsharded oracle.  The partials themselves are
produced by the oracle here; on the GPU box the same exchange carries kernel partials
(tests/test_gpu_multi.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import models_np as mm
from oracle import vb_np


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def mf_partials(impl, model, objective, alpha, lam, base):
    """numpy restatement of the partial vector vb_mf_step writes (include/viabel_b200.h)."""
    d = impl.dim
    x = impl.transform(lam, base)
    lp, g = model.logp_and_grad(x)
    eps = base['eps']
    out = np.zeros(2 * d + 4)
    if objective == 'klvi_pd':                        # VB_OBJ_KLVI_PD: the CHIVI pass with alpha = 0 (unit weights)
        objective, alpha = 'chivi', 0.0
    if objective == 'klvi':
        w = np.ones(len(lp))
        out[0], out[2] = 0.0, lp.sum()
    else:
        lw = lp - impl.logq_of_base(lam, base)
        out[0] = lw.max()
        w = np.exp(alpha * (lw - out[0]))
        out[2] = lw.sum()
    out[1], out[3] = w.sum(), len(lp)
    out[4:4 + d] = w @ g
    out[4 + d:] = w @ (g * eps)
    return out


def mf_finalize(kind, objective, alpha, lam, parts):
    """numpy restatement of vb_mf_finalize: rescale to the global max, sum in rank order, Appendix A.1."""
    parts = np.asarray(parts)
    d = (parts.shape[1] - 4) // 2
    pd = objective == 'klvi_pd'
    if pd:
        objective, alpha = 'chivi', 0.0
    M = parts[:, 0].max() if objective == 'chivi' else 0.0
    tot = np.zeros(parts.shape[1])
    for p in parts:
        f = np.exp(alpha * (p[0] - M)) if objective == 'chivi' else 1.0
        tot[1] += p[1] * f
        tot[2:4] += p[2:4]
        tot[4:] += p[4:] * f
    n = tot[3]
    sig = np.exp(lam[d:])
    gm, gl = tot[4:4 + d], tot[4 + d:] * sig
    if pd:                                            # value from the summed log weights, gradient as KLVI
        return -(tot[2] / n), np.concatenate([-gm / n, -(1 + gl / n)])
    if objective == 'klvi':
        H = lam[d:].sum() + (0.5 * d * (1 + np.log(2 * np.pi)) if kind == 'mf_gauss' else 0.0)
        return -(H + tot[2] / n), np.concatenate([-gm / n, -(1 + gl / n)])
    return np.log(tot[1] / n) / alpha + M, np.concatenate([alpha * gm / n, alpha * (gl + tot[1]) / n])


def _worker(rank, world, port, ret):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from viabel_b200 import vb
        from viabel_b200._device import all_gather_rows
        r, w, g = vb._dist_info(True)
        assert (r, w) == (rank, world)
        # sharding is opt-in: an initialised process group alone never splits the sample axis
        assert vb._dist_info(False)[:2] == (0, 1) and vb._dist_info(None)[:2] == (0, 1)
        n = 1001
        start, count = vb._shard(n, rank, world)
        d = 10
        fam = vb_np.mean_field_t_variational_family(d, 40)
        model = mm.EightSchoolsNCP()
        lam = np.concatenate([0.3 * np.random.RandomState(0).randn(d), np.full(d, -0.1)])
        full = fam.impl.draw_base(np.random.RandomState(1), n)          # every rank can regenerate it
        mine = {'eps': full['eps'][start:start + count]}
        results = {}
        for objective, alpha in (('klvi', 1.0), ('chivi', 2.0), ('klvi_pd', 1.0)):
            part = torch.from_numpy(mf_partials(fam.impl, model, objective, alpha, lam, mine))
            gathered = all_gather_rows(part, world)        # the product's own exchange helper
            results[objective] = mf_finalize('mf_t', objective, alpha, lam, gathered.numpy())
        # divergence_bound partials: [max, sum w, sum w^2, count, mean, M2] merged like bounds.cu
        lw = np.random.RandomState(2).randn(n) * 1.5 - 3.0
        x = lw[start:start + count]
        M = x.max()
        wv = np.exp(2.0 * (x - M))
        part = torch.tensor([M, wv.sum(), (wv ** 2).sum(), float(count), x.mean(), ((x - x.mean()) ** 2).sum()])
        gathered = all_gather_rows(part, world)
        results['div'] = gathered.numpy().copy()
        if rank == 0:
            ret['klvi'], ret['chivi'], ret['div'] = results['klvi'], results['chivi'], results['div']
            ret['klvi_pd'] = results['klvi_pd']
            ret['lam'], ret['lw'] = lam, lw
    finally:
        dist.destroy_process_group()


def test_shard_arithmetic():
    from viabel_b200 import vb
    for n, world in [(10, 3), (1001, 2), (8, 8), (5, 8), (2 ** 22, 8)]:
        spans = [vb._shard(n, r, world) for r in range(world)]
        assert spans[0][0] == 0 and sum(c for _, c in spans) == n
        for (s0, c0), (s1, _) in zip(spans, spans[1:]):
            assert s0 + c0 == s1                                     # contiguous, disjoint, ordered
        assert max(c for _, c in spans) - min(c for _, c in spans) <= 1


def test_two_rank_gloo_merge_equals_unsharded():
    world, port = 2, _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    lam = ret['lam']
    d = 10
    fam = vb_np.mean_field_t_variational_family(d, 40)
    model = mm.EightSchoolsNCP()
    base = fam.impl.draw_base(np.random.RandomState(1), 1001)
    v, g = vb_np.klvi_value_and_grad(fam.impl, model, lam, base)
    np.testing.assert_allclose(ret['klvi'][0], v, rtol=1e-13)
    np.testing.assert_allclose(ret['klvi'][1], g, rtol=1e-12, atol=1e-14)
    v, g = vb_np.chivi_value_and_grad(fam.impl, model, 2.0, lam, base)
    np.testing.assert_allclose(ret['chivi'][0], v, rtol=1e-13)
    np.testing.assert_allclose(ret['chivi'][1], g, rtol=1e-12, atol=1e-14)
    v, g = vb_np.klvi_pd_value_and_grad(fam.impl, model, lam, base)
    np.testing.assert_allclose(ret['klvi_pd'][0], v, rtol=1e-13)
    np.testing.assert_allclose(ret['klvi_pd'][1], g, rtol=1e-12, atol=1e-14)
    # Chan merge of (count, mean, M2) + max-rescaled weight sums reproduce the unsharded statistics
    parts, lw = ret['div'], ret['lw']
    M = parts[:, 0].max()
    sw = sum(p[1] * np.exp(2.0 * (p[0] - M)) for p in parts)
    cnt, mean, M2 = 0.0, 0.0, 0.0
    for p in parts:
        tot = cnt + p[3]
        delta = p[4] - mean
        mean += delta * p[3] / tot
        M2 += p[5] + delta * delta * cnt * p[3] / tot
        cnt = tot
    np.testing.assert_allclose(M, lw.max())
    np.testing.assert_allclose(sw / cnt, np.mean(np.exp(2.0 * (lw - lw.max()))), rtol=1e-13)
    np.testing.assert_allclose(mean, lw.mean(), rtol=1e-13)
    np.testing.assert_allclose(np.sqrt(M2 / cnt), lw.std(), rtol=1e-12)


# ------------------------------------------------------------------------------------------ sharded PSIS protocol
def _psis_worker(rank, world, port, n, ret):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from viabel_b200 import psis_dist, vb
        from psis_stages_np import NumpyStages
        u = np.random.RandomState(9).rand(n)
        lw = np.log(np.expm1(-0.5 * np.log1p(-u)) / 0.5)
        start, count = vb._shard(n, rank, world)
        mine = torch.from_numpy(lw[start:start + count].copy())
        out, k, det = psis_dist.psislw_sharded(mine, stages=NumpyStages(), want_detail=True)
        # same group, different shard sizes: the cached sizes must be detected as stale and gathered again
        cut = n // 3
        mine2 = torch.from_numpy((lw[:cut] if rank == 0 else lw[cut:]).copy())
        out2, k2, det2 = psis_dist.psislw_sharded(mine2, stages=NumpyStages(), want_detail=True)
        ret[rank] = (out.numpy(), k, det['tail_idx_sorted'], det['n2'], out2.numpy(), k2, det2['tail_idx_sorted'])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('n', [5000, 40000, 300000])      # 40000: the global (all-reduced histogram) refinement
def test_two_rank_gloo_sharded_psis_equals_unsharded(n):
    """The exchange protocol of viabel_b200/psis_dist.py over real gloo collectives (stages restated in
    numpy): both sides of the direct threshold; tail order exact, weights to summation order."""
    import warnings
    from oracle import psis_np
    sys_path_tests = os.path.dirname(os.path.abspath(__file__))
    os.environ['PYTHONPATH'] = sys_path_tests + os.pathsep + os.environ.get('PYTHONPATH', '')
    world, port = 2, _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_psis_worker, args=(world, port, n, ret), nprocs=world, join=True)
    u = np.random.RandomState(9).rand(n)
    lw = np.log(np.expm1(-0.5 * np.log1p(-u)) / 0.5)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        oout, ok, odet = psis_np.psislw_detail(lw, sort_kind='stable')
    out = np.concatenate([ret[0][0], ret[1][0]])
    out2 = np.concatenate([ret[0][4], ret[1][4]])
    np.testing.assert_allclose(out2, oout, rtol=1e-10, atol=1e-10)
    for r in range(world):
        np.testing.assert_array_equal(ret[r][6], odet['tailinds'][odet['tail_order']])
        np.testing.assert_allclose(ret[r][5], ok, rtol=1e-10)
        assert ret[r][3] == odet['n2']
        np.testing.assert_array_equal(ret[r][2], odet['tailinds'][odet['tail_order']])
        np.testing.assert_allclose(ret[r][1], ok, rtol=1e-10)
    np.testing.assert_allclose(out, oout, rtol=1e-10, atol=1e-10)


# ----------------------------------------------------------------------------------------- sharded full-rank CHIVI
def _fr_worker(rank, world, port, ret):
    """Each rank: local partials of the full-rank CHIVI step in numpy (what vb_fr_rows / vb_fr_weighted return on
    the GPU), then the product's own exchange code (viabel_b200.vb.fr_exchange_max / fr_exchange_sums) over gloo."""
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from viabel_b200 import vb
        d, n, alpha = 6, 1003, 2.0
        rs = np.random.RandomState(3)
        A = rs.randn(d, d)
        model = mm.CorrelatedGaussian(rs.randn(d), np.linalg.inv(A @ A.T / d + np.eye(d)))
        fam = vb_np.t_variational_family(d, 9)
        impl = fam.impl
        lam = 0.2 * rs.randn(impl.n_param)
        full = impl.draw_base(np.random.RandomState(4), n)
        start, count = vb._shard(n, rank, world)
        mine = {k: v[start:start + count] for k, v in full.items()}
        x = impl.transform(lam, mine)
        lp, g = model.logp_and_grad(x)
        lw = lp - impl.logq_of_base(lam, mine)
        M = lw.max()
        w = np.exp(alpha * (lw - M))                       # relative to the LOCAL max, like vb_fr_rows
        rows = torch.tensor([M, w.sum(), lp.sum(), lw.sum()])
        w_t, rows = vb.fr_exchange_max(rows, torch.from_numpy(w), alpha, None)
        wn = w_t.numpy()
        zt = impl.ztilde(mine)
        g_mu, G_S = wn @ g, zt.T @ (wn[:, None] * g)      # vb_fr_weighted with the rescaled weights
        rows, g_mu_t, G_S_t = vb.fr_exchange_sums(rows, torch.from_numpy(g_mu), torch.from_numpy(G_S), world, None, count)
        assert rows.numel() == 5 and float(rows[4]) == n        # the global sample count travels with the sums
        if rank == 0:
            ret['rows'], ret['g_mu'], ret['G_S'] = rows.numpy(), g_mu_t.numpy(), G_S_t.numpy()
            ret['lam'], ret['n'] = lam, n
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_full_rank_chivi_exchange_equals_unsharded():
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_fr_worker, args=(2, _free_port(), ret), nprocs=2, join=True)
    d, n, alpha = 6, ret['n'], 2.0
    rs = np.random.RandomState(3)
    A = rs.randn(d, d)
    model = mm.CorrelatedGaussian(rs.randn(d), np.linalg.inv(A @ A.T / d + np.eye(d)))
    impl = vb_np.t_variational_family(d, 9).impl
    lam = ret['lam']
    full = impl.draw_base(np.random.RandomState(4), n)
    ov, og = vb_np.chivi_value_and_grad(impl, model, alpha, lam, full)
    rows, g_mu, G_S = ret['rows'], ret['g_mu'], ret['G_S']
    # finish exactly as _Objective._fr does after the exchange (viabel_b200/vb.py)
    value = np.log(rows[1] / n) / alpha + rows[0]
    mu, L, Sigma = impl.fold(lam)
    S, lamb, V = impl.sqrt_sigma(Sigma)
    r = np.sqrt(lamb)
    Y = V @ ((V.T @ G_S @ V) / (r[:, None] + r[None, :])) @ V.T
    G_L = (Y + Y.T) @ L
    phi = G_L[impl.tri].copy()
    phi[impl.diag_pos] *= np.diag(L)
    phi[impl.diag_pos] += rows[1]
    grad = np.concatenate([alpha * g_mu / n, alpha * phi / n])
    np.testing.assert_allclose(value, ov, rtol=1e-12)
    np.testing.assert_allclose(grad, og, rtol=1e-10, atol=1e-13)
