"""Parity at BASELINE.json's FULL sizes through size-independent properties (the oracle cannot run there in
seconds): normalisation, body invariance, tail sortedness, shift invariance and permutation equivariance of psislw
on 1.25e8 weights (C5, one GPU's shard); agreement of the one-pass divergence kernel with a plain fp64 torch
evaluation; additivity of the fused KLVI step over the Monte Carlo sample axis at 2^22 samples (C4: two half
launches with global sample offsets merge to the single launch -- the multi-GPU sharding identity on one GPU) and
equality of the Philox-in-kernel and the streamed-base-draw modes; determinism of the d = 512 CHIVI step (C3)."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _c5_weights(n, seed):
    import torch
    g = torch.Generator(device='cuda')
    g.manual_seed(seed)
    u = torch.rand(n, dtype=torch.float64, device='cuda', generator=g)
    return torch.log(torch.expm1(-0.5 * torch.log1p(-u)) / 0.5)          # log of GPD(k = 0.5, sigma = 1) variates


def test_c5_psislw_properties_at_full_size():
    import torch
    from viabel_b200 import psis
    n = 125_000_000
    lw = _c5_weights(n, 7)
    out, k, det = psis.psislw_detail(lw)
    K1 = int(math.ceil(min(0.2 * n, 3.0 * math.sqrt(n))))                 # psis.py:157
    assert det['n2'] == K1 and det['smoothed'] and 0.4 < k < 0.6          # continuous data: no ties at the cutoff
    # (1) normalised: log-sum-exp of the result is 0 (psis.py:200)
    assert abs(torch.logsumexp(out, 0).item()) < 1e-9
    # (2) the body is the input shifted by ONE constant, exactly as the reference computes it: (x - max) - lse
    M, lse = det['max'], det['lse']
    assert M == lw.max().item()
    changed = (out != (lw - M) - lse)
    assert int(changed.sum().item()) <= K1
    tail = torch.from_numpy(det['tail_idx_sorted']).cuda()
    assert bool(changed[tail].all()) or int(changed.sum().item()) >= K1 - 8   # (a smoothed value may coincide by chance)
    # (3) tail order: ascending raw weights (ties impossible here) and ascending smoothed values; every raw tail
    # weight is strictly above the cutoff, everything else is at or below it
    raw_tail = lw[tail] - M
    assert bool((raw_tail[1:] > raw_tail[:-1]).all())
    new_tail = out[tail]
    assert bool((new_tail[1:] >= new_tail[:-1]).all())
    assert raw_tail.min().item() > det['xcutoff']
    mask = torch.ones(n, dtype=torch.bool, device='cuda')
    mask[tail] = False
    assert (lw[mask] - M).max().item() <= det['xcutoff']
    del mask, changed
    # (4) shift invariance: psislw(lw + c) == psislw(lw) (the reference subtracts the maximum first, psis.py:165)
    out_s, k_s = psis.psislw(lw + 3.25)
    assert abs(k_s - k) <= 1e-9 * abs(k)
    assert (out_s - out).abs().max().item() < 1e-9
    del out_s
    # (5) permutation equivariance (reversal): same tail set, same values at the mirrored positions
    out_r, k_r = psis.psislw(torch.flip(lw, [0]))
    assert abs(k_r - k) <= 1e-11 * abs(k)
    assert (torch.flip(out_r, [0]) - out).abs().max().item() < 1e-11


def test_c5_divergence_matches_plain_fp64_at_full_size():
    import torch
    import viabel_b200
    n = 125_000_000
    lw = _c5_weights(n, 8) - 6.0
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        d2, elbo = viabel_b200.divergence_bound(lw, alpha=2., return_log_norm_bound=True)
    cubo = 0.5 * (torch.logsumexp(2.0 * lw, 0).item() - math.log(n))      # bounds.py:169-173 in one expression
    ref_elbo = lw.mean().item()
    np.testing.assert_allclose(elbo, ref_elbo, rtol=1e-11)
    np.testing.assert_allclose(d2, 2.0 * (cubo - ref_elbo), rtol=1e-10)


def test_c4_step_is_additive_over_the_sample_axis_at_full_size():
    import torch
    from viabel_b200 import models, vb, _lib
    n, d = 1 << 22, 10
    rs = np.random.RandomState(5)
    lam = torch.from_numpy(np.concatenate([0.3 * rs.randn(d), -0.1 + 0.05 * rs.randn(d)])).cuda()
    model = models.EightSchoolsNCP()
    op = torch.ops.viabel_b200
    for objective, alpha in ((_lib.OBJ_KLVI, 1.0), (_lib.OBJ_CHIVI, 2.0)):
        whole, p_whole = op.mf_step(_lib.FAMILY_MF_T, 40.0, model.handle.value, objective, alpha, lam, n, None, 0, 3, 0)
        _, p_a = op.mf_step(_lib.FAMILY_MF_T, 40.0, model.handle.value, objective, alpha, lam, n // 2, None, 0, 3, 0)
        _, p_b = op.mf_step(_lib.FAMILY_MF_T, 40.0, model.handle.value, objective, alpha, lam, n - n // 2, None, 0, 3,
                            n // 2)
        merged = op.mf_finalize(_lib.FAMILY_MF_T, objective, alpha, lam, torch.stack([p_a, p_b]).contiguous())
        assert torch.isfinite(whole).all()
        np.testing.assert_allclose(merged.cpu().numpy(), whole.cpu().numpy(), rtol=1e-11, atol=1e-13)
        assert p_whole[3].item() == n
    # streamed base draws == Philox inside the kernel (same seed, call offset, global sample indices)
    fam = vb.mean_field_t_variational_family(d, 40)
    eps = fam.impl.base_draws(n, 0, 3)['eps']
    streamed, _ = op.mf_step(_lib.FAMILY_MF_T, 40.0, model.handle.value, _lib.OBJ_KLVI, 1.0, lam, n, eps, 0, 0, 0)
    fused, _ = op.mf_step(_lib.FAMILY_MF_T, 40.0, model.handle.value, _lib.OBJ_KLVI, 1.0, lam, n, None, 0, 3, 0)
    np.testing.assert_allclose(streamed.cpu().numpy(), fused.cpu().numpy(), rtol=1e-11, atol=1e-13)


def test_c3_full_rank_chivi_step_is_deterministic_at_full_size():
    import torch
    import bench_workloads as W
    from viabel_b200 import models, vb
    d, n = 512, 1 << 16
    mp, Pp, lam = W.c3_problem(d)
    fam = vb.t_variational_family(d, 40, dtype='f32')
    obj = vb.black_box_chivi(2.0, fam, models.CorrelatedGaussian(mp, Pp), n, group=False)
    base = fam.impl.base_draws(n, 11, 0)
    v1, g1 = obj(lam, base_draws=base)
    v2, g2 = obj(lam, base_draws=base)
    assert np.isfinite(v1) and np.all(np.isfinite(g1)) and g1.shape == (d + d * (d + 1) // 2,)
    assert v1 == v2 and np.array_equal(g1, g2)          # fixed-order reductions: bit-identical repeats
