"""GPU parity of the sharded PSIS protocol (viabel_b200/psis_dist.py): R virtual ranks run the real
staged C ABI on one GPU, exchanging through an in-process communicator (threads + barriers) with the
same three collectives NCCL provides.  The result must equal the unsharded oracle: tail index set and
order exact, k-hat / smoothed weights to summation-order tolerance."""
import os
import sys
import threading

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from oracle import psis_np           # noqa: E402
from test_gpu_bounds_psis import _quiet  # noqa: E402

pytestmark = pytest.mark.gpu


class _Shared:
    def __init__(self, world):
        self.world = world
        self.slots = [None] * world
        self.barrier = threading.Barrier(world, timeout=120)


class ThreadComm:
    """all_gather / all_reduce between threads of one process (each thread = one virtual rank)."""

    def __init__(self, shared, rank):
        self.s, self.rank, self.world = shared, rank, shared.world

    def all_gather(self, t):
        self.s.slots[self.rank] = t.clone()
        self.s.barrier.wait()
        out = torch.stack([self.s.slots[r] for r in range(self.world)])
        self.s.barrier.wait()
        return out

    def all_reduce_sum(self, t):
        t.copy_(self.all_gather(t).sum(0))
        return t

    def all_reduce_max(self, t):
        t.copy_(self.all_gather(t).max(0).values)
        return t


def run_virtual(lw, world, reff=1.0, sizes=None):
    from viabel_b200 import psis_dist
    n = len(lw)
    if sizes is None:
        base, rem = divmod(n, world)
        sizes = [base + (1 if r < rem else 0) for r in range(world)]
    starts = np.concatenate([[0], np.cumsum(sizes)])
    shared = _Shared(world)
    res, err = [None] * world, []

    def work(r):
        try:
            torch.cuda.set_device(0)
            mine = torch.from_numpy(lw[starts[r]:starts[r + 1]].copy()).cuda()
            res[r] = psis_dist.psislw_sharded(mine, Reff=reff, comm=ThreadComm(shared, r), want_detail=True)
        except Exception as e:      # noqa: BLE001
            err.append(e)
            shared.barrier.abort()

    ts = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    if err:
        raise err[0]
    out = np.concatenate([r[0].cpu().numpy() for r in res])
    return out, res[0][1], [r[2] for r in res]


def gpd_lw(n, k, seed):
    u = np.random.RandomState(seed).rand(n)
    return np.log(np.expm1(-k * np.log1p(-u)) / k)


@pytest.mark.parametrize('n,world,k', [(5000, 2, 0.5), (5000, 3, 0.7), (300000, 2, 0.5), (300001, 4, 0.7),
                                       (2000000, 8, 0.5), (1500000, 3, 0.2)])
def test_sharded_equals_unsharded_oracle(n, world, k):
    lw = gpd_lw(n, k, n % 97)
    out, kh, dets = run_virtual(lw, world)
    oout, ok, odet = _quiet(psis_np.psislw_detail, lw, sort_kind='stable')
    for det in dets:                                    # replicated tail work: identical on every rank
        assert det['n2'] == odet['n2']
        np.testing.assert_array_equal(det['tail_idx_sorted'], odet['tailinds'][odet['tail_order']])
        np.testing.assert_allclose(det['k'], ok, rtol=1e-9)
        assert det['xcutoff'] == odet['xcutoff']
    np.testing.assert_allclose(kh, ok, rtol=1e-9)
    np.testing.assert_allclose(out, oout, rtol=1e-9, atol=1e-9)


def test_sharded_uneven_shards_and_ties():
    rs = np.random.RandomState(5)
    lw = np.round(gpd_lw(400000, 0.6, 11), 2)           # heavy ties, also inside the tail
    lw[rs.randint(0, len(lw), 50)] = lw.max()           # ties at the very top
    out, kh, dets = run_virtual(lw, 3, sizes=[10, 250000, 149990])
    oout, ok, odet = _quiet(psis_np.psislw_detail, lw, sort_kind='stable')
    np.testing.assert_array_equal(dets[0]['tail_idx_sorted'], odet['tailinds'][odet['tail_order']])
    np.testing.assert_allclose(kh, ok, rtol=1e-9)
    np.testing.assert_allclose(out, oout, rtol=1e-9, atol=1e-9)


def test_sharded_matches_single_gpu_path_bitwise_on_indices():
    from viabel_b200 import psis
    lw = gpd_lw(1 << 20, 0.5, 3)
    out1, k1, det1 = psis.psislw_detail(lw)
    out, kh, dets = run_virtual(lw, 4)
    np.testing.assert_array_equal(dets[2]['tail_idx_sorted'], det1['tail_idx_sorted'])
    np.testing.assert_allclose(kh, k1, rtol=1e-12)
    np.testing.assert_allclose(out, out1, rtol=1e-11, atol=1e-11)


def test_device_driven_protocol_is_taken_and_falls_back_on_ties(monkeypatch):
    """The device-driven protocol (no host read between stages) runs for large tie-free inputs; heavy ties in one
    refinement bin make its on-device check fail and the call falls back to the host-driven protocol; VB_PSIS_HOST=1
    forces the host-driven one.  All three give the oracle's result."""
    from viabel_b200 import psis_dist
    lw = gpd_lw(2_000_000, 0.5, 5)
    oout, ok, odet = _quiet(psis_np.psislw_detail, lw, sort_kind='stable')
    out, kh, dets = run_virtual(lw, 8)
    assert psis_dist.PATH_TAKEN[0] == 'dev'
    np.testing.assert_array_equal(dets[5]['tail_idx_sorted'], odet['tailinds'][odet['tail_order']])
    np.testing.assert_allclose(out, oout, rtol=1e-9, atol=1e-9)
    monkeypatch.setenv('VB_PSIS_HOST', '1')
    out_h, kh_h, dets_h = run_virtual(lw, 8)
    assert psis_dist.PATH_TAKEN[0] == 'host'
    np.testing.assert_array_equal(dets_h[0]['tail_idx_sorted'], dets[0]['tail_idx_sorted'])
    np.testing.assert_allclose(out_h, out, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(kh_h, kh, rtol=1e-12)
    monkeypatch.delenv('VB_PSIS_HOST')
    tied = np.round(gpd_lw(400000, 0.6, 11), 2)
    out_t, kh_t, dets_t = run_virtual(tied, 3)               # ties, but no refinement bin above 512 entries
    oout_t, ok_t, odet_t = _quiet(psis_np.psislw_detail, tied, sort_kind='stable')
    np.testing.assert_array_equal(dets_t[0]['tail_idx_sorted'], odet_t['tailinds'][odet_t['tail_order']])
    tied[np.random.RandomState(1).randint(0, len(tied), 1000)] = tied.max()      # 1000 copies of the maximum: one bin overflows
    out_t, kh_t, dets_t = run_virtual(tied, 3)
    assert psis_dist.PATH_TAKEN[0] == 'dev->host'
    oout_t, ok_t, odet_t = _quiet(psis_np.psislw_detail, tied, sort_kind='stable')
    np.testing.assert_array_equal(dets_t[0]['tail_idx_sorted'], odet_t['tailinds'][odet_t['tail_order']])
    np.testing.assert_allclose(out_t, oout_t, rtol=1e-9, atol=1e-9)
