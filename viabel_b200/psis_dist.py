"""``psislw`` for log weights sharded across the ranks of a torch.distributed group.

Rank r holds the contiguous slice ``[index0_r, index0_r + n_r)`` of the global weight vector.  The
result equals the unsharded ``psislw`` of the concatenation: the tail index set, its order
(ties by ascending global index) and k-hat are the same on every rank; each rank gets the
smoothed, normalised weights of its own slice.

Exchange per call (all small; nothing bulk crosses NVLink):
  all_gather   shard sizes (1 int64, first call only: cached and re-validated), strided samples (4096 x 16 B per rank)
  all_gather   (candidate count, buffer capacity, shard size) per rank;  all_reduce  local maxima (max)
  all_gather   refinement histograms (8192 x int32 per rank): one global threshold, and every rank's kept count
  all_gather   the kept (key, index) pairs (~K+1 in TOTAL; K = 94 869 at n = 1e9: ~1.5 MB) with the two body-sum
               terms of each rank in the last row
The device work is the staged C ABI of ``include/viabel_b200.h`` (``vb_psis_sample`` ... ``vb_psis_write``);
``stages`` lets the CPU tests drive the same protocol over gloo with a numpy restatement of the stages.
"""
import math

import torch

from . import _lib

SAMPLE = 4096
DIRECT_MAX = 1 << 17


def psis_tail_count(n, Reff):
    """``ceil(min(0.2 n, 3 sqrt(n / Reff)))`` (``psis.py:157``)."""
    return int(math.ceil(min(0.2 * n, 3.0 * math.sqrt(n / Reff))))


def _pow2_at_least(x, lo=SAMPLE):
    p = lo
    while p < x:
        p <<= 1
    return p


class CudaStages:
    """The staged C ABI on the current CUDA device / stream."""

    def __init__(self):
        from ._device import Workspace, device
        self.ws = Workspace()
        self.device = device()

    def _p(self, t):
        return 0 if t is None else t.data_ptr()

    def _stream(self):
        return torch.cuda.current_stream().cuda_stream

    def empty(self, shape, dtype):
        return torch.empty(shape, dtype=dtype, device=self.device)

    def sample(self, lw, index0):
        out = self.empty((SAMPLE, 2), torch.int64)
        _lib.check(_lib.lib.vb_psis_sample(self._p(lw), lw.numel(), index0, self._p(out), self._stream()), 'vb_psis_sample')
        return out

    def pivot(self, sample_buf, count, frac, out):
        _lib.check(_lib.lib.vb_psis_pivot(self._p(sample_buf), count, sample_buf.shape[0], float(frac), self._p(out),
                                          self._stream()), 'vb_psis_pivot')

    def filter(self, lw, index0, pivot, cap):
        cand = self.empty((cap, 2), torch.int64)
        count = self.empty((1,), torch.int64)
        body_max = self.empty((2,), torch.float64)
        nb = _lib.lib.vb_psis_filter_workspace_bytes(lw.numel())
        ws = self.ws.get('filter', nb)
        _lib.check(_lib.lib.vb_psis_filter(self._p(lw), lw.numel(), index0, self._p(pivot), self._p(cand), cap,
                                           self._p(count), self._p(body_max), self._p(ws), ws.numel(), self._stream()),
                   'vb_psis_filter')
        return cand, count, body_max

    def prepare(self, cand, count, M, pivot, body_max, terms):
        _lib.check(_lib.lib.vb_psis_prepare(self._p(cand), count, self._p(M), self._p(pivot), self._p(body_max),
                                            self._p(terms), self._stream()), 'vb_psis_prepare')

    def refine(self, cand, count, M, pivot, keep, terms):
        out = self.empty((max(count, 1), 2), torch.int64)
        cnt = self.empty((1,), torch.int64)
        ws = self.ws.get('refine', _lib.lib.vb_psis_refine_workspace_bytes())
        _lib.check(_lib.lib.vb_psis_refine(self._p(cand), count, self._p(M), self._p(pivot), keep, self._p(out),
                                           self._p(cnt), self._p(terms) + 8, self._p(ws), ws.numel(), self._stream()),
                   'vb_psis_refine')
        return out, cnt

    def refine_hist(self, cand, count, M, pivot):
        """This rank's histogram of candidate keys (int32 [bins]); summed over ranks it selects ONE threshold."""
        hist = self.empty((int(_lib.lib.vb_psis_hist_bins()),), torch.int32)
        ws = self.ws.get('refine', _lib.lib.vb_psis_refine_workspace_bytes())
        _lib.check(_lib.lib.vb_psis_refine_hist(self._p(cand), count, self._p(M), self._p(pivot), self._p(hist),
                                                self._p(ws), ws.numel(), self._stream()), 'vb_psis_refine_hist')
        return hist

    def refine_select(self, cand, count, M, pivot, hist_global, keep, terms):
        out = self.empty((max(count, 1), 2), torch.int64)
        cnt = self.empty((1,), torch.int64)
        ws = self.ws.get('refine', _lib.lib.vb_psis_refine_workspace_bytes())
        _lib.check(_lib.lib.vb_psis_refine_select(self._p(cand), count, self._p(M), self._p(pivot),
                                                  self._p(hist_global), keep, self._p(out), self._p(cnt),
                                                  self._p(terms) + 8, self._p(ws), ws.numel(), self._stream()),
                   'vb_psis_refine_select')
        return out, cnt

    def tail(self, cands, count, K1, M, pivot, terms, want_tail_idx=False):
        diag = self.empty((8,), torch.float64)
        tail_idx = torch.full((K1 + 1,), -1, dtype=torch.int64, device=self.device) if want_tail_idx else None
        ws = self.ws.get('tail', _lib.lib.vb_psis_tail_workspace_bytes(count, K1))
        _lib.check(_lib.lib.vb_psis_tail(self._p(cands), count, K1, self._p(M), self._p(pivot), self._p(terms),
                                         terms.numel(), self._p(diag), self._p(tail_idx), self._p(ws), ws.numel(),
                                         self._stream()), 'vb_psis_tail')
        return diag, tail_idx, (ws, count, K1)

    # ---- device-driven variants: counts stay in device memory, nothing is read back between the stages
    def prepare_hist_dev(self, cand, count_t, M, pivot, body_max, terms):
        hist = self.empty((int(_lib.lib.vb_psis_hist_bins()),), torch.int32)
        ws = self.ws.get('refine', _lib.lib.vb_psis_refine_workspace_bytes())
        _lib.check(_lib.lib.vb_psis_prepare_hist_dev(self._p(cand), cand.shape[0], self._p(count_t), self._p(M), self._p(pivot),
                                                     self._p(body_max), self._p(terms), self._p(hist), self._p(ws),
                                                     ws.numel(), self._stream()), 'vb_psis_prepare_hist_dev')
        return hist

    def refine_select_dev(self, cand, count_t, M, pivot, hist_global, keep, capx, terms, n_local):
        send = self.empty((capx + 3, 2), torch.int64)
        ws = self.ws.get('refine', _lib.lib.vb_psis_refine_workspace_bytes())
        _lib.check(_lib.lib.vb_psis_refine_select_dev(self._p(cand), cand.shape[0], self._p(count_t), self._p(M),
                                                      self._p(pivot), self._p(hist_global), keep, self._p(send), capx,
                                                      self._p(terms), n_local, self._p(ws), ws.numel(), self._stream()),
                   'vb_psis_refine_select_dev')
        return send

    def gather_compact(self, recv, world, capx, sizes_t):
        cands = self.empty((capx, 2), torch.int64)
        total = self.empty((1,), torch.int64)
        terms = self.empty((2 * world,), torch.float64)
        status = self.empty((1,), torch.int32)
        _lib.check(_lib.lib.vb_psis_gather_compact(self._p(recv), world, capx, self._p(sizes_t), self._p(cands),
                                                   self._p(total), self._p(terms), self._p(status), self._stream()),
                   'vb_psis_gather_compact')
        return cands, total, terms, status

    def tail_dev(self, cands, total_t, K1, M, pivot, terms, status_t, want_tail_idx=False):
        cap = cands.shape[0]
        diag = self.empty((8,), torch.float64)
        tail_idx = torch.full((K1 + 1,), -1, dtype=torch.int64, device=self.device) if want_tail_idx else None
        ws = self.ws.get('tail', _lib.lib.vb_psis_tail_workspace_bytes(cap, K1))
        _lib.check(_lib.lib.vb_psis_tail_dev(self._p(cands), cap, self._p(total_t), K1, self._p(M), self._p(pivot),
                                             self._p(terms), terms.numel(), 0, 0, self._p(status_t), self._p(diag),
                                             self._p(tail_idx), self._p(ws), ws.numel(), self._stream()),
                   'vb_psis_tail_dev')
        return diag, tail_idx, (ws, cap, K1)

    def merge_div(self, gathered, world):
        """(world, 6) divergence partials -> the 8 statistics of divergence_bound(alpha = 2) (device)."""
        out = self.empty((8,), torch.float64)
        _lib.check(_lib.lib.vb_bounds_divergence_merge(self._p(gathered), int(world), 2.0, self._p(out), self._stream()),
                   'vb_bounds_divergence_merge')
        return out

    def write(self, lw, index0, M, handle):
        """S6; also leaves this shard's divergence partials (alpha = 2, relative to the global maximum) in
        ``self.last_div``: taken in the same pass over the weights."""
        ws, count, K1 = handle
        out = torch.empty_like(lw)
        part = self.empty((6,), torch.float64)
        dws = self.ws.get('div', _lib.lib.vb_psis_div_workspace_bytes())
        _lib.check(_lib.lib.vb_psis_write_div(self._p(lw), self._p(out), lw.numel(), index0, self._p(M), count, K1,
                                              self._p(ws), self._p(part), self._p(dws), dws.numel(), self._stream()),
                   'vb_psis_write_div')
        self.last_div = part
        return out


class DistComm:
    """The three collectives the protocol needs, over a torch.distributed group (NCCL on the GPU
    box, gloo in the CPU tests).  Tests substitute an in-process communicator with the same methods."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.cache_key = ('dist', id(group) if group is not None else 0, self.world)

    def all_gather(self, t):
        """(world, *t.shape): row r = rank r's tensor (equal shapes on every rank)."""
        flat = t.contiguous().view(-1)
        out = torch.empty(self.world * flat.numel(), dtype=t.dtype, device=t.device)
        self.dist.all_gather_into_tensor(out, flat, group=self.group)
        return out.view((self.world,) + tuple(t.shape))

    def all_reduce_max(self, t):
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return t

    def all_reduce_sum(self, t):
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return t


def psislw_sharded(lw_local, Reff=1.0, group=None, stages=None, want_detail=False, comm=None, src=None):
    """Sharded ``psislw`` (``psis.py:112-208``) of one weight column.

    ``lw_local``: this rank's contiguous slice (float64, on the stages' device).  Returns
    ``(lw_out_local, k)`` or, with ``want_detail``, ``(lw_out_local, k, detail)`` where detail holds
    the diag vector and the tail's global indices in ascending tail order."""
    st = stages if stages is not None else CudaStages()
    comm = comm if comm is not None else DistComm(group)
    world, rank = comm.world, comm.rank
    n_local = int(lw_local.numel())
    dev = lw_local.device
    # shard sizes: gathered once per communicator and re-used; every call re-validates them inside the candidate
    # count exchange below (each rank's n_local rides along), and starts over with a fresh gather if any changed
    key = getattr(comm, 'cache_key', None)
    sizes = _SIZES.get(key) if key is not None else None
    for _round in range(2):
        if sizes is None:
            sizes = [int(v) for v in comm.all_gather(torch.tensor([n_local], dtype=torch.int64, device=dev)).cpu().view(-1).tolist()]
            if key is not None:
                _SIZES[key] = sizes
        res = None
        if _round == 0 and _dev_path_ok(st, comm, sum(sizes)):
            res = _sharded_dev(st, comm, lw_local, Reff, sizes, want_detail, src)     # None: take the host-driven path
            PATH_TAKEN[0] = 'dev' if res is not None else 'dev->host'
        else:
            PATH_TAKEN[0] = 'host'
        if res is None:
            res = _sharded_with_sizes(st, comm, lw_local, Reff, sizes, want_detail, src)
        if res is not None:
            return res
        sizes = None                       # some rank's shard changed since the sizes were cached
        if key is not None:
            _SIZES.pop(key, None)
    raise _lib.ViabelB200Error('psislw: shard sizes changed during the call')


_SIZES = {}
_SIZES_DEV = {}
PATH_TAKEN = [None]          # which protocol the last call ran: 'dev', 'dev->host' (fallback) or 'host' (tests)


def _dev_path_ok(st, comm, n):
    import os
    return (hasattr(st, 'tail_dev') and hasattr(comm, 'all_reduce_sum') and n > DIRECT_MAX
            and os.environ.get('VB_PSIS_HOST', '0') in ('', '0'))


def exchange_capacity(K1):
    """Rows of the fixed-capacity candidate slot every rank sends: the global refinement keeps K1 + 1 candidates plus
    the rest of the threshold bin in TOTAL, whatever the number of ranks, and one rank may hold all of them."""
    return (K1 + 1) + (K1 + 1) // 4 + 8192


def _sharded_dev(st, comm, lw_local, Reff, sizes, want_detail, src):
    """Device-driven protocol: four collectives (samples, maximum, histogram, candidate slots), no host read until
    the diagnostics come back at the end.  Counts, capacities and the refinement threshold stay in device memory;
    the gathers have a fixed capacity (``exchange_capacity``), so no rank needs a host-side count.  Returns None when
    the on-device checks failed (too few candidates for the first pivot, an overflowing buffer, heavy ties, a changed
    shard size): the caller then runs the host-driven protocol, which handles those cases."""
    world, rank = comm.world, comm.rank
    n_local = int(lw_local.numel())
    n = sum(sizes)
    index0 = sum(sizes[:rank])
    K1 = psis_tail_count(n, Reff)
    capx = exchange_capacity(K1)
    key = (getattr(comm, 'cache_key', None), tuple(sizes), str(lw_local.device))
    sizes_t = _SIZES_DEV.get(key)
    if sizes_t is None:
        sizes_t = torch.tensor(sizes, dtype=torch.int64, device=lw_local.device)
        _SIZES_DEV.clear()
        _SIZES_DEV[key] = sizes_t
    pivot = st.empty((1,), torch.float64)
    mine = st.sample(lw_local, index0)
    cap_s = _pow2_at_least(world * SAMPLE)
    sample_all = st.empty((cap_s, 2), torch.int64)
    sample_all[:world * SAMPLE] = comm.all_gather(mine).view(world * SAMPLE, 2)
    cap = min(n_local, max(1 << 21, n_local // 16))
    # First pivot: 8x the tail fraction, at least sample rank 10 of the world * 4096 sampled weights.  Fewer than K1 + 1
    # true values above it needs >= rank sample hits where rank / 8 are expected (P < 1e-9); the status word catches it
    # and the host-driven protocol redoes the call from its 1 % pivot.  At 8 x 1.25e8 weights this stages 13x fewer
    # candidates per rank than the 1 % pivot.
    frac = min(1.0, max(8.0 * (K1 + 1) / n, 10.0 / (world * SAMPLE)))
    st.pivot(sample_all, world * SAMPLE, frac, pivot)
    cand, count_t, body_max = st.filter(lw_local, index0, pivot, cap)
    M = body_max[1:2].clone()
    comm.all_reduce_max(M)
    terms = st.empty((2,), torch.float64)
    hist = st.prepare_hist_dev(cand, count_t, M, pivot, body_max, terms)
    comm.all_reduce_sum(hist)                                            # 32 KB: one global threshold
    send = st.refine_select_dev(cand, count_t, M, pivot, hist, K1 + 1, capx, terms, n_local)
    recv = comm.all_gather(send)                                         # (world, capx + 3, 2)
    all_cands, total_t, all_terms, status_t = st.gather_compact(recv, world, capx, sizes_t)
    diag, tail_idx, handle = st.tail_dev(all_cands, total_t, K1, M, pivot, all_terms, status_t, want_tail_idx=want_detail)
    out = st.write(lw_local, index0, M, handle)
    host_stats = None
    if src is not None and getattr(st, 'last_div', None) is not None and hasattr(st, 'merge_div'):
        # the caller's own tensor: all_bounds(lw) usually follows (experiments.py:69,159-166).  Its alpha = 2 statistics
        # are merged here (one small all-gather) and come back with the diagnostics in the same host read
        stats = st.merge_div(comm.all_gather(st.last_div), world)
        both = torch.cat([diag, stats]).cpu().numpy()                    # the one host read of the call
        d, host_stats = both[:8], both[8:16].copy()
    else:
        d = diag.cpu().numpy()                                           # the one host read of the call
    if d[7] < 0:
        return None
    if src is not None and getattr(st, 'last_div', None) is not None:
        from .psis import _DivCache
        _DivCache.store(src, st.last_div, world, host=host_stats)
    if not want_detail:
        return out, d[0]
    n2 = int(d[3])
    detail = dict(k=d[0], sigma=d[1], xcutoff=d[2], n2=n2, smoothed=bool(d[4]), max=d[5], lse=d[6],
                  n_candidates=int(d[7]), tail_idx_sorted=tail_idx[:n2].cpu().numpy(), index0=index0, n=n)
    return out, d[0], detail


def _sharded_with_sizes(st, comm, lw_local, Reff, sizes, want_detail, src=None):
    world, rank = comm.world, comm.rank
    n_local = int(lw_local.numel())
    dev = lw_local.device
    n = sum(sizes)
    if n <= 1:
        raise ValueError("More than one log-weight needed.")
    if min(sizes) < 1:
        raise ValueError('every rank needs at least one log weight')
    index0 = sum(sizes[:rank])
    K1 = psis_tail_count(n, Reff)
    direct = n <= DIRECT_MAX

    pivot = st.empty((1,), torch.float64)
    if direct:
        pivot.fill_(-math.inf)
        cap = n_local
    else:
        mine = st.sample(lw_local, index0)
        cap_s = _pow2_at_least(world * SAMPLE)
        sample_all = st.empty((cap_s, 2), torch.int64)
        sample_all[:world * SAMPLE] = comm.all_gather(mine).view(world * SAMPLE, 2)
        cap = min(n_local, max(1 << 21, n_local // 16))
    frac = max(0.01, 4.0 * (K1 + 1) / n)
    for attempt in range(10):
        if not direct:
            st.pivot(sample_all.clone(), world * SAMPLE, frac, pivot)    # the stage may sort / pad its input
        cand, count_t, body_max = st.filter(lw_local, index0, pivot, cap)
        # one collective + one host read: every rank's candidate count, buffer capacity and shard size
        trip = torch.cat([count_t.view(1).to(torch.int64), torch.tensor([cap, n_local], dtype=torch.int64, device=dev)])
        gathered = comm.all_gather(trip).cpu().view(-1, 3).tolist()
        if [int(g[2]) for g in gathered] != list(sizes):
            return None
        all_counts = [int(g[0]) for g in gathered]
        count, total = all_counts[rank], sum(all_counts)
        if any(int(g[0]) > int(g[1]) for g in gathered):
            raise _lib.ViabelB200Error('psislw: more candidates than the buffer holds (heavy ties at the top)')
        if total >= K1 + 1:
            break
        if direct or frac >= 1.0:
            raise _lib.ViabelB200Error('psislw: could not collect the tail candidates')
        frac = min(1.0, frac * 4.0)
    M = body_max[1:2].clone()
    comm.all_reduce_max(M)
    terms = st.empty((2,), torch.float64)
    st.prepare(cand, count, M, pivot, body_max, terms)
    if world > 1 and total > 2 * (K1 + 1) and total > 8192:
        # global refinement: the bins are the same on every rank, so the ranks exchange their 8192 counters (one
        # all-gather, 32 KB per rank) and cut at ONE threshold: ~K+1 candidates in total, independent of the world
        # size -- and every rank also knows how many each rank keeps, so no second count exchange is needed
        hist_all = comm.all_gather(st.refine_hist(cand, count, M, pivot))                     # (world, bins) int32
        cand2, _ = st.refine_select(cand, count, M, pivot, hist_all.sum(0, dtype=torch.int32), K1 + 1, terms)
        h = hist_all.cpu().numpy().astype('int64')            # the one host read of this exchange
        cum = h.sum(0).cumsum()
        hit = (cum >= K1 + 1).nonzero()[0]
        bstar = int(hit[0]) if len(hit) else h.shape[1] - 1   # same rule as the device scan (hist_scan_kernel)
        counts = [int(v) for v in h[:, :bstar + 1].sum(1)]
    else:
        cand2, count2_t = st.refine(cand, count, M, pivot, K1 + 1, terms)
        counts = [int(v) for v in comm.all_gather(count2_t.view(1).to(torch.int64)).cpu().view(-1).tolist()]
    count2 = counts[rank]
    maxc = max(max(counts), 1)
    # candidates and the two body-sum terms travel in ONE all-gather: row maxc of the send buffer holds the terms
    send = st.empty((maxc + 1, 2), torch.int64)
    send[:count2] = cand2[:count2]
    if count2 < maxc:
        send[count2:maxc] = 0
    send[maxc] = terms.view(torch.int64)
    recv = comm.all_gather(send)                                      # (world, maxc + 1, 2)
    all_cands = torch.cat([recv[r, :counts[r]] for r in range(world)], dim=0).contiguous()
    all_terms = recv[:, maxc, :].contiguous().view(torch.float64).view(-1)     # rank order: deterministic sum
    total2 = sum(counts)
    diag, tail_idx, handle = st.tail(all_cands, total2, K1, M, pivot, all_terms, want_tail_idx=want_detail)
    out = st.write(lw_local, index0, M, handle)
    if src is not None and getattr(st, 'last_div', None) is not None:
        from .psis import _DivCache                      # the caller's tensor: all_bounds(lw, group=...) reuses the pass
        _DivCache.store(src, st.last_div, world)
    n = sum(sizes)
    d = diag.cpu().numpy()
    if not want_detail:
        return out, d[0]
    n2 = int(d[3])
    detail = dict(k=d[0], sigma=d[1], xcutoff=d[2], n2=n2, smoothed=bool(d[4]), max=d[5], lse=d[6],
                  n_candidates=int(d[7]), tail_idx_sorted=tail_idx[:n2].cpu().numpy(), index0=index0, n=n)
    return out, d[0], detail
