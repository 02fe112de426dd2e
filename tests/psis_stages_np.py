"""numpy restatement of the staged PSIS C ABI (include/viabel_b200.h, vb_psis_sample ... vb_psis_write)
on CPU tensors, so that the sharded protocol of viabel_b200/psis_dist.py can run over gloo without a
GPU.  Test infrastructure only.  Candidates are (n, 2) int64 tensors holding (uint64 key, index)."""
import math

import numpy as np
import torch

from oracle import psis_np

SAMPLE = 4096


def _key_of(x):
    b = np.asarray(x, dtype=np.float64).view(np.uint64)
    neg = (b >> np.uint64(63)) != 0
    return np.where(neg, ~b, b | np.uint64(1 << 63))


def _val_of(key):
    key = np.asarray(key, dtype=np.uint64)
    pos = (key >> np.uint64(63)) != 0
    return np.where(pos, key & np.uint64((1 << 63) - 1), ~key).view(np.float64)


class NumpyStages:
    def __init__(self):
        self.device = torch.device('cpu')

    def empty(self, shape, dtype):
        return torch.zeros(shape, dtype=dtype)

    def sample(self, lw, index0):
        x = lw.numpy()
        e = (np.arange(SAMPLE, dtype=np.int64) * len(x)) // SAMPLE
        out = np.stack([_key_of(x[e]).view(np.int64), index0 + e], axis=1)
        return torch.from_numpy(out)

    def pivot(self, sample_buf, count, frac, out):
        keys = np.sort(sample_buf.numpy()[:count, 0].view(np.uint64))
        rank = int(math.ceil(frac * count))
        out[0] = -math.inf if rank >= count else float(_val_of(keys[count - 1 - rank]))

    def filter(self, lw, index0, pivot, cap):
        x = lw.numpy()
        p = float(pivot[0])
        is_c = x >= p
        idx = np.nonzero(is_c)[0]
        cand = np.zeros((cap, 2), dtype=np.int64)
        k = min(len(idx), cap)
        cand[:k, 0] = x[idx[:k]].view(np.int64)
        cand[:k, 1] = index0 + idx[:k]
        with np.errstate(over='ignore', invalid='ignore'):
            body = np.exp(x[~is_c] - p).sum() if (~is_c).any() else 0.0
        cmax = x[idx].max() if len(idx) else -math.inf
        return torch.from_numpy(cand), torch.tensor([len(idx)]), torch.tensor([body, cmax], dtype=torch.float64)

    def prepare(self, cand, count, M, pivot, body_max, terms):
        c = cand.numpy()
        x = c[:count, 0].view(np.float64)
        u = float(M[0]) - x
        c[:count, 0] = (~u.view(np.uint64)).view(np.int64)
        b = float(body_max[0])
        terms[0] = 0.0 if b == 0.0 else b * math.exp(float(pivot[0]) - float(M[0]))

    def refine(self, cand, count, M, pivot, keep, terms):
        c = cand.numpy()[:count]
        if count > 2 * keep and count > 8192:               # keep a superset of the `keep` largest
            u = (~c[:, 0].view(np.uint64)).view(np.float64)
            thr = np.partition(u, keep - 1)[keep - 1]
            sel = u <= thr
            terms[1] = float(np.exp(-u[~sel]).sum())
            c = c[sel]
        else:
            terms[1] = 0.0
        out = torch.from_numpy(np.ascontiguousarray(c).copy()) if len(c) else torch.zeros((1, 2), dtype=torch.int64)
        return out, torch.tensor([len(c)])

    BINS = 8192

    def _bins(self, cand, count, M, pivot):
        c = cand.numpy()[:count]
        u = (~c[:, 0].view(np.uint64)).view(np.float64)
        pu = float(M[0]) - float(pivot[0])
        scale = self.BINS / pu if (pu > 0.0 and math.isfinite(pu)) else 0.0
        return c, u, np.minimum((u * scale).astype(np.int64), self.BINS - 1)

    def refine_hist(self, cand, count, M, pivot):
        _, _, b = self._bins(cand, count, M, pivot)
        return torch.from_numpy(np.bincount(b, minlength=self.BINS).astype(np.int32))

    def refine_select(self, cand, count, M, pivot, hist_global, keep, terms):
        c, u, b = self._bins(cand, count, M, pivot)
        cum = np.cumsum(hist_global.numpy().astype(np.int64))
        hit = np.nonzero(cum >= keep)[0]
        bstar = int(hit[0]) if len(hit) else self.BINS - 1
        sel = b <= bstar
        terms[1] = float(np.exp(-u[~sel]).sum())
        c = c[sel]
        out = torch.from_numpy(np.ascontiguousarray(c).copy()) if len(c) else torch.zeros((1, 2), dtype=torch.int64)
        return out, torch.tensor([len(c)])

    def tail(self, cands, count, K1, M, pivot, terms, want_tail_idx=False):
        c = cands.numpy()[:count]
        keys, idx = c[:, 0].view(np.uint64), c[:, 1]
        order = np.lexsort((idx, keys))                      # ascending key, ties by ascending index
        u = (~keys[order]).view(np.float64)                  # descending u == ascending shifted weight
        xs = -u
        idx = idx[order]
        xcut = max(xs[count - 1 - K1], math.log(np.finfo(float).tiny))      # psis.py:157,169-172
        p = int(np.searchsorted(xs, xcut, side='right'))     # strict tail x > xcutoff (psis.py:174)
        n2 = count - p
        tail_x, tail_i = xs[p:], idx[p:]
        k, sigma, smoothed = math.inf, math.nan, False
        new = tail_x.copy()
        if n2 > 4:                                           # psis.py:177-198
            x2 = np.exp(tail_x) - math.exp(xcut)
            k, sigma = psis_np.gpdfitnew(x2, sort=False)
            if k >= 1.0 / 3.0 and not math.isinf(k):
                new = np.log(psis_np.gpinv(np.arange(0.5, n2) / n2, k, sigma) + math.exp(xcut))
                new[new > 0] = 0
                smoothed = True
        total = float(np.sum(terms.numpy())) + float(np.exp(xs[:p]).sum()) + float(np.exp(new).sum())
        lse = math.log(total)
        diag = torch.tensor([k, sigma, xcut, n2, float(smoothed), float(M[0]), lse, count], dtype=torch.float64)
        tail_idx = torch.from_numpy(np.concatenate([tail_i, -np.ones(K1 + 1 - n2, dtype=np.int64)])) if want_tail_idx else None
        return diag, tail_idx, dict(lse=lse, idx=tail_i, new=new, smoothed=smoothed)

    def write(self, lw, index0, M, handle):
        x = lw.numpy()
        out = (x - float(M[0])) - handle['lse']
        if handle['smoothed']:
            li = handle['idx'] - index0
            ok = (li >= 0) & (li < len(x))
            out[li[ok]] = handle['new'][ok] - handle['lse']
        return torch.from_numpy(out)
