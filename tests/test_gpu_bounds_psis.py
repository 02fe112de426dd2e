"""GPU parity: viabel_b200.bounds / viabel_b200.psis against the committed reference outputs
(tests/golden, produced by the reference itself) and against the oracle on seeded inputs.
Tolerances: bounds rtol 1e-10; PSIS tail index set and sort order bit-exact, cutoff bit-exact,
k-hat rtol 1e-9, smoothed log weights atol 1e-9 (summation order is the only difference)."""
import warnings

import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

import golden_inputs as gi                      # noqa: E402
from oracle import bounds_np, psis_np           # noqa: E402


def _quiet(fn, *a, **k):
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        return fn(*a, **k)


# --------------------------------------------------------------------------------------- bounds
@pytest.mark.parametrize('name', gi.BOUNDS_CASES)
def test_all_bounds_vs_reference_outputs(golden, name):
    import viabel_b200
    lw, samples, kw = gi.bounds_case(name)
    res = _quiet(viabel_b200.all_bounds, lw, samples, **kw)
    assert list(res) == ['W1', 'W2', 'mean_error', 'std_error', 'cov_error', 'd2', 'log_norm_bound']
    for k, v in res.items():
        np.testing.assert_allclose(v, golden['bounds/%s/%s' % (name, k)], rtol=1e-10, err_msg=k)
    for alpha in (1.5, 2.0, 3.0):
        np.testing.assert_allclose(_quiet(viabel_b200.divergence_bound, lw, alpha),
                                   golden['bounds/%s/d_alpha_%g' % (name, alpha)], rtol=1e-10)


def test_reference_known_answer_tests_on_gpu():
    """tests/test_bounds.py:22-70 of the reference, run against the CUDA backend."""
    from scipy.stats import norm
    import viabel_b200
    MC, TOL = 10000000, 5 / np.sqrt(10000000)

    def a_div(alpha, v1, v2):
        tmp = alpha * v2 - (alpha - 1) * v1
        return -0.5 / (alpha - 1) * np.log(tmp) + .5 * alpha / (alpha - 1) * np.log(v2) - .5 * np.log(v1)

    def kl(v1, v2):
        return .5 * (v1 / v2 + np.log(v2 / v1) - 1)
    np.random.seed(846)
    v1, v2 = 4, 16
    s = norm(scale=np.sqrt(v2)).rvs(MC)
    lw = norm(scale=np.sqrt(v1)).logpdf(s) - norm(scale=np.sqrt(v2)).logpdf(s)
    for alpha in [1.5, 2, 3]:
        for elbo in [None, 0]:
            exp = a_div(alpha, v1, v2) + (alpha / (alpha - 1) * kl(v2, v1) if elbo is None else 0)
            np.testing.assert_allclose(_quiet(viabel_b200.divergence_bound, lw, alpha, elbo), exp, atol=TOL, rtol=TOL)
    np.random.seed(341)
    d2, sd = 5.0, 3.5
    s = norm.rvs(scale=sd, size=MC)
    res = viabel_b200.wasserstein_bounds(d2, s)
    np.testing.assert_allclose(res['W1'], 2 * sd * np.sqrt(np.expm1(d2)), rtol=TOL)
    np.testing.assert_allclose(res['W2'], 2 * sd * (3 * np.expm1(d2)) ** 0.25, rtol=TOL)
    np.random.seed(1639)
    v1, v2 = 2.5, 9.3
    s = norm(scale=np.sqrt(v2)).rvs(MC)
    lw = norm(scale=np.sqrt(v1)).logpdf(s) - norm(scale=np.sqrt(v2)).logpdf(s)
    res = _quiet(viabel_b200.all_bounds, lw, s, q_var=v2, log_norm_bound=None)
    np.testing.assert_allclose(res['d2'], a_div(2, v1, v2) + 2 * kl(v2, v1), rtol=TOL)
    np.testing.assert_allclose(res['W1'], 2 * np.sqrt(v2) * np.sqrt(np.expm1(res['d2'])), rtol=TOL)
    np.testing.assert_allclose(res['W2'], 2 * np.sqrt(v2) * (3 * np.expm1(res['d2'])) ** 0.25, rtol=TOL)


@pytest.mark.parametrize('n', [2, 3, 1000, 1001, 3000001])
def test_divergence_and_moments_vs_oracle(n):
    import torch
    import viabel_b200
    rs = np.random.RandomState(n)
    lw = rs.randn(n) * 2.0 - 500.0
    for alpha in (2.0, 4.0):
        np.testing.assert_allclose(_quiet(viabel_b200.divergence_bound, lw, alpha),
                                   _quiet(bounds_np.divergence_bound, lw, alpha), rtol=1e-10)
    d = 7
    x = rs.randn(n, d) * np.arange(1, d + 1) + 100.0
    a = _quiet(viabel_b200.all_bounds, lw, x)
    b = _quiet(bounds_np.all_bounds, lw, x)
    for k in b:
        np.testing.assert_allclose(a[k], b[k], rtol=1e-9, err_msg=k)
    # CUDA tensors are consumed in place (offset view: misaligned for 16-byte loads)
    buf = torch.empty(n + 1, dtype=torch.float64, device='cuda')
    buf[1:].copy_(torch.from_numpy(lw))
    np.testing.assert_allclose(_quiet(viabel_b200.divergence_bound, buf[1:]),
                               _quiet(bounds_np.divergence_bound, lw), rtol=1e-10)


def test_mc_error_warning_text_matches_reference():
    import viabel_b200
    lw = np.array([0.0, 10.0, -10.0, 5.0])
    with warnings.catch_warnings(record=True) as got:
        warnings.simplefilter('always')
        viabel_b200.divergence_bound(lw)
    with warnings.catch_warnings(record=True) as want:
        warnings.simplefilter('always')
        bounds_np.divergence_bound(lw)
    assert len(got) == len(want) == 2
    for g, w in zip(got, want):
        gm, wm = str(g.message), str(w.message)
        assert gm.split(' (mean')[0] == wm.split(' (mean')[0]          # same sentence, same quantity name
        import re
        num = r'= ([-+0-9.eE]+|inf|nan)'
        gv = [float(t) for t in re.findall(num, gm)]
        wv = [float(t) for t in re.findall(num, wm)]
        assert len(gv) == len(wv) == 2
        np.testing.assert_allclose(gv, wv, rtol=1e-10)


# --------------------------------------------------------------------------------------- PSIS
@pytest.mark.parametrize('name', gi.PSIS_CASES)
def test_psislw_vs_reference_outputs(golden, name):
    from viabel_b200 import psis
    lw = gi.psis_case(name)
    out, k, det = psis.psislw_detail(lw)
    pre = 'psis/%s/' % name
    gk = golden[pre + 'k']
    if np.isinf(gk):
        assert np.isinf(k)
    else:
        np.testing.assert_allclose(k, gk, rtol=1e-11)
    assert det['xcutoff'] == golden[pre + 'xcutoff']                                  # bit-exact order statistic
    if det['n2'] > 4:
        assert np.array_equal(det['tail_idx_sorted'], golden[pre + 'tail_idx_sorted'])   # bit-exact set + order
        np.testing.assert_allclose(out[det['tail_idx_sorted']], golden[pre + 'tail_values'], rtol=0, atol=1e-11)
    np.testing.assert_allclose(out[gi.probe_indices(len(lw))], golden[pre + 'probe_values'], rtol=0, atol=1e-11)
    np.testing.assert_allclose(psis.sumlogs(lw), golden[pre + 'sumlogs'], rtol=1e-12)
    out2, k2 = psis.psislw(lw)
    assert np.array_equal(out, out2) and (k == k2 or np.isinf(k))
    np.testing.assert_allclose(np.sum(np.exp(out)), 1.0, rtol=1e-10)                  # normalised weights


@pytest.mark.parametrize('n,kk,reff', [(2, 0.5, 1.0), (5, 0.5, 1.0), (26, 0.5, 1.0), (4097, 0.6, 1.0),
                                        (131072, 0.5, 1.0), (131073, 0.5, 1.0), (300000, 0.4, 0.7),
                                        (2000003, 0.5, 1.0)])
def test_psislw_vs_oracle(n, kk, reff):
    """Includes both sides of the direct-sort threshold (2^17) and the histogram-refined path."""
    from viabel_b200 import psis
    lw = gi.gpd_log_weights(n, kk, 4000 + n, offset=-77.0)
    out, k, det = psis.psislw_detail(lw, Reff=reff)
    oout, ok, odet = _quiet(psis_np.psislw_detail, lw, Reff=reff, sort_kind='stable')
    if np.isinf(ok):
        assert np.isinf(k)
    else:
        np.testing.assert_allclose(k, ok, rtol=1e-11)
    assert det['xcutoff'] == odet['xcutoff'] and det['n2'] == odet['n2']
    if odet['n2'] > 4:
        assert np.array_equal(det['tail_idx_sorted'], odet['tailinds'][odet['tail_order']])
    assert det['smoothed'] == odet['smoothed']
    np.testing.assert_allclose(out, oout, rtol=0, atol=1e-11)


def test_psislw_ties_use_ascending_index():
    """Heavy ties in the tail: the contract is numpy's kind='stable' order."""
    from viabel_b200 import psis
    rs = np.random.RandomState(3)
    lw = np.round(gi.gpd_log_weights(50000, 0.5, 77), 1)           # many exact duplicates
    lw[rs.randint(0, 50000, 200)] = lw.max()                        # ties at the very top too
    out, k, det = psis.psislw_detail(lw)
    oout, ok, odet = _quiet(psis_np.psislw_detail, lw, sort_kind='stable')
    assert det['xcutoff'] == odet['xcutoff'] and det['n2'] == odet['n2']
    assert np.array_equal(det['tail_idx_sorted'], odet['tailinds'][odet['tail_order']])
    np.testing.assert_allclose(k, ok, rtol=1e-11)
    np.testing.assert_allclose(out, oout, rtol=0, atol=1e-11)


@pytest.mark.parametrize('decimals,forced', [(3, False), (1, False), (3, True)])
def test_psislw_ties_in_the_counting_sort_path(decimals, forced, monkeypatch):
    """n above the direct-sort threshold: the candidates are ordered by the counting sort over the refinement bins
    (ties share a bin and are ranked by index); coarse rounding makes bins larger than the per-warp ranking takes and
    must fall back to the bitonic network; VB_PSIS_BITONIC=1 forces the fallback.  All three agree with the
    stable-sort oracle exactly in tail membership and order."""
    import subprocess
    import sys
    lw = np.round(gi.gpd_log_weights(400000, 0.5, 78), decimals)
    oout, ok, odet = _quiet(psis_np.psislw_detail, lw, sort_kind='stable')
    if forced:                                   # the switch is read once per process: run in a child
        code = ("import numpy as np, sys; sys.path.insert(0, %r); sys.path.insert(0, %r); import golden_inputs as gi;"
                "from viabel_b200 import psis; lw = np.round(gi.gpd_log_weights(400000, 0.5, 78), %d);"
                "out, k, det = psis.psislw_detail(lw); np.save(sys.argv[1], det['tail_idx_sorted']); print(repr(k))"
                % (ROOT, os.path.join(ROOT, 'tests'), decimals))
        import tempfile
        with tempfile.TemporaryDirectory() as td:
            path = os.path.join(td, 't.npy')
            env = dict(os.environ, VB_PSIS_BITONIC='1')
            res = subprocess.run([sys.executable, '-c', code, path], env=env, capture_output=True, text=True, check=True)
            tail = np.load(path)
        k = float(res.stdout.strip().splitlines()[-1].replace('np.float64(', '').rstrip(')'))
    else:
        from viabel_b200 import psis
        out, k, det = psis.psislw_detail(lw)
        tail = det['tail_idx_sorted']
        assert det['xcutoff'] == odet['xcutoff'] and det['n2'] == odet['n2']
        np.testing.assert_allclose(out, oout, rtol=0, atol=1e-11)
    assert np.array_equal(tail, odet['tailinds'][odet['tail_order']])
    np.testing.assert_allclose(k, ok, rtol=1e-11)


def test_psislw_two_dimensional_and_device_input():
    import torch
    from viabel_b200 import psis
    a, b = gi.psis_case('gpd05_n5000'), gi.psis_case('gpd07_n20000')[:5000]
    lw2 = np.stack([a, b], axis=1)
    out, ks = psis.psislw(lw2)
    oout, oks = _quiet(psis_np.psislw, lw2)
    assert out.shape == (5000, 2) and out.flags.f_contiguous and ks.shape == (2,)
    np.testing.assert_allclose(ks, oks, rtol=1e-11)
    np.testing.assert_allclose(out, oout, rtol=0, atol=1e-11)
    dev = torch.from_numpy(a).cuda()
    o, k = psis.psislw(dev)
    assert o.is_cuda and o.data_ptr() != dev.data_ptr()
    np.testing.assert_allclose(o.cpu().numpy(), oout[:, 0], rtol=0, atol=1e-11)
    o2, _ = psis.psislw(dev, overwrite_lw=True)
    assert o2.data_ptr() == dev.data_ptr()
    np.testing.assert_allclose(dev.cpu().numpy(), oout[:, 0], rtol=0, atol=1e-11)
    # out=: the result lands in a caller-owned (here pinned) host buffer
    pin_in = torch.empty(5000, dtype=torch.float64, pin_memory=True)
    pin_in.copy_(torch.from_numpy(a))
    pin_out = torch.empty(5000, dtype=torch.float64, pin_memory=True)
    res, k3 = psis.psislw(pin_in.numpy(), out=pin_out.numpy())
    assert res is not None and np.shares_memory(res, pin_out.numpy()) and k3 == k
    np.testing.assert_array_equal(res, o.cpu().numpy())
    with pytest.raises(ValueError):
        psis.psislw(a, out=np.empty(4999))


@pytest.mark.parametrize('name', gi.LOO_CASES)
def test_psisloo_vs_reference_outputs(golden, name):
    """Batched-column kernel (one CTA per column) against the real reference's psisloo / 2-D psislw outputs.
    Tolerances: k-hat and weights 1e-11 (summation order), loo terms rtol 1e-12."""
    import torch
    from viabel_b200 import psis
    ll = gi.loo_case(name)
    pre = 'loo/%s/' % name
    loo, loos, ks = psis.psisloo(ll)
    finite = np.isfinite(golden[pre + 'ks'])
    assert np.array_equal(np.isinf(ks), ~finite)
    np.testing.assert_allclose(ks[finite], golden[pre + 'ks'][finite], rtol=1e-11)
    np.testing.assert_allclose(loos, golden[pre + 'loos'], rtol=1e-12)
    np.testing.assert_allclose(loo, golden[pre + 'loo'], rtol=1e-12)
    out, ks2 = psis.psislw(-ll)
    assert out.shape == ll.shape and out.flags.f_contiguous
    np.testing.assert_array_equal(ks2, ks)
    probe = out[gi.probe_indices(ll.shape[0])]
    cols = [j for j in range(ll.shape[1]) if not (name == 'mixed_n300_m5' and j == 3)]   # column 3: ties, see below
    np.testing.assert_allclose(probe[:, cols], golden[pre + 'lw_probe'][:, cols], rtol=0, atol=1e-11)
    # tie-break contract (numpy kind='stable'): the whole matrix against the stable-sort oracle
    oout, oks = _quiet(psis_np.psislw, -ll, sort_kind='stable')
    np.testing.assert_allclose(out, oout, rtol=0, atol=1e-11)
    # device input stays on the device
    dloo, dloos, dks = psis.psisloo(torch.from_numpy(ll).cuda())
    assert dloos.is_cuda and dks.is_cuda
    np.testing.assert_array_equal(dloos.cpu().numpy(), loos)


@pytest.mark.parametrize('n,m', [(2, 3), (5, 2), (257, 33), (4096, 5), (8192, 3), (8193, 2)])
def test_psislw_columns_match_single_column_path(n, m):
    """Every column of the batched kernel equals the single-column pipeline (tail set and order are exact in both,
    so the values agree to summation order); n = 8193 exceeds the shared-memory kernel and takes the column loop."""
    from viabel_b200 import psis
    lw = gi.gpd_log_weights(n * m, 0.6, 500 + n).reshape(n, m)
    lw[:, 0] = np.random.RandomState(n).randn(n) * 0.3              # a light-tailed column
    lw[:, 1] = np.round(lw[:, 1], 1)                                # heavy ties: the selection mode must hand this
    #                                                                 column to the full in-CTA sort (n > 2048)
    out, ks = psis.psislw(lw)
    for j in range(m):
        o1, k1 = psis.psislw(np.ascontiguousarray(lw[:, j]))
        assert (np.isinf(k1) and np.isinf(ks[j])) or abs(k1 - ks[j]) <= 1e-11 * max(1.0, abs(k1))
        np.testing.assert_allclose(out[:, j], o1, rtol=0, atol=1e-11)
    oout, oks = _quiet(psis_np.psislw, lw, sort_kind='stable')
    np.testing.assert_allclose(out, oout, rtol=0, atol=1e-11)
    loo, loos, ks3 = psis.psisloo(-lw)
    oloo, oloos, _ = _quiet(psis_np.psisloo, -lw, sort_kind='stable')
    np.testing.assert_allclose(loos, oloos, rtol=1e-11, atol=1e-11)
    with pytest.raises(ValueError):
        psis.psisloo(lw[:, 0])
    with pytest.raises(ValueError):
        psis.psisloo(lw[:1])


@pytest.mark.parametrize('name', gi.GPD_CASES)
def test_gpdfit_gpinv_sumlogs(golden, name):
    from viabel_b200 import psis
    xs = gi.gpd_case(name)
    k, sigma = psis.gpdfitnew(xs, sort=False)
    np.testing.assert_allclose(k, golden['gpd/%s/k' % name], rtol=1e-11)
    np.testing.assert_allclose(sigma, golden['gpd/%s/sigma' % name], rtol=1e-11)
    k2, s2 = psis.gpdfitnew(xs[::-1].copy())                     # unsorted input, sort=True
    np.testing.assert_allclose([k2, s2], [k, sigma], rtol=1e-12)
    p = gi.gpinv_probs()
    for (kk, s) in gi.GPINV_PARAMS:
        np.testing.assert_allclose(psis.gpinv(p, kk, s), golden['gpinv/%g_%g' % (kk, s)], rtol=1e-12,
                                   equal_nan=True)
    x = np.random.RandomState(0).randn(100001) * 30
    np.testing.assert_allclose(psis.sumlogs(x), psis_np.sumlogs(x), rtol=1e-13)


def test_all_bounds_after_psislw_reuses_the_final_pass_and_matches_the_standalone_kernel():
    """psislw's normalise pass also takes the alpha = 2 divergence statistics of the raw weights
    (vb_psis_smooth_div); all_bounds on the same unmodified CUDA tensor must then give what the standalone one-pass
    kernel gives, and any in-place change of the tensor (or another tensor) must bypass the cached statistics."""
    import warnings
    import torch
    import viabel_b200
    from viabel_b200 import psis
    from viabel_b200.psis import _DivCache
    rs = np.random.RandomState(5)
    u = rs.rand(300_001)
    lw = torch.from_numpy(np.log(np.expm1(-0.4 * np.log1p(-u)) / 0.4) - 3.0).cuda()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        ref = viabel_b200.all_bounds(lw.clone(), moment_bound_fn=lambda p: 3.0 if p == 2 else 27.0, q_var=1.0)
        assert _DivCache.lookup(lw, 1) is None
        psis.psislw(lw)
        assert _DivCache.lookup(lw, 1) is not None
        got = viabel_b200.all_bounds(lw, moment_bound_fn=lambda p: 3.0 if p == 2 else 27.0, q_var=1.0)
        for k in ref:
            np.testing.assert_allclose(got[k], ref[k], rtol=1e-10, err_msg=k)
        oracle = bounds_np.all_bounds(lw.cpu().numpy(), moment_bound_fn=lambda p: 3.0 if p == 2 else 27.0, q_var=1.0)
        for k in ref:
            np.testing.assert_allclose(got[k], oracle[k], rtol=1e-9, err_msg=k)
        # alpha != 2 never uses it; a modified tensor never uses it
        d15 = viabel_b200.divergence_bound(lw, alpha=1.5)
        np.testing.assert_allclose(d15, bounds_np.divergence_bound(lw.cpu().numpy(), alpha=1.5), rtol=1e-10)
        lw.add_(0.25)
        assert _DivCache.lookup(lw, 1) is None
        d2 = viabel_b200.divergence_bound(lw, alpha=2.0)
        np.testing.assert_allclose(d2, bounds_np.divergence_bound(lw.cpu().numpy(), alpha=2.0), rtol=1e-10)
