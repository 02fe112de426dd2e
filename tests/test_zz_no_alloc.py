"""vb_bounds_divergence_static: the no-allocation variant of vb_bounds_divergence (scratch in a file-scope
``static __device__`` array of the library).  No counterpart in the reference.
Collected last on purpose.

Checks, through the C ABI: same d_out[8] as the existing entry point (same kernels, same grid, same fixed-order
merges -> bit-identical; asserted at the 1e-10 the divergence tests use AND with array_equal) on two calls in a
row with different inputs (the array keeps the last call's contents), no device allocation by the call, and the
size limit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _lw(n, seed, shift):
    import torch
    g = torch.Generator(device='cuda')
    g.manual_seed(seed)
    u = torch.rand(n, dtype=torch.float64, device='cuda', generator=g)
    return torch.log(torch.expm1(-0.25 * torch.log1p(-u)) / 0.25) + shift


def test_static_divergence_matches_workspace_entry_point_twice_in_a_row():
    import torch
    from viabel_b200 import _lib
    from viabel_b200._device import ptr, stream_ptr
    lib = _lib.lib
    assert lib.vb_bounds_divergence_static_max_n() == 1 << 31
    # a large call first (full 1184-CTA grid), then a small one (1 CTA): the second must not see the first's partials
    cases = [(_lw(3_000_001, 1, -4.0), 2.0), (_lw(257, 2, 1.5), 1.5), (_lw(1_000_000, 3, 0.0), 2.0)]
    ws = torch.zeros(lib.vb_bounds_divergence_workspace_bytes(1), dtype=torch.uint8, device='cuda')
    ref = torch.empty(8, dtype=torch.float64, device='cuda')
    out = torch.empty(8, dtype=torch.float64, device='cuda')
    torch.cuda.synchronize()
    for lw, alpha in cases:
        _lib.check(lib.vb_bounds_divergence(ptr(lw), lw.numel(), alpha, ptr(ref), ptr(ws), ws.numel(), stream_ptr()),
                   'vb_bounds_divergence')
        before = torch.cuda.memory_allocated()
        _lib.check(lib.vb_bounds_divergence_static(ptr(lw), lw.numel(), alpha, ptr(out), stream_ptr()),
                   'vb_bounds_divergence_static')
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        a, b = out.cpu().numpy(), ref.cpu().numpy()
        np.testing.assert_allclose(a[:7], b[:7], rtol=1e-10)
        assert np.array_equal(a[:7], b[:7])
        # and against a plain fp64 evaluation (bounds.py:169-176)
        M = lw.max().item()
        w = torch.exp(alpha * (lw - M))
        np.testing.assert_allclose(a[0], M, rtol=0)
        np.testing.assert_allclose(a[1], w.mean().item(), rtol=1e-10)
        np.testing.assert_allclose(a[3], lw.mean().item(), rtol=1e-10, atol=1e-12)
        assert a[6] == lw.numel()


def test_static_divergence_rejects_sizes_beyond_the_maximum_and_bad_arguments():
    import torch
    from viabel_b200 import _lib
    from viabel_b200._device import ptr, stream_ptr
    lib = _lib.lib
    lw = _lw(1024, 4, 0.0)
    out = torch.empty(8, dtype=torch.float64, device='cuda')
    rc = lib.vb_bounds_divergence_static(ptr(lw), (1 << 31) + 1, 2.0, ptr(out), stream_ptr())     # rejected before any launch
    assert rc == -5
    assert b'fixed scratch' in lib.vb_error_string(rc)
    with pytest.raises(_lib.ViabelB200Error, match='fixed scratch'):
        _lib.check(rc, 'vb_bounds_divergence_static')
    assert lib.vb_bounds_divergence_static(0, 1024, 2.0, ptr(out), stream_ptr()) == -1
    assert lib.vb_bounds_divergence_static(ptr(lw), 0, 2.0, ptr(out), stream_ptr()) == -1
    assert lib.vb_bounds_divergence_static(ptr(lw), 1024, 0.0, ptr(out), stream_ptr()) == -1
    torch.cuda.synchronize()
