"""GPU parity for the GEMM-shaped pieces (fp64 path): vb_gemm, the correlated-Gaussian and
Student-t regression plugins, the generic mean-field reduction, and the full-rank Student-t
family with KLVI / CHIVI (sqrtm = reference semantics, chol = documented fast mode)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import models_np as mm          # noqa: E402
from oracle import vb_np                    # noqa: E402


def rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


@pytest.mark.parametrize('m,n,k', [(1, 1, 1), (5, 7, 3), (128, 64, 16), (129, 65, 17), (300, 200, 1000)])
@pytest.mark.parametrize('ta,tb', [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_gemm_f64(m, n, k, ta, tb):
    import torch
    from viabel_b200 import _lib
    from viabel_b200._device import ptr, stream_ptr
    rs = np.random.RandomState(m * 7 + n)
    A = rs.randn(*((k, m) if ta else (m, k)))
    B = rs.randn(*((n, k) if tb else (k, n)))
    C0 = rs.randn(m, n)
    Ad, Bd, Cd = (torch.from_numpy(x).cuda() for x in (A, B, C0.copy()))
    _lib.check(_lib.lib.vb_gemm(_lib.F64, ta, tb, m, n, k, 0.7, ptr(Ad), A.shape[1], ptr(Bd), B.shape[1],
                                -0.3, ptr(Cd), n, 0, 0, stream_ptr()))
    want = 0.7 * (A.T if ta else A) @ (B.T if tb else B) - 0.3 * C0
    assert rel(Cd.cpu().numpy(), want) < 1e-13


def _gauss(d, seed=4):
    rs = np.random.RandomState(seed)
    A = rs.randn(d, d)
    Sp = A @ A.T / d + np.eye(d)
    mp = np.random.RandomState(seed + 1).randn(d)
    return mp, np.linalg.inv(Sp)


def test_gauss_plugin():
    from viabel_b200 import models
    d = 37
    mp, Pp = _gauss(d)
    model, omodel = models.CorrelatedGaussian(mp, Pp), mm.CorrelatedGaussian(mp, Pp)
    x = np.random.RandomState(0).randn(1000, d) * 2
    lp, g = model.logp_and_grad(x)
    olp, og = omodel.logp_and_grad(x)
    assert rel(lp, olp) < 1e-12 and rel(g, og) < 1e-12
    assert rel(model(x), olp) < 1e-12


def _regression(N, d, seed=1):
    rs = np.random.RandomState(seed)
    X = rs.randn(N, d) / np.sqrt(d)
    bt = np.random.RandomState(seed + 1).randn(d)
    y = X @ bt + rs.standard_t(40, N)
    return X, y - y.mean(), bt


@pytest.mark.parametrize('N,d,n', [(50, 3, 7), (1000, 16, 200), (5000, 33, 129)])
def test_regression_plugin(N, d, n):
    from viabel_b200 import models
    X, y, bt = _regression(N, d)
    model, omodel = models.StudentTRegression(X, y, df=40.0), mm.StudentTRegression(X, y, 40.0)
    b = bt + 0.1 * np.random.RandomState(3).randn(n, d)
    lp, g = model.logp_and_grad(b)
    olp, og = omodel.logp_and_grad(b)
    assert rel(lp, olp) < 1e-12 and rel(g, og) < 1e-11


@pytest.mark.parametrize('N,d,n', [(300, 8, 40), (5000, 32, 129), (40000, 128, 512), (70001, 64, 300)])
def test_regression_plugin_tensor_path(N, d, n):
    """fp32 storage: both design-matrix products on tcgen05 (3xTF32), fused residual / psi epilogue,
    fp64 accumulation across row chunks.  Stated tolerance: log p 1e-5, gradient 2e-5 (max-norm relative; the
    floor is the fp32 rounding of psi and X, not the 3xTF32 products)."""
    from viabel_b200 import models
    X, y, bt = _regression(N, d)
    X32, y32 = X.astype(np.float32), y.astype(np.float32)
    model = models.StudentTRegression(X32, y32, df=40.0, dtype='f32')
    omodel = mm.StudentTRegression(X32.astype(np.float64), y32.astype(np.float64), 40.0)
    b = (bt + 0.1 * np.random.RandomState(3).randn(n, d)).astype(np.float32)
    lp, g = model.logp_and_grad(b, dtype='f32')
    olp, og = omodel.logp_and_grad(b.astype(np.float64))
    assert rel(lp, olp) < 1e-5, rel(lp, olp)
    assert rel(g, og) < 2e-5, rel(g, og)
    np.testing.assert_allclose(lp, olp, rtol=2e-5)
    lp_only = model(b, dtype='f32')
    np.testing.assert_allclose(lp_only, lp, rtol=1e-12)


@pytest.mark.parametrize('raw', ['1', '2', '3'])
def test_regression_tensor_path_raw_operand_variants(raw, monkeypatch):
    """VB_REG_RAW selects the GEMM variants that read X as plain fp32 and split it into tf32 hi / lo on chip
    (converter warps; GEMM1 bit 0, GEMM2 bit 1).  Same stated tolerance as the default pre-split path."""
    from viabel_b200 import models
    monkeypatch.setenv('VB_REG_RAW', raw)
    monkeypatch.setenv('VB_REG_PATH', '1')          # the two-GEMM 3xTF32 pipeline (the fused fp16x3 kernel is the default)
    N, d, n = 5000, 64, 200
    X, y, bt = _regression(N, d)
    X32, y32 = X.astype(np.float32), y.astype(np.float32)
    model = models.StudentTRegression(X32, y32, df=40.0, dtype='f32')
    omodel = mm.StudentTRegression(X32.astype(np.float64), y32.astype(np.float64), 40.0)
    b = (bt + 0.1 * np.random.RandomState(3).randn(n, d)).astype(np.float32)
    lp, g = model.logp_and_grad(b, dtype='f32')
    olp, og = omodel.logp_and_grad(b.astype(np.float64))
    assert rel(lp, olp) < 1e-5, rel(lp, olp)
    assert rel(g, og) < 2e-5, rel(g, og)


def test_regression_tensor_path_wide_model():
    """d > 128: the coefficient tile no longer fits tensor memory, GEMM1 falls back to streamed split operands."""
    from viabel_b200 import models
    N, d, n = 3000, 160, 130
    X, y, bt = _regression(N, d)
    X32, y32 = X.astype(np.float32), y.astype(np.float32)
    model = models.StudentTRegression(X32, y32, df=40.0, dtype='f32')
    omodel = mm.StudentTRegression(X32.astype(np.float64), y32.astype(np.float64), 40.0)
    b = (bt + 0.1 * np.random.RandomState(3).randn(n, d)).astype(np.float32)
    lp, g = model.logp_and_grad(b, dtype='f32')
    olp, og = omodel.logp_and_grad(b.astype(np.float64))
    assert rel(lp, olp) < 1e-5, rel(lp, olp)
    assert rel(g, og) < 2e-5, rel(g, og)


def test_regression_tensor_path_long_accumulation_chains():
    """Half a million rows, draws close to the optimum (the gradient is a small difference of large sums): the
    production regime.  GEMM2's split-K chains reach their cap (256 k tiles = 8192 rows per tensor-memory
    accumulation).  The tensor core's fp32 accumulate truncates (round toward zero), which biases GEMM1's X b by
    about 7e-7 relative; near the optimum that bias is amplified by |b| / |b - b*| in the gradient.  Measured on
    B200: 4.6e-5 (3.7e-5 with 16-k-tile chains, 5.6e-5 with 1024).  Stated tolerance here: 1e-4 on the gradient,
    1e-5 on log p -- far below the Monte Carlo noise of the objective these gradients feed."""
    from viabel_b200 import models
    N, d, n = 1 << 19, 128, 128
    X, y, bt = _regression(N, d)
    X32, y32 = X.astype(np.float32), y.astype(np.float32)
    model = models.StudentTRegression(X32, y32, df=40.0, dtype='f32')
    omodel = mm.StudentTRegression(X32.astype(np.float64), y32.astype(np.float64), 40.0)
    b = (bt + 0.02 * np.random.RandomState(3).randn(n, d)).astype(np.float32)
    lp, g = model.logp_and_grad(b, dtype='f32')
    olp, og = omodel.logp_and_grad(b.astype(np.float64))
    assert rel(lp, olp) < 1e-5, rel(lp, olp)
    assert rel(g, og) < 1e-4, rel(g, og)


def test_mean_field_regression_tensor_path_klvi():
    """Config-2 in miniature through the public objective with fp32 storage."""
    from viabel_b200 import models, vb
    N, d, n = 20000, 128, 256
    X, y, bt = _regression(N, d)
    X32, y32 = X.astype(np.float32), y.astype(np.float32)
    model = models.StudentTRegression(X32, y32, df=40.0, dtype='f32')
    omodel = mm.StudentTRegression(X32.astype(np.float64), y32.astype(np.float64), 40.0)
    fam, ofam = vb.mean_field_t_variational_family(d, 40, dtype='f32'), vb_np.mean_field_t_variational_family(d, 40)
    lam = np.concatenate([bt + 0.01 * np.random.RandomState(3).randn(d), np.full(d, np.log(1e-2))])
    base = ofam.impl.draw_base(np.random.RandomState(9), n)
    base32 = {'eps': base['eps'].astype(np.float32)}
    v, g = vb.black_box_klvi(fam, model, n, group=False)(lam, base_draws=base32)
    ov, og = vb_np.klvi_value_and_grad(ofam.impl, omodel, lam, {'eps': base32['eps'].astype(np.float64)})
    np.testing.assert_allclose(v, ov, rtol=1e-6)
    assert rel(g, og) < 2e-5, rel(g, og)


@pytest.mark.parametrize('kind', ['gauss', 't'])
def test_mean_field_with_regression_plugin(kind):
    """Config-2 shape in miniature: the unfused path sample -> plugin -> vb_mf_reduce."""
    from viabel_b200 import models, vb
    N, d, n = 3000, 24, 500
    X, y, bt = _regression(N, d)
    model, omodel = models.StudentTRegression(X, y, df=40.0), mm.StudentTRegression(X, y, 40.0)
    if kind == 'gauss':
        fam, ofam = vb.mean_field_gaussian_variational_family(d), vb_np.mean_field_gaussian_variational_family(d)
    else:
        fam, ofam = vb.mean_field_t_variational_family(d, 40), vb_np.mean_field_t_variational_family(d, 40)
    lam = np.concatenate([bt + 0.01 * np.random.RandomState(3).randn(d), np.full(d, np.log(0.05))])
    base = ofam.impl.draw_base(np.random.RandomState(9), n)
    v, g = vb.black_box_klvi(fam, model, n, group=False)(lam, base_draws=base)
    ov, og = vb_np.klvi_value_and_grad(ofam.impl, omodel, lam, base)
    np.testing.assert_allclose(v, ov, rtol=1e-11)
    assert rel(g, og) < 1e-10
    v, g = vb.black_box_chivi(2.0, fam, model, n, group=False)(lam, base_draws=base)
    ov, og = vb_np.chivi_value_and_grad(ofam.impl, omodel, 2.0, lam, base)
    np.testing.assert_allclose(v, ov, rtol=1e-11)
    assert rel(g, og) < 1e-9
    # Philox draws through the same unfused path
    v, g = vb.black_box_klvi(fam, model, n, group=False)(lam)
    assert np.isfinite(v) and np.all(np.isfinite(g))


@pytest.mark.parametrize('scale', ['sqrtm', 'chol'])
@pytest.mark.parametrize('d,n', [(4, 50), (16, 1000), (70, 3001)])
def test_full_rank_t(scale, d, n):
    from viabel_b200 import models, vb
    mp, Pp = _gauss(d)
    model, omodel = models.CorrelatedGaussian(mp, Pp), mm.CorrelatedGaussian(mp, Pp)
    fam, ofam = vb.t_variational_family(d, 40, scale=scale), vb_np.t_variational_family(d, 40, scale=scale)
    assert fam.var_param_dim == d + d * (d + 1) // 2
    lam = 0.1 * np.random.RandomState(6).randn(fam.var_param_dim)
    base = ofam.impl.draw_base(np.random.RandomState(5), n)
    v, g = vb.black_box_klvi(fam, model, n, group=False)(lam, base_draws=base)
    ov, og = vb_np.klvi_value_and_grad(ofam.impl, omodel, lam, base)
    np.testing.assert_allclose(v, ov, rtol=1e-11)
    assert rel(g, og) < 1e-10
    v, g = vb.black_box_chivi(2.0, fam, model, n, group=False)(lam, base_draws=base)
    ov, og = vb_np.chivi_value_and_grad(ofam.impl, omodel, 2.0, lam, base)
    np.testing.assert_allclose(v, ov, rtol=1e-11)
    assert rel(g, og) < 1e-9


@pytest.mark.parametrize('span', [1.5, 3.0, 4.5])
def test_full_rank_chain_ill_conditioned(span):
    """The chain kernels on scale factors whose spectrum reaches cond(L) = 11, 2e2, 1.1e5 (cond(Sigma) = 1.3e10: beyond that the oracle's own eigh of Sigma has no digits left): the polar schedule is built
    for sigma_min / sigma_max >= 1e-3 and re-enters in cycles below that; the Smith doubling and the inverse iteration
    count up with the condition number.  Tolerance scales with cond(L) (the Sylvester solve is that ill-conditioned:
    the eigh formulation of the oracle loses the same digits)."""
    from viabel_b200 import models, vb
    d, n = 40, 2000
    mp, Pp = _gauss(d)
    model, omodel = models.CorrelatedGaussian(mp, Pp), mm.CorrelatedGaussian(mp, Pp)
    fam, ofam = vb.t_variational_family(d, 40), vb_np.t_variational_family(d, 40)
    rs = np.random.RandomState(8)
    lam = 0.05 * rs.randn(fam.var_param_dim)
    mu_, L, _ = ofam.impl.fold(lam)
    r, c = np.tril_indices(d)
    diag_pos = d + np.flatnonzero(r == c)
    lam[diag_pos] = np.linspace(-span, 0.5, d)[rs.permutation(d)]
    mu_, L, _ = ofam.impl.fold(lam)
    cond = np.linalg.cond(L)
    base = ofam.impl.draw_base(np.random.RandomState(5), n)
    v, g = vb.black_box_chivi(2.0, fam, model, n, group=False)(lam, base_draws=base)
    info = fam.impl.chain_info()
    assert info[3] == 0, info
    ov, og = vb_np.chivi_value_and_grad(ofam.impl, omodel, 2.0, lam, base)
    np.testing.assert_allclose(v, ov, rtol=1e-13 * max(cond, 100.0))
    assert rel(g, og) < 1e-14 * cond + 1e-12, (rel(g, og), cond, info)      # measured 7e-15, 1e-13, 4e-11
    if span > 4:
        assert info[0] > 8, info          # the spectrum reaches below 1e-3: the schedule re-entered


@pytest.mark.parametrize('d,n', [(96, 5001), (70, 3000), (5, 257)])
def test_full_rank_draws_carry_the_operand_planes(d, n):
    """fp32 storage: base_draws() writes the 3xTF32 planes of z with the draws (vb_base_draws_fr_split) and the step's
    weighted outer product consumes them (vb_fr_weighted_split).  Bit-identical to the separate draw / split passes
    (d = 70, 5: row pitch padded to a multiple of four, padding columns zero)."""
    import torch
    from viabel_b200 import _lib, models, vb
    from viabel_b200._device import ptr, stream_ptr
    fam = vb.t_variational_family(d, 40, dtype='f32')
    b = fam.impl.base_draws(n, 7, 3, index0=11)
    assert b['z_planes'].numel() * 4 == _lib.lib.vb_split_bytes(n, d)
    z2 = torch.empty_like(b['z'])
    c2 = torch.empty_like(b['chi2'])
    _lib.check(_lib.lib.vb_base_draws(_lib.FAMILY_FR_T, 40.0, _lib.F32, n, d, 7, 3, 11, ptr(z2), ptr(c2), stream_ptr()))
    assert torch.equal(z2, b['z']) and torch.equal(c2, b['chi2'])
    ref = torch.empty_like(b['z_planes'])
    _lib.check(_lib.lib.vb_split_f32(ptr(b['z']), n, d, d, ptr(ref), stream_ptr()))
    used = 2 * n * ((d + 3) // 4 * 4)                    # vb_split_bytes rounds up to 256 bytes: compare what is written
    assert torch.equal(ref[:used], b['z_planes'][:used])
    mp, Pp = _gauss(d)
    lam = 0.1 * np.random.RandomState(6).randn(fam.var_param_dim)
    for make in (lambda: vb.black_box_klvi(fam, models.CorrelatedGaussian(mp, Pp), n, group=False),
                 lambda: vb.black_box_chivi(2.0, fam, models.CorrelatedGaussian(mp, Pp), n, group=False)):
        v1, g1 = make()(lam, base_draws=b)                                       # planes used
        v2, g2 = make()(lam, base_draws={'z': b['z'], 'chi2': b['chi2']})        # z split by the step
        assert v1 == v2 and np.array_equal(g1, g2)


@pytest.mark.parametrize('d', [600, 777])
def test_full_rank_chain_beyond_one_tile_per_cta(d):
    """d > 512: more output tiles than CTAs (every CTA walks several tiles through the same shared-memory ring and
    mbarrier phases), d = 777 odd (8-byte copies, ragged edge tiles of both the full and the symmetric products)."""
    from viabel_b200 import models, vb
    rs = np.random.RandomState(d)
    A = rs.randn(d, d)
    mp, Pp = rs.randn(d), np.linalg.inv(A @ A.T / d + np.eye(d))
    fam, ofam = vb.t_variational_family(d, 40), vb_np.t_variational_family(d, 40)
    lam = 0.02 * rs.randn(fam.var_param_dim)
    n = 300
    base = ofam.impl.draw_base(np.random.RandomState(5), n)
    v, g = vb.black_box_chivi(2.0, fam, models.CorrelatedGaussian(mp, Pp), n, group=False)(lam, base_draws=base)
    assert fam.impl.chain_info()[3] == 0
    ov, og = vb_np.chivi_value_and_grad(ofam.impl, mm.CorrelatedGaussian(mp, Pp), 2.0, lam, base)
    np.testing.assert_allclose(v, ov, rtol=1e-12)
    assert rel(g, og) < 1e-11


def test_full_rank_family_closures():
    import torch
    from viabel_b200 import vb
    d = 12
    fam, ofam = vb.t_variational_family(d, 9), vb_np.t_variational_family(d, 9)
    lam = 0.2 * np.random.RandomState(2).randn(fam.var_param_dim)
    x = np.random.RandomState(3).randn(500, d) * 2
    assert rel(fam.logdensity(x, lam), ofam.logdensity(x, lam)) < 1e-11
    # sample(): Philox draws pushed through the reference's map x = mu + z sqrtm(Sigma) / s
    n = 4000
    s = fam.sample(lam, n, seed=11)
    bd = fam.impl.base_draws(n, 11, 0)
    want = ofam.impl.transform(lam, {'z': bd['z'].cpu().numpy(), 'chi2': bd['chi2'].cpu().numpy()})
    assert rel(s, want) < 1e-12
    # log q of reparameterised samples equals the base-draw shortcut used inside the step
    lq = fam.logdensity(torch.from_numpy(s).cuda(), lam).cpu().numpy()
    want_lq = ofam.impl.logq_of_base(lam, {'z': bd['z'].cpu().numpy(), 'chi2': bd['chi2'].cpu().numpy()})
    assert rel(lq, want_lq) < 1e-10
    # Philox objective runs end to end
    from viabel_b200 import models
    mp, Pp = _gauss(d)
    v, g = vb.black_box_chivi(2.0, fam, models.CorrelatedGaussian(mp, Pp), 2000, group=False)(lam)
    assert np.isfinite(v) and g.shape == (fam.var_param_dim,) and np.all(np.isfinite(g))


@pytest.mark.parametrize('m,n,k', [(128, 128, 32), (128, 128, 256), (256, 384, 96), (1000, 200, 520), (77, 130, 36)])
@pytest.mark.parametrize('ta,tb', [(0, 1), (0, 0), (1, 1), (1, 0)])
def test_gemm_tf32x3(m, n, k, ta, tb):
    """3xTF32 tcgen05 GEMM against fp64 numpy.  Stated tolerance: 2e-6 * sum_k |a||b| per entry
    (measured ~3e-7); a single-pass TF32 product would be ~5e-4."""
    import torch
    from viabel_b200 import _lib
    from viabel_b200._device import ptr, stream_ptr
    rs = np.random.RandomState(m + n + k)
    A = rs.randn(*((k, m) if ta else (m, k))).astype(np.float32)
    B = rs.randn(*((n, k) if tb else (k, n))).astype(np.float32)
    C0 = rs.randn(m, n).astype(np.float32)
    Ad, Bd, Cd = (torch.from_numpy(x).cuda() for x in (A, B, C0.copy()))
    nb = _lib.lib.vb_gemm_workspace_bytes(_lib.F32, m, n, k)
    ws = torch.zeros(nb, dtype=torch.uint8, device='cuda')
    _lib.check(_lib.lib.vb_gemm(_lib.F32, ta, tb, m, n, k, 0.5, ptr(Ad), A.shape[1], ptr(Bd), B.shape[1],
                                2.0, ptr(Cd), n, ptr(ws), ws.numel(), stream_ptr()))
    torch.cuda.synchronize()
    opA = (A.T if ta else A).astype(np.float64)
    opB = (B.T if tb else B).astype(np.float64)
    want = 0.5 * opA @ opB + 2.0 * C0
    bound = 2e-6 * (np.abs(opA) @ np.abs(opB)) + 1e-6 * np.abs(want)
    err = np.abs(Cd.cpu().numpy() - want)
    assert np.all(err <= bound), (err.max(), (err / bound).max())


@pytest.mark.parametrize('scale', ['sqrtm', 'chol'])
@pytest.mark.parametrize('d,n', [(16, 1000), (96, 5000), (200, 4097)])
def test_full_rank_t_f32_tensor_path(scale, d, n):
    """fp32 storage + 3xTF32 tcgen05 products + fp64 reductions against the fp64 oracle on the same
    (fp32-rounded) base draws.  Stated tolerance: value rtol 5e-6 (KLVI) / 2e-5 (CHIVI), gradient 1e-4 / 2e-3 of its max entry."""
    from viabel_b200 import models, vb
    mp, Pp = _gauss(d)
    model, omodel = models.CorrelatedGaussian(mp, Pp), mm.CorrelatedGaussian(mp, Pp)
    fam = vb.t_variational_family(d, 40, dtype='f32', scale=scale)
    ofam = vb_np.t_variational_family(d, 40, scale=scale)
    lam = 0.1 * np.random.RandomState(6).randn(fam.var_param_dim)
    base = ofam.impl.draw_base(np.random.RandomState(5), n)
    b32 = {'z': base['z'].astype(np.float32), 'chi2': base['chi2'].astype(np.float32)}
    br = {'z': b32['z'].astype(np.float64), 'chi2': b32['chi2'].astype(np.float64)}
    v, g = vb.black_box_klvi(fam, model, n, group=False)(lam, base_draws=b32)
    ov, og = vb_np.klvi_value_and_grad(ofam.impl, omodel, lam, br)
    np.testing.assert_allclose(v, ov, rtol=5e-6)
    assert np.max(np.abs(g - og)) < 1e-4 * np.max(np.abs(og))
    v, g = vb.black_box_chivi(2.0, fam, model, n, group=False)(lam, base_draws=b32)
    ov, og = vb_np.chivi_value_and_grad(ofam.impl, omodel, 2.0, lam, br)
    np.testing.assert_allclose(v, ov, rtol=2e-5)
    assert np.max(np.abs(g - og)) < 2e-3 * np.max(np.abs(og))
