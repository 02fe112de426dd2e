"""Stress test standing in for compute-sanitizer's racecheck (closed on this GPU pool): every alternative code path
is forced through its environment switch and must reproduce the default path -- bit for bit where the arithmetic is
the same, within the stated tolerance where it is not -- and the kernels with atomic tickets, last-CTA merges, grid
barriers or mbarrier pipelines are run repeatedly on the same input and must return identical bits every time.
Switches read once per process are exercised in subprocesses."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import sys, warnings
import numpy as np, torch
sys.path.insert(0, %(root)r)
from viabel_b200 import psis, models, vb
import viabel_b200
out = {}
rs = np.random.RandomState(11)
# ---- PSIS, three sizes: direct sort, bitonic-sized candidate set, counting-sort-sized candidate set
for n in (3000, 400000, 9000000):
    u = rs.rand(n)
    lw = np.log(np.expm1(-0.5 * np.log1p(-u)) / 0.5)
    o, k, det = psis.psislw_detail(torch.from_numpy(lw).cuda())
    out['psis_%%d' %% n] = o.cpu().numpy(); out['psis_k_%%d' %% n] = np.array([k]); out['psis_t_%%d' %% n] = det['tail_idx_sorted']
    # misaligned input AND output base (8-byte but not 16-byte aligned): the scalar-load variants
    buf = torch.empty(n + 1, dtype=torch.float64, device='cuda'); buf[1:] = torch.from_numpy(lw).cuda()
    o2, k2 = psis.psislw(buf[1:])
    out['psis_mis_%%d' %% n] = o2.cpu().numpy()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        out['div_%%d' %% n] = np.array([viabel_b200.divergence_bound(buf[1:].clone(), alpha=2.0), viabel_b200.divergence_bound(buf[1:], alpha=2.0)])
# ---- psisloo (columns kernel: selection mode vs full sort)
ll = -np.log(np.expm1(-0.5 * np.log1p(-rs.rand(4000, 64))) / 0.5)
loo, loos, ks = psis.psisloo(ll)
out['loo'] = loos; out['loo_k'] = ks
# ---- regression plugin: fused fp16x3 kernel vs the two-GEMM 3xTF32 pipeline
N, d, n = 30000, 64, 300
X = (rs.randn(N, d) / np.sqrt(d)).astype(np.float32); bt = rs.randn(d)
y = (X.astype(np.float64) @ bt + rs.standard_t(40, N)).astype(np.float32)
b = (bt + 0.1 * rs.randn(n, d)).astype(np.float32)
m = models.StudentTRegression(X, y, df=40.0, dtype='f32')
lp, g = m.logp_and_grad(b, dtype='f32')
out['reg_lp'] = lp; out['reg_g'] = g
# ---- full-rank CHIVI, d = 40: chain kernels vs the eigh formulation
dd, nn = 40, 20000
A = rs.randn(dd, dd); mp = rs.randn(dd); Pp = np.linalg.inv(A @ A.T / dd + np.eye(dd))
fam = vb.t_variational_family(dd, 40)
lam = 0.1 * rs.randn(fam.var_param_dim)
base = {'chi2': rs.chisquare(40, nn), 'z': rs.randn(nn, dd)}
v, gr = vb.black_box_chivi(2.0, fam, models.CorrelatedGaussian(mp, Pp), nn, group=False)(lam, base_draws=base)
out['fr_v'] = np.array([v]); out['fr_g'] = gr
# ---- repeat-and-compare (tickets, last-CTA merges, grid barriers, mbarrier pipelines): identical bits every time
lw = torch.from_numpy(np.log(np.expm1(-0.5 * np.log1p(-rs.rand(2000000))) / 0.5)).cuda()
ref_o, ref_k = psis.psislw(lw)
f8 = vb.mean_field_t_variational_family(10, 40)
lam8 = np.concatenate([0.3 * rs.randn(10), np.full(10, -0.1)])
eps8 = f8.impl.base_draws(1 << 20, 3, 0)['eps']
obj8 = vb.black_box_chivi(2.0, f8, models.EightSchoolsNCP(), 1 << 20, group=False)
r8 = obj8(lam8, base_draws=eps8)
objfr = vb.black_box_chivi(2.0, fam, models.CorrelatedGaussian(mp, Pp), nn, group=False)
same = True
for rep in range(6):
    o, k = psis.psislw(lw)
    same &= bool(torch.equal(o, ref_o)) and k == ref_k
    lp2, g2 = m.logp_and_grad(b, dtype='f32')
    same &= np.array_equal(lp2, lp) and np.array_equal(g2, g)
    v2, gr2 = objfr(lam, base_draws=base)
    same &= (v2 == v) and np.array_equal(gr2, gr)
    a8 = obj8(lam8, base_draws=eps8)
    same &= (a8[0] == r8[0]) and np.array_equal(a8[1], r8[1])
out['repeat_ok'] = np.array([int(same)])
np.savez(sys.argv[1], **out)
'''


def _run(tmp_path, name, env_extra):
    env = dict(os.environ)
    env.update(env_extra)
    path = str(tmp_path / (name + '.npz'))
    res = subprocess.run([sys.executable, '-c', WORKER % {'root': ROOT}, path], cwd=ROOT, env=env, capture_output=True,
                         text=True, timeout=900)
    assert res.returncode == 0, name + ': ' + res.stdout[-1500:] + res.stderr[-3000:]
    return dict(np.load(path))


def test_every_fallback_path_reproduces_the_default_path(tmp_path):
    base = _run(tmp_path, 'default', {})
    assert base['repeat_ok'][0] == 1
    for n in (3000, 400000, 9000000):                   # misaligned base pointers: same bits as the aligned run
        assert np.array_equal(base['psis_mis_%d' % n], base['psis_%d' % n])
        assert base['div_%d' % n][0] == base['div_%d' % n][1]
    exact = [k for k in base if k.startswith(('psis_', 'loo'))]
    variants = {
        'host_driven_psis': ({'VB_PSIS_HOST': '1'}, exact),
        'bitonic_sort': ({'VB_PSIS_BITONIC': '1'}, exact),
        'bitonic_host': ({'VB_PSIS_BITONIC': '1', 'VB_PSIS_HOST': '1'}, exact),
        'columns_full_sort': ({'VB_PSIS_COLUMNS_FULL': '1'}, exact),
    }
    for name, (env, keys) in variants.items():
        got = _run(tmp_path, name, env)
        assert got['repeat_ok'][0] == 1, name
        for k in keys:
            if k.startswith('psis_k') or k == 'loo_k' or k == 'loo':
                np.testing.assert_allclose(got[k], base[k], rtol=1e-12, err_msg=name + ' ' + k)      # summation order of the fit
            elif k.startswith('psis_t'):
                assert np.array_equal(got[k], base[k]), name + ' ' + k                               # tail permutation: exact
            else:
                np.testing.assert_allclose(got[k], base[k], rtol=1e-12, atol=1e-12, err_msg=name + ' ' + k)
    # different arithmetic: stated tolerances
    tf32 = _run(tmp_path, 'two_gemm_tf32', {'VB_REG_PATH': '1'})
    np.testing.assert_allclose(tf32['reg_lp'], base['reg_lp'], rtol=2e-5)
    assert np.max(np.abs(tf32['reg_g'] - base['reg_g'])) <= 4e-5 * np.max(np.abs(base['reg_g']))
    for raw in ('1', '2', '3'):
        r = _run(tmp_path, 'tf32_raw' + raw, {'VB_REG_PATH': '1', 'VB_REG_RAW': raw})
        assert np.max(np.abs(r['reg_g'] - base['reg_g'])) <= 4e-5 * np.max(np.abs(base['reg_g']))
    for dg in ('1', '2'):                               # other (draw group x row range) schedules of the fused kernel
        r = _run(tmp_path, 'fused_dg' + dg, {'VB_REG_DG': dg, 'VB_REG_FLUSH': '7'})
        np.testing.assert_allclose(r['reg_lp'], base['reg_lp'], rtol=1e-6)
        assert np.max(np.abs(r['reg_g'] - base['reg_g'])) <= 2e-5 * np.max(np.abs(base['reg_g']))
        assert r['repeat_ok'][0] == 1
    eigh = _run(tmp_path, 'fr_eigh', {'VB_FR_EIGH': '1'})
    np.testing.assert_allclose(eigh['fr_v'], base['fr_v'], rtol=1e-11)
    assert np.max(np.abs(eigh['fr_g'] - base['fr_g'])) <= 1e-9 * np.max(np.abs(base['fr_g']))
