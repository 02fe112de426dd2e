"""CPU-only checks: the C-ABI library loads and exports every symbol include/viabel_b200.h
declares; host-only entry points (model handles, sizing queries, argument validation) behave;
the Python surface keeps the reference's names, defaults and error behaviour."""
import ctypes as C
import inspect
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    hdr = open(os.path.join(ROOT, 'include', 'viabel_b200.h')).read()
    hdr = re.sub(r'/\*.*?\*/', '', hdr, flags=re.S)
    return sorted(set(re.findall(r'\b(vb_[a-z0-9_]+)\s*\(', hdr)))


def test_library_exports_every_declared_symbol():
    from viabel_b200 import _lib
    lib = C.CDLL(_lib.LIB_PATH)
    names = _header_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), 'missing export ' + n
    assert sorted(_lib.SIGNATURES) == names          # the ctypes table mirrors the header exactly
    assert _lib.lib.vb_abi_version() == 1
    assert _lib.lib.vb_compiled_sm() == 100
    assert _lib.lib.vb_error_string(-3).decode() == 'workspace too small'


def test_library_is_sm100a_sass():
    """The shared object carries sm_100a SASS (and uses the TMA unit's bulk copy in the fused step)."""
    import subprocess
    from viabel_b200 import _lib
    out = subprocess.run(['cuobjdump', '-lelf', _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert 'sm_100a' in out


def test_model_handles_and_argument_errors():
    from viabel_b200 import _lib
    lib = _lib.lib
    h = C.c_void_p()
    assert lib.vb_model_create_funnel(1.35, C.byref(h)) == 0
    assert lib.vb_model_kind(h) == _lib.MODEL_FUNNEL and lib.vb_model_dim(h) == 2
    lib.vb_model_destroy(h)
    assert lib.vb_model_create_funnel(-1.0, C.byref(h)) == -1
    assert lib.vb_model_create_mixture(2.0, C.byref(h)) == 0
    assert lib.vb_model_dim(h) == 1
    lib.vb_model_destroy(h)
    y = (C.c_double * 8)(28, 8, -3, 7, -1, 1, 18, 12)
    s = (C.c_double * 8)(15, 10, 16, 11, 9, 11, 10, 18)
    assert lib.vb_model_create_eight_schools(y, s, 8, C.byref(h)) == 0
    assert lib.vb_model_dim(h) == 10
    # argument validation happens before any CUDA call
    assert lib.vb_mf_step(0, 0.0, h, 0, 1.0, 1, None, 10, None, 0, 0, 0, None, None, None, None, 0, None) == -1
    lib.vb_model_destroy(h)
    assert lib.vb_model_create_eight_schools(y, s, 5, C.byref(h)) == -2       # only J = 8 is compiled in
    s[3] = 0.0
    assert lib.vb_model_create_eight_schools(y, s, 8, C.byref(h)) == -1
    assert lib.vb_psis_smooth(None, None, 10, 1.0, None, None, None, 0, None) == -1
    assert lib.vb_bounds_divergence(None, 0, 2.0, None, None, 0, None) == -1
    assert lib.vb_max(None, 5, None, None, 0, None) == -1


def test_sizing_queries():
    from viabel_b200 import _lib
    lib = _lib.lib
    assert lib.vb_psis_tail_capacity(10 ** 9, 1.0) == 94869          # ceil(3 sqrt(1e9))
    assert lib.vb_psis_tail_capacity(100, 1.0) == 20                 # min(0.2 n, 3 sqrt n)
    assert lib.vb_psis_tail_capacity(1, 1.0) == 0
    assert lib.vb_psis_smooth_workspace_bytes(10 ** 6, 1.0) > 16 * 10 ** 4
    assert lib.vb_mf_step_workspace_bytes(10) >= 148 * 24 * 8
    assert lib.vb_mf_reduce_workspace_bytes(4096, 128) >= 2 * 4096 * 8
    assert lib.vb_bounds_divergence_workspace_bytes(10 ** 9) < 1 << 20


def test_python_surface_matches_reference_signatures():
    """Names, argument order and defaults of the kept surface."""
    import viabel_b200
    from viabel_b200 import bounds, psis, vb

    def params(f):
        return [(p.name, p.default) for p in inspect.signature(f).parameters.values()]
    E = inspect.Parameter.empty
    assert viabel_b200.bounds.__all__ == ['all_bounds', 'error_bounds', 'wasserstein_bounds', 'divergence_bound']
    assert params(bounds.all_bounds)[:6] == [('log_weights', E), ('samples', None), ('moment_bound_fn', None),
                                             ('q_var', None), ('p_var', None), ('log_norm_bound', None)]
    assert params(bounds.divergence_bound)[:4] == [('log_weights', E), ('alpha', 2.), ('log_norm_bound', None),
                                                   ('return_log_norm_bound', False)]
    assert params(bounds.wasserstein_bounds)[:3] == [('d2', E), ('samples', None), ('moment_bound_fn', None)]
    assert params(bounds.error_bounds) == [('W1', np.inf), ('W2', np.inf), ('q_var', np.inf), ('p_var', np.inf)]
    assert params(psis.psislw)[:3] == [('lw', E), ('Reff', 1.0), ('overwrite_lw', False)]      # + group=None
    # + device_loop=None, base_draws=None, graph_iters=None (extensions after the reference's parameters)
    assert params(vb.adagrad_optimize)[:8] == [('n_iters', E), ('objective_and_grad', E), ('init_param', E),
                                               ('has_log_norm', False), ('window', 10), ('learning_rate', .01),
                                               ('epsilon', .1), ('learning_rate_end', None)]
    assert params(vb.black_box_klvi)[:3] == [('var_family', E), ('logdensity', E), ('n_samples', E)]
    assert params(vb.black_box_chivi)[:4] == [('alpha', E), ('var_family', E), ('logdensity', E), ('n_samples', E)]
    assert vb.VariationalFamily._fields == ('sample', 'entropy', 'logdensity', 'mean_and_cov', 'pth_moment',
                                            'var_param_dim')


def test_family_host_closures_and_errors():
    """entropy / mean_and_cov / pth_moment are host arithmetic: check against the oracle on CPU."""
    from oracle import vb_np
    from viabel_b200 import vb
    rs = np.random.RandomState(0)
    with pytest.raises(ValueError, match='df must be greater than 2'):
        vb.mean_field_t_variational_family(3, 2)
    with pytest.raises(ValueError, match='df must be greater than 2'):
        vb.t_variational_family(3, 2)
    for mk, omk, P in [(lambda: vb.mean_field_gaussian_variational_family(4),
                        lambda: vb_np.mean_field_gaussian_variational_family(4), 8),
                       (lambda: vb.mean_field_t_variational_family(4, 9),
                        lambda: vb_np.mean_field_t_variational_family(4, 9), 8),
                       (lambda: vb.t_variational_family(4, 9), lambda: vb_np.t_variational_family(4, 9), 14)]:
        fam, ofam = mk(), omk()
        assert fam.var_param_dim == ofam.var_param_dim == P
        lam = 0.3 * rs.randn(P)
        np.testing.assert_allclose(fam.entropy(lam), ofam.entropy(lam), rtol=1e-13)
        for a, b in zip(fam.mean_and_cov(lam), ofam.mean_and_cov(lam)):
            np.testing.assert_allclose(a, b, rtol=1e-13)
        for p in (2, 4):
            np.testing.assert_allclose(fam.pth_moment(p, lam), ofam.pth_moment(p, lam), rtol=1e-12)
        with pytest.raises(ValueError, match='only p = 2 or 4 supported'):
            fam.pth_moment(3, lam)
    with pytest.raises(ValueError, match='df must be greater than p'):
        vb.mean_field_t_variational_family(2, 3.5).pth_moment(4, np.zeros(4))
    with pytest.raises(TypeError, match='registered viabel_b200.models plugin'):
        vb.black_box_klvi(vb.mean_field_gaussian_variational_family(2), lambda x: x.sum(1), 10)


def test_adagrad_matches_oracle_loop():
    from oracle import vb_np
    from viabel_b200 import vb
    target = np.array([1.0, -2.0, 0.5])
    rs = np.random.RandomState(1)
    noise = rs.randn(400, 3) * 0.1
    calls = {'a': 0, 'b': 0}

    def make(key):
        def obj(p):
            g = p - target + noise[calls[key]]
            calls[key] += 1
            return 0.5 * np.sum((p - target) ** 2), g
        return obj
    a = vb.adagrad_optimize(400, make('a'), np.zeros(3), learning_rate=0.05, learning_rate_end=0.005)
    b = vb_np.adagrad_optimize(400, make('b'), np.zeros(3), learning_rate=0.05, learning_rate_end=0.005)
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)
    assert a[1].shape == (100, 3)
    assert list(vb.learning_rate_schedule(8, .1, .01)) == list(vb_np.learning_rate_schedule(8, .1, .01))
    with pytest.raises(ValueError, match='learning rate must be positive'):
        list(vb.learning_rate_schedule(4, -1, None))
    with pytest.raises(ValueError, match='initial learning rate must be greater than final learning rate'):
        list(vb.learning_rate_schedule(4, .1, .1))


def test_bounds_and_psis_argument_errors_without_gpu():
    from viabel_b200 import bounds, psis
    with pytest.raises(ValueError, match='alpha must be greater than 1'):
        bounds.divergence_bound(np.zeros(4), alpha=1.0)
    with pytest.raises(ValueError, match='must provides samples if moment_bound_fn not given'):
        bounds.wasserstein_bounds(1.0)
    with pytest.raises(ValueError, match='1 or 2 dimensional'):
        psis.psislw(np.zeros((2, 2, 2)))
    with pytest.raises(ValueError, match='More than one log-weight needed'):
        psis.psislw(np.zeros(1))
    with pytest.raises(ValueError, match='Invalid input array'):
        psis.gpdfitnew(np.zeros((2, 2)))
    res = bounds.error_bounds(W1=1.0, W2=2.0, q_var=np.diag([4.0, 9.0]), p_var=None)
    assert res == {'mean_error': 1.0, 'std_error': 2.0, 'cov_error': 2 * (3.0 * 2.0 + 4.0)}
    res = bounds.wasserstein_bounds(0.5, moment_bound_fn=lambda p: 2.0 ** p)
    np.testing.assert_allclose(res['W1'], 2 * 2.0 * np.sqrt(np.expm1(0.5)))


def test_torch_custom_ops_registered_cuda_only():
    """torch.ops.viabel_b200.* exist and have no CPU kernel (no CPU fallback)."""
    import torch
    import viabel_b200
    from viabel_b200 import _ops
    for name in _ops.OPS:
        assert hasattr(torch.ops.viabel_b200, name)
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops.viabel_b200.logsumexp(torch.zeros(4, dtype=torch.float64))
