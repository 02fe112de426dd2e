"""GPU parity of the device-resident AdaGrad loop (reference viabel/vb.py:345-389).

Three pins: (i) against the CPU oracle's adagrad_optimize on identical external base draws (rtol 1e-9: fp64
kernels, error compounding over the iterations); (ii) bit-for-bit against this package's own host loop, which
calls the same CUDA objective once per iteration with the same Philox streams; (iii) CUDA-graph replay against
plain launches, bit for bit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import models_np as mm          # noqa: E402
from oracle import vb_np                    # noqa: E402


def _setup(name):
    from viabel_b200 import models, vb
    if name == 'funnel':
        return (lambda: vb.mean_field_gaussian_variational_family(2), vb_np.mean_field_gaussian_variational_family(2),
                models.Funnel(), mm.Funnel(), np.array([0., -1., 1., 1.]))
    if name == 'mixture':
        return (lambda: vb.mean_field_t_variational_family(1, 40), vb_np.mean_field_t_variational_family(1, 40),
                models.NormalMixture(), mm.NormalMixture(), np.array([0.3, 0.5]))
    rs = np.random.RandomState(3)
    init = np.concatenate([0.3 * rs.randn(10), -0.2 + 0.1 * rs.randn(10)])
    return (lambda: vb.mean_field_t_variational_family(10, 40), vb_np.mean_field_t_variational_family(10, 40),
            models.EightSchoolsNCP(), mm.EightSchoolsNCP(), init)


@pytest.mark.parametrize('name', ['funnel', 'mixture', 'eight_schools'])
@pytest.mark.parametrize('objective', ['klvi', 'chivi'])
@pytest.mark.parametrize('graph_iters', [0, 7])
@pytest.mark.parametrize('mode', [True, 'two-launch'])
def test_device_loop_matches_oracle_adagrad_on_external_draws(name, objective, graph_iters, mode):
    from viabel_b200 import vb
    make, ofam, model, omodel, init = _setup(name)
    n_iters, n = 37, 513                                     # ragged: 5 graph replays of 7 + 2 plain iterations
    base = np.stack([ofam.impl.draw_base(np.random.RandomState(100 + i), n)['eps'] for i in range(n_iters)])
    it = iter(range(n_iters))
    if objective == 'klvi':
        obj = vb.black_box_klvi(make(), model, n, group=False)
        oobj = lambda lam: vb_np.klvi_value_and_grad(ofam.impl, omodel, lam, {'eps': base[next(it)]})        # noqa: E731
    else:
        obj = vb.black_box_chivi(2.0, make(), model, n, group=False)
        oobj = lambda lam: vb_np.chivi_value_and_grad(ofam.impl, omodel, 2.0, lam, {'eps': base[next(it)]})  # noqa: E731
    kw = dict(window=5, learning_rate=0.05, learning_rate_end=0.01, epsilon=0.1)
    got = vb.adagrad_optimize(n_iters, obj, init, device_loop=mode, base_draws=base, graph_iters=graph_iters, **kw)
    want = vb_np.adagrad_optimize(n_iters, oobj, init, **kw)
    assert got[1].shape == want[1].shape == (n_iters - 3 * n_iters // 4, init.size)
    np.testing.assert_allclose(got[2], want[2], rtol=1e-9, atol=1e-12)        # value history
    np.testing.assert_allclose(got[1], want[1], rtol=1e-9, atol=1e-12)        # kept iterates
    np.testing.assert_allclose(got[0], want[0], rtol=1e-9, atol=1e-12)        # averaged iterate
    assert np.array_equal(got[3], want[3])


@pytest.mark.parametrize('name,objective', [('funnel', 'klvi'), ('eight_schools', 'klvi'), ('eight_schools', 'chivi'),
                                            ('mixture', 'chivi')])
@pytest.mark.parametrize('n', [1000, 5000])      # one CTA: iterations loop inside one launch; several CTAs: graph replay
def test_device_loop_equals_host_loop_bit_for_bit(name, objective, n):
    """Same Philox streams (KLVI: the family's persistent stream; CHIVI: numpy's global RNG, vb.py:258), same
    kernel for the gradient, explicitly rounded update: every iterate and value is identical."""
    from viabel_b200 import vb
    make, _, model, _, init = _setup(name)
    n_iters = 130
    runs = []
    for device_loop, graph_iters in ((False, None), (True, 0), (True, 50), ('two-launch', 50), ('two-launch', 0)):
        fam = make()
        if objective == 'klvi':
            obj = vb.black_box_klvi(fam, model, n, group=False)
        else:
            obj = vb.black_box_chivi(2.0, fam, model, n, group=False)
        np.random.seed(11)
        runs.append(vb.adagrad_optimize(n_iters, obj, init, learning_rate=.02, learning_rate_end=.002,
                                        device_loop=device_loop, graph_iters=graph_iters))
        if objective == 'klvi':
            assert fam.impl.offset == n_iters                 # one call offset consumed per iteration
    host = runs[0]
    for dev in runs[1:]:
        for a, b in zip(host, dev):
            assert np.array_equal(np.asarray(a), np.asarray(b))


def test_device_loop_is_the_default_for_fused_objectives_and_validates():
    from viabel_b200 import models, vb
    fam = vb.mean_field_gaussian_variational_family(2)
    obj = vb.black_box_klvi(fam, models.Funnel(), 256, group=False)
    assert vb._device_loop_ok(obj, False) and not vb._device_loop_ok(obj, True)
    assert not vb._device_loop_ok(lambda lam: (0.0, lam), False)
    with pytest.raises(TypeError):
        vb.adagrad_optimize(10, lambda lam: (0.0, lam), np.zeros(2), device_loop=True)
    with pytest.raises(ValueError, match='learning rate must be positive'):
        vb.adagrad_optimize(10, obj, np.zeros(4), learning_rate=-1.0)
    with pytest.raises(ValueError, match='greater than final'):
        vb.adagrad_optimize(10, obj, np.zeros(4), learning_rate=0.01, learning_rate_end=0.1)
    opt, kept, vals, norms = vb.adagrad_optimize(3, obj, np.zeros(4))          # n_iters < 4: keep_from = 2
    assert kept.shape == (1, 4) and vals.shape == (3,) and norms.shape == (3,)
    np.testing.assert_array_equal(opt, kept[0])
