"""The six minimax quintics hard-coded in csrc/fr_chain.cu (kPolarSched) are what scripts/polar_schedule.py computes:
polynomial k is the best odd-quintic approximation of 1 on the image interval of its predecessor, starting from
[1e-3, 1].  Also checks, in exact numpy arithmetic, the property the kernel relies on: six steps take every singular
value in [1e-3, 1] to 1 +- 1e-3, the cubic Newton-Schulz steps that follow converge, and values BELOW 1e-3 are only
lifted (never pushed out of the basin), which is what makes the cyclic re-entry safe."""
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _kernel_schedule():
    src = open(os.path.join(ROOT, 'viabel_b200', 'csrc', 'fr_chain.cu')).read()
    body = src[src.index('kPolarSched[6][3] = {'):]
    body = body[:body.index('};')]
    nums = [float(x) for x in re.findall(r'-?\d+\.\d+', body)]
    assert len(nums) == 18
    return np.array(nums).reshape(6, 3)


def test_kernel_constants_are_the_minimax_schedule():
    sys.path.insert(0, os.path.join(ROOT, 'scripts'))
    import polar_schedule
    sched = _kernel_schedule()
    lo, hi = 1e-3, 1.0
    for k in range(6):
        (a, b, c), t = polar_schedule.opt_quintic(lo, hi, ngrid=3000)
        np.testing.assert_allclose(sched[k], [a, b, c], rtol=2e-4, atol=2e-4)
        lo, hi = 1 - t, 1 + t


def test_schedule_maps_the_design_interval_to_one_and_only_lifts_below_it():
    sched = _kernel_schedule()
    x = np.concatenate([np.geomspace(1e-3, 1.0, 20001), np.linspace(1e-3, 1.0, 20001)])
    small = np.geomspace(1e-9, 1e-3, 2001)
    for a, b, c in sched:
        x = a * x + b * x ** 3 + c * x ** 5
        new = a * small + b * small ** 3 + c * small ** 5
        assert np.all(new > small) and np.all(new < 2.0)          # lifted, still inside [0, 1.99]
        small = new
    assert np.max(np.abs(x - 1.0)) < 1.2e-3
    for _ in range(3):                                              # cubic Newton-Schulz: 1.5 x - 0.5 x^3
        x = 1.5 * x - 0.5 * x ** 3
    assert np.max(np.abs(x - 1.0)) < 1e-15
    # re-entry: polynomials 2..6 applied to anything in [0, 1.99] stay in [0, 1.99] and end within 1e-3 of 1 for
    # inputs that entered the cycle above 0.0085 (the design interval of the second polynomial)
    y = np.linspace(0.0085, 1.99, 40001)
    z = np.geomspace(1e-12, 1.99, 4001)
    for a, b, c in sched[1:]:
        y = a * y + b * y ** 3 + c * y ** 5
        z = a * z + b * z ** 3 + c * z ** 5
        assert np.all(z >= 0) and np.all(z < 2.0)
    assert np.max(np.abs(y - 1.0)) < 1.2e-3
