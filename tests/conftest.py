import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'tests')):
    if p not in sys.path:
        sys.path.insert(0, p)

REFERENCE = os.environ.get('VIABEL_REFERENCE', '/root/reference')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, 'tests', 'golden', 'reference_outputs.npz'))


@pytest.fixture(scope='session')
def reference_modules():
    """The reference's own bounds.py / psis.py, importable only in the build container."""
    import importlib.util
    if not os.path.isdir(REFERENCE):
        pytest.skip('reference checkout not present on this machine')
    sys.dont_write_bytecode = True

    def load(name, path):
        spec = importlib.util.spec_from_file_location(name, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    return (load('ref_bounds', os.path.join(REFERENCE, 'viabel', 'bounds.py')),
            load('ref_psis', os.path.join(REFERENCE, 'notebooks', 'psis.py')))


@pytest.fixture(scope='session', autouse=True)
def _built_library():
    """Every test session starts from a built library (nvcc cross-compiles without a GPU)."""
    from viabel_b200.build import build
    build()
