"""viabel_b200 — B200-native Monte Carlo variational hot path behind viabel's Python surface.

``from viabel_b200 import all_bounds, error_bounds, wasserstein_bounds, divergence_bound``
mirrors ``viabel/__init__.py:1``; ``viabel_b200.vb`` mirrors ``viabel/vb.py``'s kept surface;
``viabel_b200.psis`` mirrors ``notebooks/psis.py``; ``viabel_b200.models`` holds the model plugins.
Importing needs the compiled ``libviabel_b200.so`` (``python viabel_b200/build.py``): there is no
CPU fallback.
"""
from .bounds import *          # noqa: F401,F403
from . import bounds, functions, models, psis, vb   # noqa: F401

__version__ = '0.1.0'
from . import _ops             # noqa: F401,E402  (registers torch.ops.viabel_b200.*)
