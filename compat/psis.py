"""``import psis`` as the notebooks do (``notebooks/psis.py``), served by ``viabel_b200.psis``."""
from viabel_b200.psis import *            # noqa: F401,F403
from viabel_b200.psis import __all__      # noqa: F401
