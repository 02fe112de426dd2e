"""``viabel.vb`` (``viabel/vb.py:28-40``) served by ``viabel_b200.vb``; the model plugins ride along."""
from viabel_b200.vb import *              # noqa: F401,F403
from viabel_b200.vb import __all__        # noqa: F401
from viabel_b200 import models            # noqa: F401
