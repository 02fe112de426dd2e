"""``viabel.bounds`` (``viabel/bounds.py:13-213``) served by ``viabel_b200.bounds``."""
from viabel_b200.bounds import *          # noqa: F401,F403
from viabel_b200.bounds import __all__    # noqa: F401
