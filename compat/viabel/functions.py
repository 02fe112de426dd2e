"""``viabel.functions`` (``viabel/functions.py:8-77``) served by ``viabel_b200.functions``."""
from viabel_b200.functions import *       # noqa: F401,F403
