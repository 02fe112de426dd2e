"""``import viabel`` drop-in: the reference's package surface (``viabel/__init__.py:1``) served by ``viabel_b200``.

Put ``<repo>/compat`` and ``<repo>`` on ``PYTHONPATH``; code written against kunalghosh/viabel then runs unchanged
on the CUDA path, provided its ``logdensity`` is one of the registered model plugins (``viabel_b200.models``)."""
from viabel_b200.bounds import *          # noqa: F401,F403
from . import bounds, functions, vb       # noqa: F401
