"""``viabel._distributions`` (``viabel/_distributions.py:8-38``)."""
from viabel_b200.vb import multivariate_t_logpdf   # noqa: F401
