"""CPU oracle for the viabel Monte Carlo VI / PSIS / bounds hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in ``viabel_b200/`` imports this package.  The only
legitimate callers are ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py``; there it is the checker (or the timed CPU
baseline), never the product path.

Contents (every function cites the reference ``file:line`` it restates; citations are into
``/root/reference``):

* ``bounds_np``  – numpy restatement of ``viabel/bounds.py``.                PINNED: against
  the reference's own ``tests/test_bounds.py`` known answers, the normal-mixture notebook
  golden numbers, and A/B against the imported reference module (``tests/golden``).
* ``psis_np``    – numpy restatement of ``notebooks/psis.py``.               PINNED: A/B
  against the imported reference module run in the build container (``tests/golden``,
  generator ``scripts/make_golden.py``).
* ``models_np``  – the model log densities with analytic gradients
  (funnel, normal mixture, Student-t regression, eight-schools ncp, correlated Gaussian).
* ``vb_np``      – closed-form numpy restatement of ``viabel/vb.py`` families, KLVI, CHIVI,
  ``adagrad_optimize``.  This is the timed CPU baseline.
* ``vb_autograd``– literal torch-CPU-fp64-autograd transcription of ``viabel/vb.py`` /
  ``viabel/_distributions.py`` (autograd → torch.autograd); used to validate ``vb_np``.

PARITY STATUS of the VI path (``vb_np`` / ``vb_autograd``): **parity unpinned**.
``viabel/vb.py`` cannot be imported anywhere we can run (its dependencies ``autograd``,
``paragami`` and ``pystan`` are absent and there is no network) and the reference ships no
test, fixture or golden vector for it (``README.md:66``).  What *is* pinned: family
``logdensity`` against ``scipy.stats`` (the arithmetic ``autograd.scipy`` wraps), and the
closed-form gradients against torch autograd of the literal transcription.
"""
