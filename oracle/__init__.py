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

* ``functions_np`` – numpy restatement of ``viabel/functions.py`` (R-hat family, iterate
  averaging) and of the RMSProp / Adam R-hat optimisers of ``viabel/vb.py:392-712``.   PINNED:
  golden outputs of the reference's own functions (``tests/golden/optimizers.npz``, generator
  ``scripts/make_golden_optim.py``), which also pin ``vb_np.adagrad_optimize`` and
  ``vb_np.learning_rate_schedule``.

PARITY STATUS of the VI path (``vb_np`` / ``vb_autograd``):
* mean-field families, KLVI and CHIVI ``objective_and_grad``: **PINNED** on golden vectors made
  by running the reference's own ``viabel/vb.py`` in the build container
  (``tests/golden/vb_reference.npz``, generator ``scripts/make_golden_vb.py``): ``autograd`` is
  absent, so its ``value_and_grad`` / ``vector_jacobian_product`` are stood in by complex-step
  differentiation of the reference's own objective functions (exact to rounding), see
  ``scripts/ref_stand_ins.py``.
* full-rank Student-t family: pinned on golden vectors of the reference's own
  ``t_variational_family`` / ``multivariate_t_logpdf`` / ``black_box_*`` (same file and
  generator): values, samples, log density and moments at 1e-11; gradients through a 4-point
  central difference of the reference's objective (``eigh`` / ``sqrtm`` are not complex-analytic
  as called), tolerance 1e-7, observed 1e-11.  NOT pinned: ``paragami``'s flattening layout --
  the package is absent and its three patterns are stood in from its documentation.
"""
