"""CPU oracle for the HMC/NUTS hot path -- TEST INFRASTRUCTURE ONLY.

This package is a NumPy (plus a little C, see ``potentials.c``) restatement of
the reference algorithm (NumPyro 0.18.0, ``numpyro/infer/hmc_util.py``,
``numpyro/infer/hmc.py``, ``numpyro/infer/util.py``) and of the third-party
``jax.random`` Threefry-2x32 generator the reference draws its randomness from.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it.  The product (``numpyro_b200``) never
does: it fails loudly when the CUDA library is missing.

Parity status
-------------
* Sampler (everything in ``hmc.py`` and ``hmc_util.py``): PINNED to the reference's OWN
  source.  ``tests/golden/refshim.py`` executes the unmodified ``numpyro/util.py``,
  ``numpyro/infer/hmc_util.py``, ``numpyro/infer/hmc.py`` and ``numpyro/diagnostics.py`` from
  /root/reference over a NumPy stand-in for the slice of JAX they use;
  ``tests/golden/make_refpin.py`` drives the reference's ``HMC``/``NUTS`` classes through 28
  cases (fixed parameters, diagonal / dense / block-structured warm-ups, HMC, both Threefry
  layouts, four model families) and commits the traces as ``tests/golden/refpin.npz``.  The
  oracle reproduces every field of every transition of every case BIT FOR BIT
  (``tests/test_refpin.py``).  Not reference code in those runs, and therefore restated:
  ``jax.random`` (Threefry, below), autodiff (closed-form gradients, checked against finite
  differences and float64), and the model tracing / init search of ``infer/util.py``.
* Potentials, init search (``find_valid_initial_params``), ``postprocess_fn`` and the whole
  ``MCMC.run`` stack: PINNED to the reference's own model code -- the entire unmodified
  ``numpyro`` package is imported over the stand-in (``tests/golden/refshim_full.py``) and the
  model functions of its example files are evaluated by its ``potential_energy`` /
  ``initialize_model`` / ``constrain_fn`` / ``MCMC`` (``tests/golden/make_refmodels.py`` ->
  ``refmodels.npz``; ``tests/test_refmodels.py``): potential values to f32 round-off, initial
  points and MCMC draws exactly.  Gradients are closed forms checked against central
  differences of an independent ``scipy.stats`` statement.
* Also pinned against the golden vectors held by the reference's own tests
  (``test/infer/test_hmc_util.py``, ``test/test_diagnostics.py``): adaptation schedule,
  checkpoint indices, iterative U-turn truth table, Welford regulariser, dual averaging,
  velocity Verlet, diagnostics known answers -- ``tests/test_oracle_golden.py``.
* Threefry-2x32 (third party, ``jax>=0.4.25``, not under /root/reference, not importable
  here): restated from the published algorithm, pinned against the Random123 known-answer
  vectors and the values printed in JAX's documentation.  The random STREAMS of a live
  JAX run are the one thing that stays unpinned.
"""
