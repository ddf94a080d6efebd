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
* Sampler numerics (adaptation schedule, checkpoint indices, iterative U-turn
  truth table, Welford regulariser, dual averaging, velocity Verlet,
  diagnostics known answers): PINNED against the golden vectors held by the
  reference's own tests (``test/infer/test_hmc_util.py``,
  ``test/test_diagnostics.py``) -- see ``tests/test_oracle_golden.py``.
* Threefry-2x32: pinned against the Random123 known-answer vectors and against
  the values printed in JAX's own documentation.  ``jax``/``jaxlib`` are not
  vendored under ``/root/reference`` (``setup.py`` pins only ``jax>=0.4.25``)
  and cannot be imported in the build container, so exact *sample streams*
  (tier 3) are "parity unpinned" against a live reference run: they are checked
  CUDA-vs-oracle only.
"""
