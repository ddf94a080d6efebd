"""numpyro_b200: a B200-native HMC/NUTS sampling hot path behind NumPyro's kernel API.

``infer.NUTS`` / ``infer.HMC`` / ``infer.MCMC`` mirror ``numpyro.infer``; the transition
kernel, the potentials of the registered model families (``models``) and the Threefry
generator (``random``) run in hand-written CUDA (``csrc/``) behind the C ABI declared in
``include/b200mc.h``.  There is no CPU path.
"""
from . import diagnostics, distributed, handlers, infer, models, random  # noqa: F401

__version__ = "0.1.0"
