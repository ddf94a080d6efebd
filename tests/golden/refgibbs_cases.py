"""Shared by ``make_refgibbs.py`` (reference side, NumPy stand-in for jax) and
``tests/test_refgibbs.py`` (oracle and CUDA side): the eight-schools Gibbs update of ``mu`` and
the prior draw ``init_to_sample`` takes for it -- the same float32 NumPy arithmetic on both sides,
only the source of the standard normal differs (``jax.random.normal`` of the stand-in / the
oracle's Threefry / the device RNG)."""
import numpy as np

import refpin_cases as rc

NUM_WARMUP, NUM_STEPS = 12, 22
SEEDS = (0, 1)
KERNEL_KW = dict(max_tree_depth=6)
# the same wrapper around HMC (test_hmc_gibbs.py parametrises its models over [HMC, NUTS])
HMC_KW = dict(trajectory_length=0.4, step_size=0.02)
HMC_SEED, HMC_WARMUP, HMC_STEPS = 3, 6, 14


def make_gibbs_fn(normal, wrap=lambda v: v):
    """mu | tau, theta_base, y  ~  Normal(m, 1/sqrt(p)),  p = 1/25 + sum 1/sigma^2,
    m = sum((y - tau theta_base) / sigma^2) / p   (prior Normal(0, 5), README.md:101-110)."""
    f = np.float32
    s2 = (rc.S8.astype(np.float32) ** 2).astype(np.float32)

    def gibbs_fn(rng_key, gibbs_sites, hmc_sites):
        tau = np.asarray(hmc_sites["tau"], np.float32)
        tb = np.asarray(hmc_sites["theta_base"], np.float32)
        prec = f(1.0 / 25.0) + (f(1) / s2).sum(dtype=np.float32)
        mean = ((rc.Y8.astype(np.float32) - tau * tb) / s2).sum(dtype=np.float32) / prec
        eps = np.asarray(normal(rng_key), np.float32)
        return {"mu": wrap(np.float32(mean + eps / np.sqrt(prec)))}

    return gibbs_fn


def prior_sample_mu(normal1):
    """``init_to_sample`` on ``Normal(0, 5)``: ``normal1(key)`` = ``normal(key, (1,))``."""
    return np.float32(0.0) + np.float32(5.0) * np.asarray(normal1, np.float32)[0]
