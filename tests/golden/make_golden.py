"""Generate the committed golden fixtures under tests/golden/.

The reference (NumPyro 0.18.0) cannot be imported in this image (hard dependency on
jax/jaxlib, absent and not installable), so these vectors come from the CPU oracle
(``oracle/``), which is itself pinned to the golden vectors held by the reference's own
tests (tests/test_oracle_golden.py).  They are "parity unpinned" with respect to a live
reference run and serve as (a) a regression pin of the oracle and (b) fixed inputs/outputs
for the CUDA parity tests, which must not read /root/reference at run time.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import hmc as ohmc  # noqa: E402
from oracle import potentials as opot  # noqa: E402
from oracle import threefry as T  # noqa: E402
from oracle.hmc_util import Config  # noqa: E402

Y8 = np.array([28.0, 8.0, -3.0, 7.0, -1.0, 1.0, 18.0, 12.0])
S8 = np.array([15.0, 10.0, 16.0, 11.0, 9.0, 11.0, 10.0, 18.0])
SEEDS = (0, 1, 42, 2**31 + 5, 2**40 + 3)


def small_logistic(N=257, D=7, seed=3):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(N, D)).astype(np.float32)
    X[:, -1] = 1.0
    w = rng.normal(size=D) * 0.5
    y = (rng.random(N) < 1 / (1 + np.exp(-X @ w))).astype(np.uint8)
    return X, y


def small_sv(Tn=30, seed=21):
    rng = np.random.default_rng(seed)
    s = np.cumsum(rng.normal(size=Tn) * 0.02) - 4.5
    return (np.exp(s) * rng.standard_t(10, size=Tn)).astype(np.float32)


def threefry_fixture():
    out = {"seeds": np.array(SEEDS, np.uint64)}
    for part in (False, True):
        tag = "part" if part else "legacy"
        for i, seed in enumerate(SEEDS):
            key = T.PRNGKey(seed)
            out[f"{tag}_{i}_key"] = key
            out[f"{tag}_{i}_split2"] = T.split(key, 2, part)
            out[f"{tag}_{i}_split3"] = T.split(key, 3, part)
            out[f"{tag}_{i}_split7"] = T.split(key, 7, part)
            out[f"{tag}_{i}_bits10"] = T.random_bits(key, (10,), part)
            out[f"{tag}_{i}_uniform55"] = T.uniform(key, (55,), -2.0, 2.0, part)
            out[f"{tag}_{i}_normal55"] = T.normal(key, (55,), part)
    return out


def potential_fixture():
    X, y = small_logistic()
    r = small_sv()
    rng = np.random.default_rng(7)
    out = {"logistic_X": X, "logistic_y": y, "sv_returns": r}
    fams = {"logistic": opot.Logistic(X, y, np.float64),
            "eight_schools": opot.EightSchools(S8, Y8, np.float64),
            "sv": opot.StochasticVolatility(r, np.float64)}
    for name, fam in fams.items():
        zs = (rng.normal(size=(8, fam.dim)) * 0.7).astype(np.float32)
        if name == "sv":
            zs[:, 1:-1] = zs[:, 1:-1] * 0.1 - 4.5
            zs[:, -1] -= 3.5
        U = np.zeros(8)
        g = np.zeros((8, fam.dim))
        for i in range(8):
            U[i], g[i] = fam.potential_and_grad(zs[i].astype(np.float64))
        out[f"{name}_z"], out[f"{name}_U"], out[f"{name}_grad"] = zs, U, g
    return out


def trace_fixture():
    X, y = small_logistic()
    out = {}
    cases = {"eight_schools": (opot.EightSchools(S8, Y8), 0.3, 4),
             "logistic": (opot.Logistic(X, y), 0.05, 1)}
    for name, (fam, eps, chains) in cases.items():
        for part in (False, True):
            res, states, _ = ohmc.run_mcmc(fam, T.PRNGKey(11), 0, 50, num_chains=chains,
                                           cfg=Config(np.float32, part), step_size=eps,
                                           adapt_step_size=False, adapt_mass_matrix=False)
            tag = f"{name}_{'part' if part else 'legacy'}"
            for f in ("num_steps", "diverging", "accept_prob", "potential_energy"):
                out[f"{tag}_{f}"] = res[f]
            out[f"{tag}_z_last"] = np.stack([s.z for s in states])
        out[f"{name}_step_size"] = np.float32(eps)
    return out


if __name__ == "__main__":
    np.savez_compressed(os.path.join(HERE, "threefry.npz"), **threefry_fixture())
    np.savez_compressed(os.path.join(HERE, "potentials.npz"), **potential_fixture())
    np.savez_compressed(os.path.join(HERE, "traces.npz"), **trace_fixture())
    print("wrote", [f for f in os.listdir(HERE) if f.endswith(".npz")])
