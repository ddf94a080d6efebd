"""Case table shared by ``make_refpin.py`` (runs the reference's own hmc.py/hmc_util.py over
``refshim``) and the tests that replay the cases with the oracle and the CUDA path.

TEST INFRASTRUCTURE.  A case = (family, kernel class, constructor kwargs exactly as NumPyro's
``HMC``/``NUTS`` take them, seed, num_warmup, num_samples, threefry layout).
"""
import numpy as np

Y8 = np.array([28.0, 8.0, -3.0, 7.0, -1.0, 1.0, 18.0, 12.0])
S8 = np.array([15.0, 10.0, 16.0, 11.0, 9.0, 11.0, 10.0, 18.0])


def make_data():
    """Small synthetic data sets (also stored inside refpin.npz, which the tests read)."""
    rng = np.random.default_rng(3)
    X = rng.normal(size=(257, 7)).astype(np.float32)
    X[:, -1] = 1.0
    w = rng.normal(size=7) * 0.5
    y = (rng.random(257) < 1 / (1 + np.exp(-X @ w))).astype(np.uint8)
    rng = np.random.default_rng(5)
    Xh = rng.normal(size=(48, 6)).astype(np.float32)
    mu = Xh[:, :3] @ np.array([2.0, -1.0, 0.5])
    Yn = (mu - mu.mean() + 0.05 * rng.normal(size=48)).astype(np.float32)
    Yb = (rng.random(48) < 1 / (1 + np.exp(-mu))).astype(np.float32)
    rng = np.random.default_rng(21)
    s = np.cumsum(rng.normal(size=30) * 0.02) - 4.5
    r = (np.exp(s) * rng.standard_t(10, size=30)).astype(np.float32)
    return {"logistic_X": X, "logistic_y": y, "hs_X": Xh, "hs_Yn": Yn, "hs_Yb": Yb,
            "sv_returns": r}


def families(data):
    from oracle import potentials as opot

    return {"es": opot.EightSchools(S8, Y8),
            "logit": opot.Logistic(data["logistic_X"], data["logistic_y"]),
            "hsb": opot.Horseshoe(data["hs_X"], data["hs_Yb"], "bernoulli"),
            "hsn": opot.Horseshoe(data["hs_X"], data["hs_Yn"], "normal"),
            "sv": opot.StochasticVolatility(data["sv_returns"])}


def init_point(fam, seed):
    """Deterministic unconstrained start (the reference's init search is not part of the
    pinned code, see refshim.py)."""
    rng = np.random.default_rng(1000 + seed)
    z = (rng.uniform(-1, 1, size=fam.dim)).astype(np.float32)
    if fam.name == "stochastic_volatility" or fam.__class__.__name__ == "StochasticVolatility":
        off = fam.offsets()
        z[off["s"][0]:off["s"][0] + off["s"][1]] = z[off["s"][0]:off["s"][0] + off["s"][1]] * 0.1 - 4.5
        z[off["sigma"][0]] = -3.5
        z[off["nu"][0]] = 2.0
    return z


_IMM10 = (0.5 + np.arange(10) * 0.25).astype(np.float32)
_A = np.random.default_rng(9).normal(size=(10, 10))
_IMM10D = (_A @ _A.T / 10 + np.eye(10)).astype(np.float32)

# name: (family, algo, kwargs, seed, num_warmup, num_samples, partitionable)
CASES = {
    "es_fixed_part": ("es", "NUTS", dict(step_size=0.3, adapt_step_size=False, adapt_mass_matrix=False), 11, 0, 50, True),
    "es_fixed_legacy": ("es", "NUTS", dict(step_size=0.3, adapt_step_size=False, adapt_mass_matrix=False), 11, 0, 50, False),
    "es_adapt_diag": ("es", "NUTS", dict(), 1, 150, 30, True),
    "es_adapt_diag_legacy": ("es", "NUTS", dict(), 2, 150, 30, False),
    "es_adapt_dense": ("es", "NUTS", dict(dense_mass=True), 3, 150, 30, True),
    "es_adapt_dense_noreg": ("es", "NUTS", dict(dense_mass=True, regularize_mass_matrix=False), 4, 150, 20, True),
    "es_short_warmup": ("es", "NUTS", dict(), 5, 15, 20, True),
    "es_target09": ("es", "NUTS", dict(target_accept_prob=0.9, step_size=0.5), 6, 120, 20, True),
    "es_depth_tuple": ("es", "NUTS", dict(max_tree_depth=(3, 6), step_size=0.05, adapt_step_size=False), 7, 40, 30, True),
    "es_frss": ("es", "NUTS", dict(find_heuristic_step_size=True), 8, 120, 20, True),
    "es_imm_diag": ("es", "NUTS", dict(inverse_mass_matrix=_IMM10, adapt_mass_matrix=False), 9, 60, 20, True),
    "es_imm_dense": ("es", "NUTS", dict(inverse_mass_matrix=_IMM10D, dense_mass=True, adapt_mass_matrix=False), 10, 60, 20, True),
    "es_hmc_traj": ("es", "HMC", dict(), 12, 100, 20, True),
    "es_hmc_numsteps": ("es", "HMC", dict(num_steps=7, trajectory_length=None, step_size=0.2), 13, 60, 20, True),
    "es_hmc_fixed": ("es", "HMC", dict(step_size=0.2, trajectory_length=1.5, adapt_step_size=False, adapt_mass_matrix=False), 14, 0, 40, True),
    "logit_adapt": ("logit", "NUTS", dict(), 15, 100, 30, True),
    "logit_fixed": ("logit", "NUTS", dict(step_size=0.05, adapt_step_size=False, adapt_mass_matrix=False), 16, 0, 50, True),
    "hsb_dense": ("hsb", "NUTS", dict(dense_mass=True, max_tree_depth=6), 17, 120, 20, True),
    "hsn_diag": ("hsn", "NUTS", dict(max_tree_depth=6), 18, 100, 20, True),
    "sv_adapt": ("sv", "NUTS", dict(max_tree_depth=7), 19, 80, 20, True),
    # structured mass matrices (hmc_util.py:439-486, test_mcmc.py:988-1050)
    "es_block_mu_tau": ("es", "NUTS", dict(dense_mass=[("mu", "tau")]), 20, 150, 30, True),
    "es_block_reversed": ("es", "NUTS", dict(dense_mass=[("theta_base", "mu")]), 21, 150, 30, True),
    "es_block_two": ("es", "NUTS", dict(dense_mass=[("tau",), ("theta_base", "mu")]), 22, 150, 30, True),
    "es_block_two_legacy": ("es", "NUTS", dict(dense_mass=[("tau",), ("theta_base", "mu")]), 23, 150, 30, False),
    "es_block_imm_dict": ("es", "NUTS", dict(inverse_mass_matrix={("mu",): np.array([2.0], np.float32), ("theta_base", "tau"): (0.5 + 0.125 * np.arange(9)).astype(np.float32)}, adapt_mass_matrix=False), 24, 60, 20, True),
    "es_block_imm_dict_dense": ("es", "NUTS", dict(inverse_mass_matrix={("tau", "mu"): np.array([[2.0, 0.3], [0.3, 1.0]], np.float32)}, dense_mass=[("tau", "mu")]), 25, 150, 20, True),
    "es_block_hmc": ("es", "HMC", dict(dense_mass=[("mu", "tau")], trajectory_length=2.0), 26, 100, 20, True),
    "hsn_block": ("hsn", "NUTS", dict(dense_mass=[("tau", "prec_obs"), ("unscaled_betas",)], max_tree_depth=6), 27, 120, 20, True),
}

FIELDS = ("z", "potential_energy", "energy", "num_steps", "accept_prob", "mean_accept_prob",
          "diverging", "step_size")


def block_tag(names):
    return "+".join(names)


# fori_collect (numpyro/util.py:277-407): (lower, upper, thinning, collection_size)
COLLECT_CASES = [(0, 10, 1, None), (5, 20, 3, None), (7, 20, 4, None), (0, 9, 2, None), (3, 3, 1, None),
                 (10, 37, 5, 8), (0, 1, 1, None), (4, 30, 7, None)]


def collect_tag(c):
    return "collect/" + "_".join("x" if v is None else str(v) for v in c)
