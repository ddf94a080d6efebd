"""GPU parity of the many-chain lock-step engine (csrc/batch_kernel.cuh): from 64 chains on
the horseshoe potential runs as two split-precision tensor-core contractions over all chains
(csrc/gemm3x_kernel.cuh) and every chain advances by one gradient per global step.

Checked against (a) the float64 oracle (tier 2, potential and gradient), (b) the oracle's
sampler chain by chain (tier 3: identical tree sizes with fixed step size and mass matrix),
(c) the CTA-per-chain CUDA-core path of the same library on the same keys, and (d) the known
coefficients of the synthetic regression after a dense-mass warm-up (tier 4).
"""
import os

import numpy as np
from numpy.testing import assert_allclose
import pytest

pytestmark = pytest.mark.gpu

from oracle import hmc as ohmc  # noqa: E402
from oracle import potentials as opot  # noqa: E402
from oracle import threefry as T  # noqa: E402
from oracle.hmc_util import Config  # noqa: E402


@pytest.fixture(scope="module")
def nb():
    import __graft_entry__ as g

    return g.build()


def make_horseshoe(N, D, likelihood, seed=0):
    # examples/horseshoe_regression.py:105-125
    rng = np.random.RandomState(seed)
    X = rng.randn(N, D).astype(np.float32)
    Y = X[:, :3] @ np.array([2.0, -1.0, 0.50])
    Y = Y - Y.mean()
    if likelihood == "normal":
        Y = Y + 0.05 * rng.randn(N)
    else:
        Y = rng.binomial(1, 1 / (1 + np.exp(-Y))).astype(np.float64)
    return X, Y.astype(np.float32)


def model_of(nb, likelihood):
    return nb.models.horseshoe_bernoulli if likelihood == "bernoulli" else nb.models.horseshoe_normal


@pytest.mark.parametrize("likelihood", ["bernoulli", "normal"])
@pytest.mark.parametrize("N,D,n_pts", [(300, 40, 64), (1000, 130, 256), (130, 7, 300)])
def test_batched_potential_and_grad(nb, likelihood, N, D, n_pts):
    X, Y = make_horseshoe(N, D, likelihood)
    k = nb.infer.NUTS(model_of(nb, likelihood))
    k.init(T.split(T.PRNGKey(0), 64, True), 0, None, (X, Y), {})  # 64 chains: tensor-core engine
    fam = opot.Horseshoe(X, Y, likelihood, np.float64)
    rng = np.random.default_rng(1)
    zs = (rng.normal(size=(n_pts, fam.dim)) * 0.5).astype(np.float32)
    U, g = k.potential_and_grad(zs)
    for i in range(0, n_pts, 7):
        U0, g0 = fam.potential_and_grad(zs[i].astype(np.float64))
        assert_allclose(U[i], U0, rtol=2e-5)
        assert_allclose(g[i], g0, rtol=1e-4, atol=1e-4 * np.abs(g0).max())


@pytest.mark.parametrize("likelihood", ["bernoulli", "normal"])
def test_tier3_batched_chains_match_oracle(nb, likelihood):
    X, Y = make_horseshoe(100, 5, likelihood)
    fam = opot.Horseshoe(X, Y, likelihood)
    kw = dict(step_size=0.02, adapt_step_size=False, adapt_mass_matrix=False, max_tree_depth=6)
    chains, n = 64, 12
    mc = nb.infer.MCMC(nb.infer.NUTS(model_of(nb, likelihood), **kw), num_warmup=0, num_samples=n,
                       num_chains=chains, chain_method="vectorized")
    key = T.PRNGKey(11)
    mc.run(key, X, Y, extra_fields=("num_steps", "potential_energy"))
    ex = mc.get_extra_fields(group_by_chain=True)
    keys = T.split(key, chains, True)
    for c in (0, 1, 17, 63):
        ok = ohmc.HMCKernel(fam, cfg=Config(np.float32, True), **kw)
        r, _ = ohmc.run_chain(ok, keys[c], 0, n)
        assert np.array_equal(ex["num_steps"][c], r["num_steps"]), c
        assert_allclose(ex["potential_energy"][c], r["potential_energy"], rtol=5e-4, atol=2e-3)


def test_batched_engine_matches_cta_per_chain_path(nb):
    X, Y = make_horseshoe(200, 12, "normal")
    kw = dict(step_size=0.01, adapt_step_size=False, adapt_mass_matrix=False, max_tree_depth=5)
    out = {}
    for name, env in (("batch", None), ("cta", "100000")):
        if env is None:
            os.environ.pop("B200MC_BATCH_MIN_CHAINS", None)
        else:
            os.environ["B200MC_BATCH_MIN_CHAINS"] = env
        try:
            mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.horseshoe_normal, **kw), num_warmup=0,
                               num_samples=8, num_chains=64, chain_method="vectorized")
            mc.run(T.PRNGKey(5), X, Y, extra_fields=("num_steps",))
            out[name] = (mc.get_extra_fields(group_by_chain=True)["num_steps"],
                         mc.get_samples(group_by_chain=True)["betas"])
        finally:
            os.environ.pop("B200MC_BATCH_MIN_CHAINS", None)
    assert np.array_equal(out["batch"][0], out["cta"][0])
    assert_allclose(out["batch"][1], out["cta"][1], rtol=2e-3, atol=2e-3)


def test_batched_dense_warmup_recovers_coefficients(nb):
    X, Y = make_horseshoe(150, 6, "normal")
    mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.horseshoe_normal, dense_mass=True), num_warmup=150,
                       num_samples=60, num_chains=64, chain_method="vectorized")
    mc.run(T.PRNGKey(1), X, Y, extra_fields=("num_steps",))
    b = mc.get_samples()["betas"].mean(0)
    assert_allclose(b[:3], [2.0, -1.0, 0.5], atol=0.15)
    assert int(np.asarray(mc.last_state.i).min()) == 210
    # the factors the window-end factorisation left behind belong to the adapted covariance:
    # M^{1/2} M^{1/2}^T = (M^-1)^-1, M^{-1/2} = (M^{1/2})^-1, both lower triangular
    ad = mc.last_state.adapt_state
    imm = next(iter(ad.inverse_mass_matrix.values()))
    msq = next(iter(ad.mass_matrix_sqrt.values()))
    msqi = next(iter(ad.mass_matrix_sqrt_inv.values()))
    D = imm.shape[-1]
    assert imm.shape == (64, D, D)
    for c in (0, 31, 63):
        assert not np.allclose(imm[c], np.eye(D))
        assert_allclose(msq[c] @ msq[c].T @ imm[c], np.eye(D), atol=5e-3)
        assert_allclose(msqi[c] @ msq[c], np.eye(D), atol=2e-3)
        assert np.allclose(np.triu(msq[c], 1), 0) and np.allclose(np.triu(msqi[c], 1), 0)
        assert int(np.asarray(ad.window_idx)[c]) > 1


# ---------------------------------------------------------------- many-chain logistic regression
def make_logistic(N, D, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(N, D)).astype(np.float32)
    X[:, -1] = 1.0
    y = rng.random(N) < 1 / (1 + np.exp(-X @ (rng.normal(size=D) * 0.5)))
    return X, y


@pytest.mark.parametrize("N,D", [(1000, 5), (5000, 55), (333, 70)])
def test_batched_logistic_potential_and_grad(nb, N, D):
    X, y = make_logistic(N, D, seed=N)
    k = nb.infer.NUTS(nb.models.logistic_regression)
    k.init(T.split(T.PRNGKey(0), 64, True), 0, None, (X, y), {})
    fam = opot.Logistic(X, y, np.float64)
    zs = (np.random.default_rng(2).normal(size=(100, D)) * 0.7).astype(np.float32)
    U, g = k.potential_and_grad(zs)
    for i in range(0, 100, 9):
        U0, g0 = fam.potential_and_grad(zs[i].astype(np.float64))
        assert_allclose(U[i], U0, rtol=1e-5)
        assert_allclose(g[i], g0, rtol=1e-5, atol=1e-5 * np.abs(g0).max())


def test_tier3_batched_logistic_matches_oracle_and_streaming_kernel(nb):
    X, y = make_logistic(3000, 5, seed=5)
    fam = opot.Logistic(X, y)
    kw = dict(step_size=0.01, adapt_step_size=False, adapt_mass_matrix=False)
    chains, n = 64, 10
    key = T.PRNGKey(7)
    out = {}
    for name, env in (("batch", None), ("stream", "100000")):
        if env is not None:
            os.environ["B200MC_BATCH_MIN_CHAINS"] = env
        try:
            mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.logistic_regression, **kw), num_warmup=0,
                               num_samples=n, num_chains=chains, chain_method="vectorized")
            mc.run(key, X, y, extra_fields=("num_steps", "potential_energy"))
            out[name] = (mc.get_extra_fields(group_by_chain=True), mc.get_samples(group_by_chain=True))
        finally:
            os.environ.pop("B200MC_BATCH_MIN_CHAINS", None)
    assert np.array_equal(out["batch"][0]["num_steps"], out["stream"][0]["num_steps"])
    assert_allclose(out["batch"][1]["coefs"], out["stream"][1]["coefs"], rtol=1e-3, atol=1e-4)
    keys = T.split(key, chains, True)
    for c in (0, 40):
        ok = ohmc.HMCKernel(fam, cfg=Config(np.float32, True), **kw)
        r, _ = ohmc.run_chain(ok, keys[c], 0, n)
        assert np.array_equal(out["batch"][0]["num_steps"][c], r["num_steps"]), c
        assert_allclose(out["batch"][0]["potential_energy"][c], r["potential_energy"], rtol=1e-4)


def test_more_chains_than_operand_columns_run_as_column_groups(nb):
    """300 chains = a group of 256 operand columns + a group of 44: every chain keeps its global
    index (keys, state, collected draws) and matches the oracle's chain on its own key."""
    X, y = make_logistic(2000, 5, seed=6)
    fam = opot.Logistic(X, y)
    kw = dict(step_size=0.02, adapt_step_size=False, adapt_mass_matrix=False)
    chains, n = 300, 8
    key = T.PRNGKey(11)
    mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.logistic_regression, **kw), num_warmup=0,
                       num_samples=n, num_chains=chains, chain_method="vectorized")
    mc.run(key, X, y, extra_fields=("num_steps", "potential_energy"))
    ex = mc.get_extra_fields(group_by_chain=True)
    z = mc.get_samples(group_by_chain=True)["coefs"]
    assert z.shape == (chains, n, 5) and np.isfinite(z).all()
    keys = T.split(key, chains, True)
    for c in (0, 255, 256, 299):
        ok = ohmc.HMCKernel(fam, cfg=Config(np.float32, True), **kw)
        r, _ = ohmc.run_chain(ok, keys[c], 0, n)
        assert np.array_equal(ex["num_steps"][c], r["num_steps"]), c
        assert_allclose(ex["potential_energy"][c], r["potential_energy"], rtol=1e-4)
        assert_allclose(z[c], r["z"], rtol=2e-3, atol=2e-4)
    # adaptation on, all chains: the second group adapts like the first
    mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.logistic_regression), num_warmup=100, num_samples=50,
                       num_chains=chains, chain_method="vectorized")
    mc.run(T.PRNGKey(12), X, y)
    s = np.asarray(mc.get_samples(group_by_chain=True)["coefs"], np.float64)
    assert nb.diagnostics.split_gelman_rubin(s).max() < 1.02
    m = s.mean(1)
    assert np.abs(m[256:].mean(0) - m[:256].mean(0)).max() < 0.02


def test_batched_logistic_posterior(nb):
    X, y = make_logistic(4000, 6, seed=9)
    mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.logistic_regression), num_warmup=200, num_samples=100,
                       num_chains=64, chain_method="vectorized")
    mc.run(T.PRNGKey(3), X, y)
    s = np.asarray(mc.get_samples(group_by_chain=True)["coefs"], np.float64)
    assert nb.diagnostics.split_gelman_rubin(s).max() < 1.02
    # posterior mean close to the maximum-likelihood-ish truth scale: compare with a long CPU-free
    # check: the gradient of the potential at the posterior mean is small relative to its scale
    fam = opot.Logistic(X, y, np.float64)
    U0, g0 = fam.potential_and_grad(s.reshape(-1, 6).mean(0))
    assert np.abs(g0).max() < 0.05 * 4000 ** 0.5
