"""Several chains on one stream of X (csrc/glm_mc_kernel.cuh): 2..16 chains of the tall-data
logistic model share each pass over the design matrix, four per launch (the reference's
chain_method="vectorized" for this model, numpyro/infer/hmc.py:790-798).

Checked: tier 2 against the float64 oracle (all four chain slots, ragged shapes, every column
count class), tier 3 against the oracle's sampler chain by chain, bit-identity of a chain run
alone through the same kernel with the same chain inside a group of four (and of five: two
launches), and the covtype-shape potential.
"""
import os
import sys

import numpy as np
from numpy.testing import assert_allclose
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import hmc as ohmc  # noqa: E402
from oracle import potentials as opot  # noqa: E402
from oracle import threefry as T  # noqa: E402
from oracle.hmc_util import Config  # noqa: E402


@pytest.fixture(scope="module")
def nb():
    import __graft_entry__ as g

    return g.build()


def make_logistic(N, D, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(N, D)).astype(np.float32)
    if D > 1:
        X[:, -1] = 1.0
    y = rng.random(N) < 1 / (1 + np.exp(-X @ (rng.normal(size=D) * 0.5)))
    return X, y


def _layout(nb, kern):
    import ctypes as C

    from numpyro_b200 import _lib

    out = (C.c_int * 8)()
    _lib.check(_lib.lib().b200mc_glm_layout(kern._handle, out))
    return list(out)


@pytest.mark.parametrize("N,D", [(1, 1), (31, 3), (33, 4), (1000, 8), (3000, 13), (4133, 55),
                                 (70001, 55), (5000, 64), (2048, 17), (999, 29), (640, 48)])
def test_multichain_potential_and_grad(nb, N, D):
    X, y = make_logistic(N, D, seed=N + D)
    k = nb.infer.NUTS(nb.models.logistic_regression)
    k.init(T.split(T.PRNGKey(0), 4, True), 0, {"coefs": np.zeros((4, D), np.float32)}, (X, y), {})
    assert _layout(nb, k)[6] == 4  # the multi-chain kernel serves this handle
    fam = opot.Logistic(X, y, np.float64)
    rng = np.random.default_rng(1)
    zs = (rng.normal(size=(7, D)) * np.array([0.1, 1.0, 3.0, 0.5, 0.01, 2.0, 1.0])[:, None]).astype(np.float32)
    U, g = k.potential_and_grad(zs)  # 4 + 3 points: a full and a partial launch
    for i in range(7):
        U0, g0 = fam.potential_and_grad(zs[i].astype(np.float64))
        assert_allclose(U[i], U0, rtol=1e-5, err_msg=f"point {i}")
        assert_allclose(g[i], g0, rtol=1e-5, atol=1e-5 * np.abs(g0).max() + 1e-6, err_msg=f"point {i}")


def _run(nb, X, y, key, chains, n, nw=0, **kw):
    mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.logistic_regression, **kw), num_warmup=nw, num_samples=n,
                       num_chains=chains, chain_method="vectorized")
    mc.run(key, X, y, extra_fields=("num_steps", "potential_energy", "accept_prob"))
    return mc.get_samples(group_by_chain=True)["coefs"], mc.get_extra_fields(group_by_chain=True), mc


@pytest.mark.parametrize("part", [True, False])
def test_tier3_multichain_matches_oracle(nb, part):
    X, y = make_logistic(3000, 5, seed=5)
    fam = opot.Logistic(X, y)
    kw = dict(step_size=0.01, adapt_step_size=False, adapt_mass_matrix=False)
    chains, n = 4, 30
    key = T.PRNGKey(11)
    kern = nb.infer.NUTS(nb.models.logistic_regression, **kw)
    kern.partitionable = part
    mc = nb.infer.MCMC(kern, num_warmup=0, num_samples=n, num_chains=chains, chain_method="vectorized")
    mc.run(key, X, y, extra_fields=("num_steps", "potential_energy", "accept_prob"))
    ex = mc.get_extra_fields(group_by_chain=True)
    res, states, _ = ohmc.run_mcmc(fam, key, 0, n, num_chains=chains, cfg=Config(np.float32, part), **kw)
    assert np.array_equal(ex["num_steps"], res["num_steps"])
    assert np.array_equal(ex["diverging"], res["diverging"])
    assert ex["num_steps"].max() > 3
    assert_allclose(ex["accept_prob"], res["accept_prob"], rtol=2e-3, atol=2e-4)
    assert_allclose(ex["potential_energy"], res["potential_energy"], rtol=1e-4, atol=1e-3)
    assert_allclose(mc.get_samples(group_by_chain=True)["coefs"], res["z"], rtol=5e-3, atol=5e-3)


def test_chain_alone_equals_chain_in_a_group(nb):
    """Per-chain arithmetic does not depend on its neighbours: chain c of a 5-chain run (launches
    of 4 + 1) is bit-identical to the 1-chain run with split(key, 5)[c] through the same kernel
    (B200MC_GLM_MC=1 forces the multi-chain kernel for a single chain), warm-up included."""
    X, y = make_logistic(20000, 55, seed=6)
    key = T.PRNGKey(3)
    z5, e5, _ = _run(nb, X, y, key, 5, 25, nw=40)
    keys = nb.random.split(key, 5)
    os.environ["B200MC_GLM_MC"] = "1"
    try:
        for c in (0, 2, 3, 4):
            z1, e1, mc1 = _run(nb, X, y, keys[c], 1, 25, nw=40)
            assert _layout(nb, mc1.sampler)[6] == 4
            assert np.array_equal(e1["num_steps"][0], e5["num_steps"][c]), c
            assert np.array_equal(z1[0], z5[c]), c
    finally:
        os.environ.pop("B200MC_GLM_MC", None)


def test_multichain_agrees_with_single_chain_kernel(nb):
    """Same chains through glm_kernel (one launch per chain, exact expf / log1pf) and through the
    multi-chain kernel (MUFU-only Bernoulli term, rows split over four lanes): identical tree
    sizes over a fixed-parameter run, positions to f32 round-off growth."""
    X, y = make_logistic(6000, 21, seed=8)
    kw = dict(step_size=0.02, adapt_step_size=False, adapt_mass_matrix=False)
    key = T.PRNGKey(5)
    zm, em, _ = _run(nb, X, y, key, 3, 20, **kw)
    os.environ["B200MC_GLM_MC"] = "0"
    try:
        zs, es, mcs = _run(nb, X, y, key, 3, 20, **kw)
        assert _layout(nb, mcs.sampler)[6] == 1
    finally:
        os.environ.pop("B200MC_GLM_MC", None)
    assert np.array_equal(em["num_steps"], es["num_steps"])
    assert_allclose(zm, zs, rtol=2e-3, atol=2e-4)
