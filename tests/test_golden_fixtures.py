"""Committed golden fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py).

CPU: the oracle still reproduces them bit for bit (regression pin of the checker).
GPU: the CUDA path, through the C ABI, against the same fixtures -- no oracle import needed
at run time and nothing read from /root/reference.
"""
import os
import sys

import numpy as np
from numpy.testing import assert_allclose
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden as mg  # noqa: E402

G = {n: np.load(os.path.join(HERE, "golden", n + ".npz")) for n in ("threefry", "potentials",
                                                                     "traces")}


def test_oracle_reproduces_threefry_fixture():
    new = mg.threefry_fixture()
    for k in G["threefry"].files:
        assert np.array_equal(new[k], G["threefry"][k]), k


def test_oracle_reproduces_potential_fixture():
    new = mg.potential_fixture()
    for k in G["potentials"].files:
        assert_allclose(new[k], G["potentials"][k], rtol=1e-12, atol=0, err_msg=k)


def test_oracle_reproduces_trace_fixture():
    new = mg.trace_fixture()
    for k in G["traces"].files:
        if k.endswith(("num_steps", "diverging")):
            assert np.array_equal(new[k], G["traces"][k]), k


# ------------------------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def nb():
    import __graft_entry__ as g

    return g.build()


@pytest.mark.gpu
@pytest.mark.parametrize("part", [False, True])
def test_cuda_threefry_against_fixture(nb, part):
    R, F = nb.random, G["threefry"]
    tag = "part" if part else "legacy"
    for i, seed in enumerate(F["seeds"]):
        key = R.PRNGKey(int(seed))
        assert np.array_equal(key, F[f"{tag}_{i}_key"])
        for n in (2, 3, 7):
            assert np.array_equal(R.split(key, n, part), F[f"{tag}_{i}_split{n}"])
        assert np.array_equal(R.bits(key, (10,), part), F[f"{tag}_{i}_bits10"])
        assert np.array_equal(R.uniform(key, (55,), -2.0, 2.0, part), F[f"{tag}_{i}_uniform55"])
        assert_allclose(R.normal(key, (55,), part), F[f"{tag}_{i}_normal55"], rtol=3e-7, atol=1e-7)


@pytest.mark.gpu
def test_cuda_potentials_against_fixture(nb):
    F = G["potentials"]
    cases = [("logistic", nb.models.logistic_regression, (F["logistic_X"], F["logistic_y"]), {}),
             ("eight_schools", nb.models.eight_schools_noncentered, (8, mg.S8), {"y": mg.Y8}),
             ("sv", nb.models.stochastic_volatility, (F["sv_returns"],), {})]
    for name, model, args, kwargs in cases:
        k = nb.infer.NUTS(model)
        k.init(nb.random.PRNGKey(0), 0, None, args, kwargs)
        U, g = k.potential_and_grad(F[f"{name}_z"])
        assert_allclose(U, F[f"{name}_U"], rtol=1e-5, err_msg=name)
        g0 = F[f"{name}_grad"]
        for i in range(g0.shape[0]):
            assert_allclose(g[i], g0[i], rtol=2e-5, atol=2e-5 * np.abs(g0[i]).max(), err_msg=name)


@pytest.mark.gpu
@pytest.mark.parametrize("part", [False, True])
def test_cuda_traces_against_fixture(nb, part):
    F = G["traces"]
    P = G["potentials"]
    tag = "part" if part else "legacy"
    cases = [("eight_schools", nb.models.eight_schools_noncentered, (8, mg.S8), {"y": mg.Y8}, 4),
             ("logistic", nb.models.logistic_regression, (P["logistic_X"], P["logistic_y"]), {}, 1)]
    for name, model, args, kwargs, chains in cases:
        kern = nb.infer.NUTS(model, step_size=float(F[f"{name}_step_size"]),
                             adapt_step_size=False, adapt_mass_matrix=False)
        kern.partitionable = part
        mc = nb.infer.MCMC(kern, num_warmup=0, num_samples=50, num_chains=chains,
                           chain_method="vectorized")
        mc.run(nb.random.PRNGKey(11), *args, extra_fields=("num_steps", "accept_prob",
                                                            "potential_energy"), **kwargs)
        ex = mc.get_extra_fields(group_by_chain=True)
        assert np.array_equal(ex["num_steps"], F[f"{name}_{tag}_num_steps"]), name
        assert np.array_equal(ex["diverging"], F[f"{name}_{tag}_diverging"]), name
        assert_allclose(ex["accept_prob"], F[f"{name}_{tag}_accept_prob"], rtol=2e-3, atol=2e-4)
        assert_allclose(ex["potential_energy"], F[f"{name}_{tag}_potential_energy"], rtol=1e-4,
                        atol=1e-3)
