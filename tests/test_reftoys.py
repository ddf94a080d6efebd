"""The Beta-Bernoulli model of the reference's own test-suite (``test/infer/test_mcmc.py:231-275``;
SURVEY 8f-2): oracle and CUDA potential against the REFERENCE's ``potential_energy`` /
``initialize_model`` / ``postprocess_fn``, executed from /root/reference over the NumPy stand-in
(``tests/golden/make_reftoys.py``), and the reference test's own assertion on the posterior."""
import os

import numpy as np
import pytest

from oracle import hmc as ohmc
from oracle.hmc_util import Config
from oracle.potentials import BetaBernoulli
from oracle.threefry import PRNGKey


@pytest.fixture(scope="module")
def fx():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "reftoys.npz"))


def test_oracle_matches_reference_beta_bernoulli(fx):
    fam = BetaBernoulli(fx["data"])
    U = np.array([fam.potential_and_grad(z)[0] for z in fx["z"]])
    np.testing.assert_allclose(U, fx["U"], rtol=5e-7)  # counts x log p vs 2000 summed log-probs
    np.testing.assert_allclose(fam.constrain(fx["z"][0])["p_latent"], fx["constrained/p_latent"], rtol=1e-6)
    f64 = BetaBernoulli(fx["data"], dtype=np.float64)
    for z in fx["z"]:
        z = z.astype(np.float64)
        g = f64.potential_and_grad(z)[1]
        gn = np.array([(f64.potential_and_grad(z + 1e-6 * e)[0] - f64.potential_and_grad(z - 1e-6 * e)[0]) / 2e-6
                       for e in np.eye(2)])
        np.testing.assert_allclose(g, gn, rtol=1e-6, atol=1e-5)
    for part in (True, False):
        for seed in (0, 1, 2):
            (z0, _, _), _ = ohmc.find_valid_initial_params(Config(np.float32, part), PRNGKey(seed), fam)
            assert np.array_equal(z0, fx[f"init_{'part' if part else 'legacy'}_{seed}"])


@pytest.mark.gpu
def test_cuda_beta_bernoulli_potential_and_init(fx):
    from numpyro_b200 import models
    from numpyro_b200.infer import NUTS

    data = fx["data"]
    fam, f64 = BetaBernoulli(data), BetaBernoulli(data, dtype=np.float64)
    k = NUTS(models.beta_bernoulli)
    for part in (True, False):
        k.partitionable = part
        for seed in (0, 1, 2):
            chain_key = PRNGKey(100 + seed)
            cfg = Config(np.float32, part)
            (z_or, _, _), _ = ohmc.find_valid_initial_params(cfg, cfg.split(chain_key)[1], fam)
            st = k.init(np.asarray(chain_key, np.uint32), 10, model_args=(data,))
            assert np.array_equal(st.z["p_latent"], z_or)
    U, g = k.potential_and_grad(fx["z"])
    np.testing.assert_allclose(U, fx["U"], rtol=2e-6)
    for i, z in enumerate(fx["z"]):
        np.testing.assert_allclose(g[i], f64.potential_and_grad(z.astype(np.float64))[1], rtol=2e-5, atol=1e-3)
    p = k.postprocess_fn((data,), {})({"p_latent": fx["z"][0]})["p_latent"]
    np.testing.assert_allclose(p, fx["constrained/p_latent"], rtol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("algo", ["HMC", "NUTS"])
def test_cuda_beta_bernoulli_posterior(fx, algo):
    """test_mcmc.py:263-275: `kernel_cls(model, trajectory_length=0.1)`, 500 + 20,000, posterior mean
    within 0.05 of the generating probabilities -- and, tighter, the conjugate posterior's moments."""
    from numpyro_b200 import models
    from numpyro_b200.infer import HMC, MCMC, NUTS
    from numpyro_b200.random import PRNGKey as DevKey

    data = fx["data"]
    kernel = (HMC if algo == "HMC" else NUTS)(models.beta_bernoulli, trajectory_length=0.1)
    mcmc = MCMC(kernel, num_warmup=500, num_samples=20000)
    mcmc.run(DevKey(2), data)
    x = mcmc.get_samples()["p_latent"]
    assert x.shape == (20000, 2)
    np.testing.assert_allclose(x.mean(0), [0.9, 0.1], atol=0.05)
    if algo == "NUTS":
        # (HMC with trajectory_length=0.1 takes ONE leapfrog of length 0.1 per transition once the
        # adapted step size exceeds it, hmc.py:399-401 -- the oracle does the same and mixes
        # slowly; the reference test asserts the loose bound only)
        s, n = data.sum(0), data.shape[0]
        a, b = 1.1 + s, 1.1 + n - s
        np.testing.assert_allclose(x.mean(0), a / (a + b), atol=1.5e-3)
        np.testing.assert_allclose(x.std(0), np.sqrt(a * b / ((a + b) ** 2 * (a + b + 1))), rtol=0.08)


# ---------------------------------------------------------------------------- Dirichlet-Categorical
def test_oracle_matches_reference_dirichlet_categorical(fx):
    from oracle.potentials import DirichletCategorical

    fam = DirichletCategorical(fx["dc/data"])
    U = np.array([fam.potential_and_grad(z)[0] for z in fx["dc/z"]])
    np.testing.assert_allclose(U, fx["dc/U"], rtol=5e-7)  # counts x log p vs 2000 gathered log-probs
    np.testing.assert_allclose(fam.constrain(fx["dc/z"][0])["p_latent"], fx["dc/constrained/p_latent"], rtol=1e-6)
    f64 = DirichletCategorical(fx["dc/data"], dtype=np.float64)
    for z in fx["dc/z"]:
        z = z.astype(np.float64)
        g = f64.potential_and_grad(z)[1]
        gn = np.array([(f64.potential_and_grad(z + 1e-6 * e)[0] - f64.potential_and_grad(z - 1e-6 * e)[0]) / 2e-6
                       for e in np.eye(2)])
        np.testing.assert_allclose(g, gn, rtol=1e-6, atol=1e-5)
    for part in (True, False):
        for seed in (0, 1, 2):
            (z0, _, _), _ = ohmc.find_valid_initial_params(Config(np.float32, part), PRNGKey(seed), fam)
            assert np.array_equal(z0, fx[f"dc/init_{'part' if part else 'legacy'}_{seed}"])


@pytest.mark.gpu
def test_cuda_dirichlet_categorical_potential_and_init(fx):
    from numpyro_b200 import models
    from numpyro_b200.infer import NUTS
    from oracle.potentials import DirichletCategorical

    data = fx["dc/data"]
    fam, f64 = DirichletCategorical(data), DirichletCategorical(data, dtype=np.float64)
    k = NUTS(models.dirichlet_categorical)
    for part in (True, False):
        k.partitionable = part
        for seed in (0, 1, 2):
            chain_key = PRNGKey(100 + seed)
            cfg = Config(np.float32, part)
            (z_or, _, _), _ = ohmc.find_valid_initial_params(cfg, cfg.split(chain_key)[1], fam)
            st = k.init(np.asarray(chain_key, np.uint32), 10, model_args=(data,))
            assert np.array_equal(st.z["p_latent"], z_or)
    U, g = k.potential_and_grad(fx["dc/z"])
    np.testing.assert_allclose(U, fx["dc/U"], rtol=2e-6)
    for i, z in enumerate(fx["dc/z"]):
        np.testing.assert_allclose(g[i], f64.potential_and_grad(z.astype(np.float64))[1], rtol=2e-5, atol=2e-3)
    p = k.postprocess_fn((data,), {})({"p_latent": fx["dc/z"][0]})["p_latent"]
    np.testing.assert_allclose(p, fx["dc/constrained/p_latent"], rtol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("dense_mass", [False, True])
@pytest.mark.parametrize("algo", ["HMC", "NUTS"])
def test_cuda_dirichlet_categorical_posterior(fx, algo, dense_mass):
    """test_mcmc.py:290-305: `kernel_cls(model, trajectory_length=1.0, dense_mass=...)`, 100 + 20,000,
    posterior mean within 0.02 of the generating probabilities -- and the conjugate posterior's mean."""
    from numpyro_b200 import models
    from numpyro_b200.infer import HMC, MCMC, NUTS
    from numpyro_b200.random import PRNGKey as DevKey

    data = fx["dc/data"]
    kernel = (HMC if algo == "HMC" else NUTS)(models.dirichlet_categorical, trajectory_length=1.0,
                                               dense_mass=dense_mass)
    mcmc = MCMC(kernel, num_warmup=100, num_samples=20000)
    mcmc.run(DevKey(2), data)
    x = mcmc.get_samples()["p_latent"]
    assert x.shape == (20000, 3)
    np.testing.assert_allclose(x.sum(-1), 1.0, rtol=1e-5)
    np.testing.assert_allclose(x.mean(0), [0.1, 0.6, 0.3], atol=0.02)
    if algo == "NUTS":
        a = 1.0 + np.bincount(data, minlength=3)
        np.testing.assert_allclose(x.mean(0), a / a.sum(), atol=2e-3)
