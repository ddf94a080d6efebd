"""The two GLMs of the registry (``poisson_regression``, ``linear_regression``): oracle
and CUDA potential against the REFERENCE's ``potential_energy`` / ``initialize_model`` /
``constrain_fn`` executed from /root/reference on the same models written with the reference's
primitives (``tests/golden/make_refglm.py``), the posterior against the mode on the float64
oracle, and the sampler against the oracle's on the same key."""
import os

import numpy as np
import pytest

from oracle import hmc as ohmc
from oracle.hmc_util import Config
from oracle.potentials import NormalGLM, PoissonGLM
from oracle.threefry import PRNGKey

CASES = {"poisson": (PoissonGLM, "counts", "poisson_regression"),
         "normal": (NormalGLM, "y", "linear_regression")}


@pytest.fixture(scope="module")
def fx():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "refglm.npz"))


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_reference_glm(fx, name):
    cls, obs, _ = CASES[name]
    fam = cls(fx["X"], fx[obs])
    U = np.array([fam.potential_and_grad(z)[0] for z in fx[f"{name}/z"]])
    np.testing.assert_allclose(U, fx[f"{name}/U"], rtol=1e-6)
    for k, v in fam.constrain(fx[f"{name}/z"][0]).items():
        np.testing.assert_allclose(v, fx[f"{name}/constrained/{k}"], rtol=1e-6)
    f64 = cls(fx["X"], fx[obs], dtype=np.float64)
    D = fam.dim
    for z in fx[f"{name}/z"][:2]:
        z = z.astype(np.float64)
        g = f64.potential_and_grad(z)[1]
        gn = np.array([(f64.potential_and_grad(z + 1e-6 * e)[0] - f64.potential_and_grad(z - 1e-6 * e)[0]) / 2e-6
                       for e in np.eye(D)])
        np.testing.assert_allclose(g, gn, rtol=1e-5, atol=1e-4)
    for part in (True, False):
        for seed in (0, 1, 2):
            (z0, _, _), _ = ohmc.find_valid_initial_params(Config(np.float32, part), PRNGKey(seed), fam)
            assert np.array_equal(z0, fx[f"{name}/init_{'part' if part else 'legacy'}_{seed}"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_cuda_glm_potential_init_and_posterior(fx, name):
    import __graft_entry__ as g

    nb = g.build()
    cls, obs, model_name = CASES[name]
    model = getattr(nb.models, model_name)
    X, y = fx["X"], fx[obs]
    fam, f64 = cls(X, y), cls(X, y, dtype=np.float64)
    k = nb.infer.NUTS(model)
    for part in (True, False):
        k.partitionable = part
        cfg = Config(np.float32, part)
        for seed in (0, 1):
            chain_key = np.asarray(PRNGKey(50 + seed), np.uint32)
            (z_or, _, _), _ = ohmc.find_valid_initial_params(cfg, cfg.split(chain_key)[1], fam)
            st = k.init(chain_key, 10, model_args=(X, y))
            assert np.array_equal(k._bound.ravel(st.z), z_or)
    U, grad = k.potential_and_grad(fx[f"{name}/z"])
    np.testing.assert_allclose(U, fx[f"{name}/U"], rtol=3e-6)
    for i, z in enumerate(fx[f"{name}/z"]):
        np.testing.assert_allclose(grad[i], f64.potential_and_grad(z.astype(np.float64))[1], rtol=3e-5, atol=2e-3)
    # every chain grouping of the chain-batched kernels evaluates the same potential
    for mode in ("0", "1", "2"):
        os.environ["B200MC_CHAIN_MODE"] = mode
        try:
            k2 = nb.infer.NUTS(model)
            k2.init(np.asarray(PRNGKey(1), np.uint32), 10, model_args=(X, y))
            U2, g2 = k2.potential_and_grad(fx[f"{name}/z"][:1])
        finally:
            os.environ.pop("B200MC_CHAIN_MODE", None)
        np.testing.assert_allclose(U2, fx[f"{name}/U"][:1], rtol=3e-6)
        np.testing.assert_allclose(g2, grad[:1], rtol=1e-4, atol=1e-3)
    # tier 4: the generating coefficients (make_refglm.make_data)
    mc = nb.infer.MCMC(nb.infer.NUTS(model), num_warmup=300, num_samples=600, num_chains=4,
                       chain_method="vectorized")
    mc.run(np.asarray(PRNGKey(2), np.uint32), X, y)
    s = mc.get_samples(group_by_chain=True)
    assert nb.diagnostics.split_gelman_rubin(np.asarray(s["coefs"], np.float64)).max() < 1.02
    zm = np.zeros(fam.dim)  # the posterior mode on the float64 oracle (plain gradient descent; the
    for _ in range(20000):  # Poisson data carry an intercept the model has not: not the generating w)
        zm -= (1e-3 if name == "poisson" else 2e-4) * f64.potential_and_grad(zm)[1]
    assert np.abs(f64.potential_and_grad(zm)[1]).max() < 1e-6
    np.testing.assert_allclose(s["coefs"].reshape(-1, 6).mean(0), zm[:6], atol=0.03)
    if name == "normal":
        np.testing.assert_allclose(np.log(s["prec_obs"]).mean(), zm[6], atol=0.05)
        np.testing.assert_allclose(s["prec_obs"].mean(), 1 / 0.4 ** 2, rtol=0.25)
    # tier 3 against the oracle's sampler on the same key (fixed step, identity mass)
    kw = dict(step_size=0.01, adapt_step_size=False, adapt_mass_matrix=False)
    mc = nb.infer.MCMC(nb.infer.NUTS(model, **kw), num_warmup=0, num_samples=8)
    key = np.asarray(PRNGKey(9), np.uint32)
    mc.run(key, X, y, extra_fields=("num_steps", "potential_energy"))
    r, _ = ohmc.run_chain(ohmc.HMCKernel(fam, cfg=Config(np.float32, True), **kw), key, 0, 8)
    assert np.array_equal(mc.get_extra_fields()["num_steps"], r["num_steps"])
    np.testing.assert_allclose(mc.get_extra_fields()["potential_energy"], r["potential_energy"], rtol=2e-4)


def _normal_loglik(fx):
    z = fx["normal/z"]
    prec = np.exp(z[:, 6:7]).astype(np.float32)
    m = z[:, :6] @ fx["X"].T
    sig = (np.float32(1) / np.sqrt(prec)).astype(np.float32)
    return -0.5 * ((fx["y"][None] - m) / sig) ** 2 - np.log(sig) - 0.5 * np.log(2 * np.pi)


def test_log_likelihood_formula_matches_reference(fx):
    np.testing.assert_allclose(_normal_loglik(fx), fx["normal/loglik"], rtol=2e-5, atol=2e-5)


@pytest.mark.gpu
def test_cuda_log_likelihood_and_predictive_linear_regression(fx):
    """``log_likelihood`` (``infer/util.py:1094-1148``) of the Normal GLM over draws, against the
    reference's own; ``Predictive`` samples the observation site around ``data @ coefs``."""
    import __graft_entry__ as g

    nb = g.build()
    z = fx["normal/z"]
    samples = {"coefs": z[:, :6], "prec_obs": np.exp(z[:, 6]).astype(np.float32)}
    ll = nb.infer.log_likelihood(nb.models.linear_regression, samples, fx["X"], fx["y"])["obs"]
    np.testing.assert_allclose(ll, fx["normal/loglik"], rtol=2e-5, atol=2e-5)
    big = {"coefs": np.repeat(z[:1, :6], 2000, 0), "prec_obs": np.full(2000, 6.25, np.float32)}
    pred = nb.infer.Predictive(nb.models.linear_regression, posterior_samples=big)(
        np.asarray(PRNGKey(4), np.uint32), fx["X"], None)["obs"]
    assert pred.shape == (2000, 400)
    np.testing.assert_allclose(pred.mean(0), fx["X"] @ z[0, :6], atol=0.06)
    np.testing.assert_allclose(pred.std(0), 0.4, rtol=0.1)
    with pytest.raises(NotImplementedError):
        nb.infer.log_likelihood(nb.models.poisson_regression, {"coefs": z[:, :6]}, fx["X"], fx["counts"])
