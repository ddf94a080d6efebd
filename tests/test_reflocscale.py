"""Eight schools under ``LocScaleReparam`` (``numpyro/infer/reparam.py:64-145``):
oracle and CUDA potential against the REFERENCE's own ``potential_energy`` /
``initialize_model`` / ``postprocess_fn`` on the reparametrised model, executed from
/root/reference over the NumPy stand-in (``tests/golden/make_reflocscale.py``)."""
import os

import numpy as np
import pytest

from oracle import hmc as ohmc
from oracle.hmc_util import Config
from oracle.potentials import EightSchoolsLocScale
from oracle.threefry import PRNGKey

Y8 = np.array([28.0, 8.0, -3.0, 7.0, -1.0, 1.0, 18.0, 12.0], np.float32)
S8 = np.array([15.0, 10.0, 16.0, 11.0, 9.0, 11.0, 10.0, 18.0], np.float32)
CASES = [(0.0, "c0"), (0.3, "c0.3"), (1.0, "c1")]


@pytest.fixture(scope="module")
def fx():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "reflocscale.npz"))


@pytest.mark.parametrize("c,tag", CASES)
def test_oracle_matches_reference_locscale(fx, c, tag):
    fam = EightSchoolsLocScale(S8, Y8, c)
    U = np.array([fam.potential_and_grad(z)[0] for z in fx[tag + "/z"]])
    np.testing.assert_allclose(U, fx[tag + "/U"], rtol=2e-6)
    con = fam.constrain(fx[tag + "/z"][0])
    assert set(con) == {k.split("/")[-1] for k in fx.files if k.startswith(tag + "/constrained/")}
    for k, v in con.items():
        np.testing.assert_allclose(v, fx[f"{tag}/constrained/{k}"], rtol=2e-6)
    fam64 = EightSchoolsLocScale(S8, Y8, c, np.float64)
    for z in fx[tag + "/z"][:3]:
        z = z.astype(np.float64)
        g = fam64.potential_and_grad(z)[1]
        gn = np.array([(fam64.potential_and_grad(z + 1e-6 * e)[0] - fam64.potential_and_grad(z - 1e-6 * e)[0]) / 2e-6
                       for e in np.eye(10)])
        np.testing.assert_allclose(g, gn, rtol=1e-5, atol=1e-6)
    for part in (True, False):  # init search: site order and key consumption
        for seed in (0, 1, 2):
            (z, _, _), _ = ohmc.find_valid_initial_params(Config(np.float32, part), PRNGKey(seed), fam)
            assert np.array_equal(z, fx[f"{tag}/init_{'part' if part else 'legacy'}_{seed}"])


def test_reparam_surface():
    from numpyro_b200 import handlers, models
    from numpyro_b200.infer.reparam import LocScaleReparam

    with pytest.raises(ValueError, match="does not satisfy `0 <= centered <= 1`"):  # reparam.py:95-99
        LocScaleReparam(1.5)
    m = handlers.reparam(models.eight_schools, config={"theta": LocScaleReparam(0.3)})
    b = m.bind(8, S8, y=Y8)
    assert [k for k, _ in b.trace_sites] == ["mu", "tau", "theta_decentered"]
    assert [k for k, _ in models.eight_schools.bind(8, S8, y=Y8).trace_sites] == ["mu", "tau", "theta"]
    assert models.lookup(m.__name__) is m
    with pytest.raises(NotImplementedError):
        handlers.reparam(models.logistic_regression, config={"coefs": LocScaleReparam(0.0)})


@pytest.mark.gpu
@pytest.mark.parametrize("c,tag", CASES)
def test_cuda_locscale_potential_init_and_postprocess(fx, c, tag):
    from numpyro_b200 import handlers, models
    from numpyro_b200.infer import NUTS
    from numpyro_b200.infer.reparam import LocScaleReparam

    model = handlers.reparam(models.eight_schools, config={"theta": LocScaleReparam(c)})
    fam64 = EightSchoolsLocScale(S8, Y8, c, np.float64)
    k = NUTS(model)
    for part in (True, False):
        k.partitionable = part
        for seed in (0, 1, 2):
            chain_key = PRNGKey(100 + seed)
            cfg = Config(np.float32, part)
            (z_or, _, _), _ = ohmc.find_valid_initial_params(cfg, cfg.split(chain_key)[1],
                                                             EightSchoolsLocScale(S8, Y8, c))
            st = k.init(np.asarray(chain_key, np.uint32), 10, model_args=(8, S8), model_kwargs={"y": Y8})
            assert np.array_equal(k._bound.ravel(st.z), z_or)
    U, g = k.potential_and_grad(fx[tag + "/z"])
    np.testing.assert_allclose(U, fx[tag + "/U"], rtol=3e-6)
    for i, z in enumerate(fx[tag + "/z"]):
        g64 = fam64.potential_and_grad(z.astype(np.float64))[1]
        np.testing.assert_allclose(g[i], g64, rtol=1e-4, atol=1e-5 * np.abs(g64).max())
    zd = k._bound.unravel(fx[tag + "/z"][0])
    for name, v in k.postprocess_fn((8, S8), {"y": Y8})(zd).items():
        np.testing.assert_allclose(v, fx[f"{tag}/constrained/{name}"], rtol=2e-6)


@pytest.mark.gpu
def test_cuda_locscale_posteriors_agree_across_parametrisations():
    """tier 4: theta's posterior is the same whichever `centered` samples it."""
    from numpyro_b200 import handlers, models
    from numpyro_b200.diagnostics import split_gelman_rubin
    from numpyro_b200.infer import MCMC, NUTS
    from numpyro_b200.infer.reparam import LocScaleReparam
    from numpyro_b200.random import PRNGKey as DevKey

    got = {}
    for c in (0.0, 0.3):
        model = handlers.reparam(models.eight_schools, config={"theta": LocScaleReparam(c)})
        m = MCMC(NUTS(model, target_accept_prob=0.9), num_warmup=1000, num_samples=4000, num_chains=8,
                 chain_method="vectorized")
        m.run(DevKey(int(c * 10) + 1), 8, S8, y=Y8)
        s = m.get_samples(group_by_chain=True)
        assert "theta" in s and "theta_decentered" in s
        assert float(np.max(split_gelman_rubin(s["theta"].astype(np.float64)))) < 1.02
        got[c] = np.concatenate([s["theta"].reshape(-1, 8), s["mu"].reshape(-1, 1)], axis=1)
    ref = MCMC(NUTS(models.eight_schools_noncentered, target_accept_prob=0.9), num_warmup=1000,
               num_samples=4000, num_chains=8, chain_method="vectorized")
    ref.run(DevKey(7), 8, S8, y=Y8)
    r = ref.get_samples()
    base = np.concatenate([r["theta"], r["mu"][:, None]], axis=1)
    sd = base.std(0)
    for c, x in got.items():
        assert np.all(np.abs(x.mean(0) - base.mean(0)) < 0.12 * sd), c
        assert np.all(np.abs(x.std(0) / sd - 1) < 0.12), c
