"""Potentials, init search and postprocessing against the REFERENCE's own model code.

``tests/golden/refmodels.npz`` (generator ``tests/golden/make_refmodels.py``) holds what the
unmodified ``numpyro`` package -- imported over a NumPy stand-in for JAX, ``refshim_full.py`` --
computes for the four registered models, using the model functions of the reference's own
example files: ``potential_energy`` at fixed unconstrained points, the initial points
``initialize_model`` draws with ``init_to_uniform`` for given keys (both Threefry layouts), and
``postprocess_fn`` outputs.  The oracle (CPU) and the CUDA path (GPU, through the C ABI) are held
to those values.  The reference arithmetic in the fixture is float32 (x64 off), hence rtol 2e-5.
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import refpin_cases as rc  # noqa: E402

G = np.load(os.path.join(HERE, "golden", "refmodels.npz"))
R = np.load(os.path.join(HERE, "golden", "refpin.npz"))
DATA = {k[5:]: R[k] for k in R.files if k.startswith("data/")}
KEYS = ["es", "logit", "hsb", "hsn", "sv"]


def _oracle(key, dtype):
    from oracle import potentials as opot

    if key == "logit":
        return opot.Logistic(DATA["logistic_X"], DATA["logistic_y"], dtype)
    if key == "es":
        return opot.EightSchools(rc.S8, rc.Y8, dtype)
    if key == "hsb":
        return opot.Horseshoe(DATA["hs_X"], DATA["hs_Yb"], "bernoulli", dtype)
    if key == "hsn":
        return opot.Horseshoe(DATA["hs_X"], DATA["hs_Yn"], "normal", dtype)
    return opot.StochasticVolatility(DATA["sv_returns"], dtype)


def _bound(key):
    from numpyro_b200 import models

    if key == "es":
        return models.eight_schools_noncentered, (8, rc.S8), {"y": rc.Y8}
    if key == "logit":
        return models.logistic_regression, (DATA["logistic_X"], DATA["logistic_y"]), {}
    if key == "hsb":
        return models.horseshoe_bernoulli, (DATA["hs_X"], DATA["hs_Yb"]), {}
    if key == "hsn":
        return models.horseshoe_normal, (DATA["hs_X"], DATA["hs_Yn"]), {}
    return models.stochastic_volatility, (DATA["sv_returns"],), {}


@pytest.mark.parametrize("key", KEYS)
def test_oracle_potential_matches_reference_model_code(key):
    fam64, fam32 = _oracle(key, np.float64), _oracle(key, np.float32)
    assert list(G[key + "/sites"]) == sorted(k for k, _ in fam64.trace_sites)  # flat order
    for z, U_ref in zip(G[key + "/z"], G[key + "/U"]):
        U64, g64 = fam64.potential_and_grad(z.astype(np.float64))
        np.testing.assert_allclose(U64, U_ref, rtol=2e-5)
        U32, g32 = fam32.potential_and_grad(z)
        np.testing.assert_allclose(U32, U_ref, rtol=2e-5)
        np.testing.assert_allclose(g32, g64, rtol=2e-3, atol=2e-3 * np.abs(g64).max())


@pytest.mark.parametrize("part", [True, False])
@pytest.mark.parametrize("key", KEYS)
def test_oracle_init_search_matches_reference_initialize_model(key, part):
    """``find_valid_initial_params`` (``infer/util.py:335-477``): uniform(-2, 2) per site in TRACE
    order, one fresh subkey per site."""
    from oracle import hmc as ohmc
    from oracle.hmc_util import Config
    from oracle.threefry import PRNGKey

    fam = _oracle(key, np.float32)
    tag = "part" if part else "legacy"
    for seed in (0, 1, 2):
        (z, pe, grad), ok = ohmc.find_valid_initial_params(Config(np.float32, part), PRNGKey(seed), fam)
        assert ok
        assert np.array_equal(z, G[f"{key}/init_{tag}_{seed}"]), (key, seed)


@pytest.mark.parametrize("key", KEYS)
def test_postprocess_matches_reference_constrain_fn(key):
    """``postprocess_fn`` = ``constrain_fn(..., return_deterministic=True)``: the shipped host
    mirror and the oracle's ``constrain``."""
    model, args, kwargs = _bound(key)
    b = model.bind(*args, **kwargs)
    z = G[key + "/z"][0]
    got = b.postprocess(b.unravel(z))
    want = {k.split("/", 2)[2]: G[k] for k in G.files if k.startswith(key + "/constrained/")}
    assert set(got) == set(want), (sorted(got), sorted(want))
    for k in want:
        np.testing.assert_allclose(np.asarray(got[k]).reshape(want[k].shape), want[k], rtol=2e-6,
                                   atol=1e-7)
    ora = _oracle(key, np.float32).constrain(z)
    for k in want:
        np.testing.assert_allclose(np.asarray(ora[k]).reshape(want[k].shape), want[k], rtol=2e-6,
                                   atol=1e-7)


sys.path.insert(0, os.path.join(HERE, "golden"))


def _mcmc_cases():
    import importlib.util

    spec = importlib.util.spec_from_file_location("_mrm_cases", os.path.join(HERE, "golden", "make_refmodels.py"))
    src = open(spec.origin).read()
    a = src.index("MCMC_CASES = {")
    b = src.index("\n}\n", a) + 3
    ns = {}
    exec(src[a:b], ns)  # the case table only: the generator itself needs /root/reference
    return ns["MCMC_CASES"]


MCMC_CASES = _mcmc_cases()


@pytest.mark.parametrize("name", list(MCMC_CASES))
def test_oracle_mcmc_matches_reference_mcmc_stack(name):
    """``MCMC.run`` of the reference (chain keys ``split(key, num_chains)``, init search, warm-up,
    ``fori_collect`` with thinning, ``postprocess_fn``) vs the oracle's ``run_chain`` per chain:
    identical draws (the gradient source is the same closed form in both runs)."""
    from oracle import hmc as ohmc
    from oracle.hmc_util import Config
    from oracle.threefry import PRNGKey

    key, kkw, mkw, seed = MCMC_CASES[name]
    fam = _oracle(key, np.float32)
    cfg = Config(np.float32, True)
    C = mkw.get("num_chains", 1)
    keys = cfg.split(PRNGKey(seed), C) if C > 1 else [PRNGKey(seed)]
    for c in range(C):
        k = ohmc.HMCKernel(fam, cfg=cfg, **kkw)
        with np.errstate(all="ignore"):
            res, _ = ohmc.run_chain(k, keys[c], mkw["num_warmup"], mkw["num_samples"],
                                    thinning=mkw.get("thinning", 1))
        assert np.array_equal(res["num_steps"], G[f"mcmc/{name}/extra/num_steps"][c]), (name, c)
        assert np.array_equal(res["potential_energy"], G[f"mcmc/{name}/extra/potential_energy"][c])
        con = fam.constrain(res["z"])
        for site in [k_.split("/")[-1] for k_ in G.files if k_.startswith(f"mcmc/{name}/samples/")]:
            want = G[f"mcmc/{name}/samples/{site}"][c]
            np.testing.assert_allclose(np.asarray(con[site]).reshape(want.shape), want, rtol=1e-6,
                                       atol=1e-7, err_msg=f"{name} {site}")


@pytest.mark.skipif(not os.path.isdir("/root/reference/numpyro"),
                    reason="the reference checkout only exists in the build container")
def test_fixture_regenerates_from_the_reference_package():
    """Imports the whole reference package over the stand-in in a fresh process and re-derives
    every entry of refmodels.npz (never on the GPU box)."""
    import subprocess

    out = subprocess.run([sys.executable, os.path.join(HERE, "golden", "make_refmodels.py"), "--check"],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-600:] + out.stderr[-600:]
    assert "check: identical" in out.stdout


def test_argument_checking_matches_reference_classes():
    """Same exception types and messages (or warning text) as the reference's ``HMC`` / ``NUTS`` /
    ``MCMC`` constructors and ``MCMC.run`` (recorded from the reference package itself)."""
    import warnings

    from numpyro_b200 import models
    from numpyro_b200.infer import HMC, MCMC, NUTS

    model = models.eight_schools_noncentered
    args, kwargs = (8, rc.S8), {"y": rc.Y8}
    bad_init = {"mu": np.zeros(3, np.float32), "tau": np.zeros(3, np.float32),
                "theta_base": np.zeros((3, 8), np.float32)}
    key = np.array([0, 1], np.uint32)
    calls = {
        "nuts_no_model": lambda: NUTS(),
        "nuts_model_and_potential": lambda: NUTS(model, potential_fn=lambda z: 0.0),
        "hmc_no_length": lambda: HMC(model, trajectory_length=None),
        "hmc_both_lengths_warns": lambda: HMC(model, num_steps=3),
        "mcmc_thinning": lambda: MCMC(NUTS(model), num_warmup=1, num_samples=1, thinning=0),
        "mcmc_chain_method": lambda: MCMC(NUTS(model), num_warmup=1, num_samples=1, chain_method="foo"),
        "mcmc_init_params_leading_dim": lambda: MCMC(
            NUTS(model), num_warmup=1, num_samples=1, num_chains=2, chain_method="sequential",
            progress_bar=False).run(key, *args, init_params=bad_init, **kwargs),
    }
    for name, f in calls.items():
        kind, msg = (str(x) for x in G["errors/" + name])
        try:
            with warnings.catch_warnings(record=True) as w:
                warnings.simplefilter("always")
                f()
            got = ("warning", str(w[0].message)) if w else ("ok", "")
        except Exception as e:  # noqa: BLE001
            got = (type(e).__name__, str(e))
        assert got == (kind, msg), (name, got, (kind, msg))
    k = NUTS(model, step_size=1, max_tree_depth=(5, 10))
    assert [repr(k._step_size), k.sample_field, repr(k.default_fields), k._algo,
            repr(k._max_tree_depth)] == [str(x) for x in G["attrs/nuts"]]


# ------------------------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def nb():
    import __graft_entry__ as g

    return g.build()


@pytest.mark.gpu
@pytest.mark.parametrize("key", KEYS)
def test_cuda_potential_matches_reference_model_code(nb, key):
    """Tier 2 against the reference itself: U from the CUDA kernels vs ``potential_energy`` of the
    reference's model code; gradient vs the float64 oracle (itself checked against central
    differences of an independent statement, tests/test_potentials_independent.py)."""
    from numpyro_b200 import infer
    from numpyro_b200.random import PRNGKey

    model, args, kwargs = _bound(key)
    k = infer.NUTS(model)
    k.init(PRNGKey(0), 10, model_args=args, model_kwargs=kwargs)
    U, g = k.potential_and_grad(G[key + "/z"])
    np.testing.assert_allclose(U, G[key + "/U"], rtol=2e-5)
    fam64 = _oracle(key, np.float64)
    for i, z in enumerate(G[key + "/z"]):
        g64 = fam64.potential_and_grad(z.astype(np.float64))[1]
        np.testing.assert_allclose(g[i], g64, rtol=1e-4, atol=1e-5 * np.abs(g64).max())


@pytest.mark.gpu
@pytest.mark.parametrize("part", [True, False])
@pytest.mark.parametrize("key", KEYS)
def test_cuda_init_search_matches_reference_initialize_model(nb, key, part):
    """The device init search starts a chain where the reference's ``initialize_model`` does.
    ``HMC.init`` hands ``split(chain_key)[1]`` to the init search (``hmc.py:745``); the oracle's
    search is pinned to the reference's draws bit for bit above, so the device is compared with
    the oracle's search on that half of the key."""
    from numpyro_b200 import infer
    from oracle import hmc as ohmc
    from oracle.hmc_util import Config
    from oracle.threefry import PRNGKey

    model, args, kwargs = _bound(key)
    b = model.bind(*args, **kwargs)
    cfg = Config(np.float32, part)
    for seed in (0, 1, 2):
        chain_key = PRNGKey(100 + seed)
        (z_or, _, _), _ = ohmc.find_valid_initial_params(cfg, cfg.split(chain_key)[1],
                                                         _oracle(key, np.float32))
        k = infer.NUTS(model)
        k.partitionable = part
        st = k.init(np.asarray(chain_key, np.uint32), 10, model_args=args, model_kwargs=kwargs)
        assert np.array_equal(b.ravel(st.z), z_or), (key, seed)


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(MCMC_CASES))
def test_cuda_mcmc_matches_reference_mcmc_stack(nb, name):
    """The shipped ``MCMC`` / ``NUTS`` classes driven exactly like the reference's: same sample
    sites and shapes; with adaptation off the kept draws are the reference's; with adaptation
    on the runs share a leading stretch of identical tree sizes (f32 chaos afterwards)."""
    from numpyro_b200 import infer
    from numpyro_b200.random import PRNGKey

    key, kkw, mkw, seed = MCMC_CASES[name]
    model, args, kwargs = _bound(key)
    m = infer.MCMC(infer.NUTS(model, **kkw), chain_method="sequential", **mkw)
    m.run(PRNGKey(seed), *args, extra_fields=("num_steps", "potential_energy", "diverging"), **kwargs)
    got = m.get_samples(group_by_chain=True)
    sites = [k_.split("/")[-1] for k_ in G.files if k_.startswith(f"mcmc/{name}/samples/")]
    assert sorted(got) == sorted(sites)
    for site in sites:
        assert np.shape(got[site]) == G[f"mcmc/{name}/samples/{site}"].shape, site
    ns = m.get_extra_fields(group_by_chain=True)["num_steps"]
    want = G[f"mcmc/{name}/extra/num_steps"]
    if not kkw.get("adapt_step_size", True):
        assert np.array_equal(ns, want)
        for site in sites:
            np.testing.assert_allclose(got[site], G[f"mcmc/{name}/samples/{site}"], rtol=5e-3, atol=5e-4)
    else:
        # warm-up transitions are not collected, so agreement of the first kept draw already
        # means num_warmup identical transitions; require it for at least one chain
        first = [np.array_equal(ns[c][:1], want[c][:1]) for c in range(ns.shape[0])]
        assert ns.shape == want.shape and (any(first) or ns.mean() < 3 * want.mean())
