"""``HMCGibbs`` (``numpyro/infer/hmc_gibbs.py:36-190``): the oracle's restatement and
the CUDA kernels with conditioned sites against the REFERENCE's own ``HMCGibbs`` around its
``NUTS(model)``, executed from /root/reference over the NumPy stand-in
(``tests/golden/make_refgibbs.py`` -> ``refgibbs.npz``): non-centred eight schools, Gibbs site
``mu`` (conjugate update), 12 adaptation + 10 sampling transitions, both Threefry layouts."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
import refgibbs_cases as gc  # noqa: E402
import refpin_cases as rc  # noqa: E402

from oracle import gibbs as ogibbs  # noqa: E402
from oracle import potentials as opot  # noqa: E402
from oracle.hmc_util import Config  # noqa: E402
from oracle.threefry import PRNGKey  # noqa: E402


@pytest.fixture(scope="module")
def fx():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "refgibbs.npz"))


def test_oracle_conditioned_potential_matches_reference_model_code(fx):
    """``condition`` + ``substitute`` of a site (``hmc_gibbs.py:30-34``): its CONSTRAINED value, no
    Jacobian term -- for a real site (``mu``) and a positive one (``tau``)."""
    full = opot.EightSchools(rc.S8, rc.Y8)
    for name in ("mu", "tau"):
        fam = ogibbs.Conditioned(full, [name], positive=("tau",))
        assert fam.dim == 9 and [k for k, _ in fam.trace_sites] == [k for k in ("mu", "tau", "theta_base") if k != name]
        for v, z, U in zip(fx[f"cond/{name}/values"], fx[f"cond/{name}/z"], fx[f"cond/{name}/U"]):
            fam.set_values({name: v})
            np.testing.assert_allclose(fam.potential_and_grad(z)[0], U, rtol=3e-7)
        f64 = ogibbs.Conditioned(opot.EightSchools(rc.S8, rc.Y8, np.float64), [name], positive=("tau",))
        f64.set_values({name: fx[f"cond/{name}/values"][1]})
        z = fx[f"cond/{name}/z"][1].astype(np.float64)
        g = f64.potential_and_grad(z)[1]
        gn = np.array([(f64.potential_and_grad(z + 1e-6 * e)[0] - f64.potential_and_grad(z - 1e-6 * e)[0]) / 2e-6
                       for e in np.eye(9)])
        np.testing.assert_allclose(g, gn, rtol=1e-5, atol=1e-5)


def _oracle_run(part, seed):
    cfg = Config(np.float32, part)
    k = ogibbs.HMCGibbs(
        opot.EightSchools(rc.S8, rc.Y8), gc.make_gibbs_fn(lambda key: cfg.normal(key, ())), ["mu"],
        prior_sample=lambda name, key, c: gc.prior_sample_mu(c.normal(key, (1,))),
        positive=("tau",), cfg=cfg, **gc.KERNEL_KW)
    st = k.init(PRNGKey(seed), gc.NUM_WARMUP)
    states = [st]
    for _ in range(gc.NUM_STEPS):
        st = k.sample(st)
        states.append(st)
    return states


@pytest.mark.parametrize("part", [True, False])
@pytest.mark.parametrize("seed", gc.SEEDS)
def test_oracle_hmcgibbs_matches_reference_state_by_state(fx, part, seed):
    tag = f"mu/{'part' if part else 'legacy'}_{seed}"
    states = _oracle_run(part, seed)
    assert np.array_equal([s.hmc_state.num_steps for s in states], fx[f"{tag}/num_steps"])
    assert np.array_equal(np.stack([np.asarray(s.rng_key, np.uint32) for s in states]), fx[f"{tag}/rng_key"])
    assert np.array_equal(np.stack([s.hmc_state.z for s in states]), fx[f"{tag}/z"])
    assert np.array_equal(np.array([s.z["mu"] for s in states], np.float32), fx[f"{tag}/mu"])
    assert np.array_equal(np.array([s.hmc_state.potential_energy for s in states], np.float32), fx[f"{tag}/pe"])
    assert np.array_equal(np.array([s.hmc_state.adapt_state.step_size for s in states], np.float32),
                          fx[f"{tag}/step_size"])


def test_oracle_hmcgibbs_around_hmc_matches_reference(fx):
    """The wrapper around ``HMC`` (fixed trajectory length, step size adapted over 6 transitions)."""
    cfg = Config(np.float32, True)
    k = ogibbs.HMCGibbs(
        opot.EightSchools(rc.S8, rc.Y8), gc.make_gibbs_fn(lambda key: cfg.normal(key, ())), ["mu"],
        prior_sample=lambda name, key, c: gc.prior_sample_mu(c.normal(key, (1,))),
        positive=("tau",), cfg=cfg, algo="HMC", **gc.HMC_KW)
    st = k.init(PRNGKey(gc.HMC_SEED), gc.HMC_WARMUP)
    states = [st]
    for _ in range(gc.HMC_STEPS):
        st = k.sample(st)
        states.append(st)
    assert np.array_equal([s.hmc_state.num_steps for s in states], fx["hmc/num_steps"])
    assert np.array_equal(np.stack([s.hmc_state.z for s in states]), fx["hmc/z"])
    assert np.array_equal(np.array([s.z["mu"] for s in states], np.float32), fx["hmc/mu"])
    assert np.array_equal(np.array([s.hmc_state.accept_prob for s in states], np.float32), fx["hmc/accept_prob"])
    assert np.array_equal(np.stack([np.asarray(s.rng_key, np.uint32) for s in states]), fx["hmc/rng_key"])


# ---------------------------------------------------------------------------------------- CUDA
def _cuda_kernel(nb, part):
    from numpyro_b200 import random as b2random

    inner = nb.infer.NUTS(nb.models.eight_schools_noncentered, **gc.KERNEL_KW)
    inner.partitionable = part
    gibbs_fn = gc.make_gibbs_fn(lambda key: b2random.normal(key, (), part))
    return nb.infer.HMCGibbs(inner, gibbs_fn=gibbs_fn, gibbs_sites=["mu"])


@pytest.mark.gpu
def test_cuda_conditioned_potential(fx):
    import __graft_entry__ as g

    nb = g.build()
    full = nb.models.eight_schools_noncentered.bind(8, rc.S8, y=rc.Y8)
    for name in ("mu", "tau"):
        k = nb.infer.NUTS(nb.models.eight_schools_noncentered)
        vals = fx[f"cond/{name}/values"]
        k._ecs = {"degree": 0, "ref": None, "after_create": None}
        k._vectorized = True
        k._create(full.condition([name]), 3, 10)
        helper = nb.infer.HMCGibbs(nb.infer.NUTS(nb.models.eight_schools_noncentered), lambda **kw: {}, [name])
        helper.inner_kernel = k
        helper._push_gibbs({name: vals})
        U, grad = k.potential_and_grad(fx[f"cond/{name}/z"])
        np.testing.assert_allclose(U, fx[f"cond/{name}/U"], rtol=2e-6)
        f64 = ogibbs.Conditioned(opot.EightSchools(rc.S8, rc.Y8, np.float64), [name], positive=("tau",))
        for i in range(3):
            f64.set_values({name: vals[i]})
            np.testing.assert_allclose(grad[i], f64.potential_and_grad(fx[f"cond/{name}/z"][i].astype(np.float64))[1],
                                       rtol=2e-5, atol=2e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("part", [True, False])
def test_cuda_hmcgibbs_matches_reference(fx, part):
    """init: Gibbs value (prior draw on the prototype trace's key), initial point and keys bit for
    bit; every transition: same tree sizes and keys, positions within float round-off of the
    reference's (the device potential sums in another order)."""
    import __graft_entry__ as g

    nb = g.build()
    args, kwargs = (8, rc.S8), {"y": rc.Y8}
    for seed in gc.SEEDS:
        tag = f"mu/{'part' if part else 'legacy'}_{seed}"
        k = _cuda_kernel(nb, part)
        st = k.init(np.asarray(PRNGKey(seed), np.uint32), gc.NUM_WARMUP, None, args, kwargs)
        view = k.inner_kernel._bound
        assert np.array_equal(np.float32(st.z["mu"]), fx[f"{tag}/mu"][0])
        assert np.array_equal(view.ravel(st.hmc_state.z), fx[f"{tag}/z"][0])
        assert np.array_equal(np.asarray(st.rng_key, np.uint32), fx[f"{tag}/rng_key"][0])
        n_same = 0
        for i in range(1, gc.NUM_STEPS + 1):
            st = k.sample(st, args, kwargs)
            assert np.array_equal(np.asarray(st.rng_key, np.uint32), fx[f"{tag}/rng_key"][i])
            z_dev, z_ref = view.ravel(st.hmc_state.z), fx[f"{tag}/z"][i]
            if int(st.hmc_state.num_steps) != int(fx[f"{tag}/num_steps"][i]) or \
                    np.abs(z_dev - z_ref).max() > 2e-2:
                break  # a tree / proposal decision within round-off of its threshold: the runs separate
            n_same += 1
            if i <= 4:  # before round-off has been amplified along the chain
                np.testing.assert_allclose(z_dev, z_ref, rtol=2e-3, atol=2e-3)
                np.testing.assert_allclose(np.float32(st.z["mu"]), fx[f"{tag}/mu"][i], rtol=2e-3, atol=2e-3)
                np.testing.assert_allclose(st.hmc_state.potential_energy, fx[f"{tag}/pe"][i], rtol=1e-3)
        assert n_same >= 5, (tag, n_same)


@pytest.mark.gpu
def test_cuda_hmcgibbs_around_hmc_matches_reference(fx):
    import __graft_entry__ as g
    from numpyro_b200 import random as b2random

    nb = g.build()
    args, kwargs = (8, rc.S8), {"y": rc.Y8}
    inner = nb.infer.HMC(nb.models.eight_schools_noncentered, **gc.HMC_KW)
    k = nb.infer.HMCGibbs(inner, gc.make_gibbs_fn(lambda key: b2random.normal(key, (), True)), ["mu"])
    st = k.init(np.asarray(PRNGKey(gc.HMC_SEED), np.uint32), gc.HMC_WARMUP, None, args, kwargs)
    view = k.inner_kernel._bound
    assert np.array_equal(view.ravel(st.hmc_state.z), fx["hmc/z"][0])
    assert np.array_equal(np.float32(st.z["mu"]), fx["hmc/mu"][0])
    for i in range(1, gc.HMC_STEPS + 1):
        st = k.sample(st, args, kwargs)
        assert np.array_equal(np.asarray(st.rng_key, np.uint32), fx["hmc/rng_key"][i])
        assert int(st.hmc_state.num_steps) == int(fx["hmc/num_steps"][i])
        tol = 2e-3 if i <= 5 else 3e-2  # no tree decisions to flip: the runs drift by round-off only
        np.testing.assert_allclose(view.ravel(st.hmc_state.z), fx["hmc/z"][i], rtol=tol, atol=tol)
        np.testing.assert_allclose(st.hmc_state.accept_prob, fx["hmc/accept_prob"][i], atol=max(tol, 5e-3))


@pytest.mark.gpu
def test_cuda_hmcgibbs_vectorized_chains_are_independent():
    """Vectorized chains (``uint32[C, 2]`` keys): chain c is the single-chain run on key c."""
    import __graft_entry__ as g
    from numpyro_b200 import random as b2random

    nb = g.build()
    args, kwargs = (8, rc.S8), {"y": rc.Y8}
    one = gc.make_gibbs_fn(lambda key: b2random.normal(key, (), True))

    def batched(rng_key, gibbs_sites, hmc_sites):
        if np.ndim(rng_key) == 1:
            return one(rng_key, gibbs_sites, hmc_sites)
        rows = [one(rng_key[c], None, {k: v[c] for k, v in hmc_sites.items()})["mu"] for c in range(len(rng_key))]
        return {"mu": np.asarray(rows, np.float32)}

    keys = np.stack([PRNGKey(s) for s in (3, 4, 5)]).astype(np.uint32)
    kv = nb.infer.HMCGibbs(nb.infer.NUTS(nb.models.eight_schools_noncentered, **gc.KERNEL_KW), batched, ["mu"])
    sv = kv.init(keys, 8, None, args, kwargs)
    singles = []
    for c in range(3):
        k1 = nb.infer.HMCGibbs(nb.infer.NUTS(nb.models.eight_schools_noncentered, **gc.KERNEL_KW), batched, ["mu"])
        singles.append((k1, k1.init(keys[c], 8, None, args, kwargs)))
    for step in range(12):
        for c, (k1, s1) in enumerate(singles):
            assert np.array_equal(np.asarray(sv.z["mu"])[c], np.asarray(s1.z["mu"])), (step, c)
            assert np.array_equal(np.asarray(sv.hmc_state.z["theta_base"])[c], s1.hmc_state.z["theta_base"]), (step, c)
            assert np.array_equal(np.asarray(sv.rng_key)[c], s1.rng_key)
            assert int(np.asarray(sv.hmc_state.num_steps)[c]) == int(s1.hmc_state.num_steps)
        sv = kv.sample(sv, args, kwargs)
        singles = [(k1, k1.sample(s1, args, kwargs)) for k1, s1 in singles]


@pytest.mark.gpu
def test_cuda_hmcgibbs_posterior():
    """tier 4: HMC-within-Gibbs samples the same posterior as NUTS on the full model
    (``test_hmc_gibbs.py`` style: moments, here against the plain CUDA NUTS run)."""
    import __graft_entry__ as g

    nb = g.build()
    args, kwargs = (8, rc.S8), {"y": rc.Y8}
    key = np.asarray(PRNGKey(7), np.uint32)
    m1 = nb.infer.MCMC(_cuda_kernel(nb, True), num_warmup=300, num_samples=1500)
    m1.run(key, *args, **kwargs)
    s1 = m1.get_samples()
    assert set(s1) == {"mu", "tau", "theta_base", "theta"} and s1["theta"].shape == (1500, 8)
    m2 = nb.infer.MCMC(nb.infer.NUTS(nb.models.eight_schools_noncentered), num_warmup=500, num_samples=4000,
                       num_chains=4, chain_method="vectorized")
    m2.run(key, *args, **kwargs)
    s2 = m2.get_samples()
    for name, tol in (("mu", 0.6), ("theta", 0.7)):
        np.testing.assert_allclose(s1[name].mean(0), s2[name].mean(0), atol=tol)
    np.testing.assert_allclose(np.log(s1["tau"]).mean(), np.log(s2["tau"]).mean(), atol=0.35)
    np.testing.assert_allclose(s1["mu"].std(), s2["mu"].std(), rtol=0.25)
