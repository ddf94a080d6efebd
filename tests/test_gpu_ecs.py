"""HMCECS on the device (``numpyro/infer/hmc_gibbs.py:502-665``,
``contrib/ecs_proxies.py``) against ``oracle/ecs.py`` (itself held to the reference's own
``_update_block`` / ``_subsample_fn`` / ``perturbed_method`` in ``tests/test_refecs.py``):
integer work bit for bit, the subsampled potential to a stated float32 tolerance, whole runs
decision for decision, and the posterior against full-data NUTS."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _problem(n=20000, D=8, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, D)).astype(np.float32)
    X[:, -1] = 1.0
    th = rng.standard_normal(D) * 0.5
    y = (rng.random(n) < 1 / (1 + np.exp(-X @ th))).astype(np.uint8)
    return X, y, th


def _mode(X, y, iters=25):
    """posterior mode by Newton steps in float64 (the reference example gets it from SVI)"""
    X64, th = X.astype(np.float64), np.zeros(X.shape[1])
    for _ in range(iters):
        s = 1 / (1 + np.exp(-X64 @ th))
        g = th + X64.T @ (s - y)
        H = np.eye(th.size) + (X64.T * (s * (1 - s))) @ X64
        th = th - np.linalg.solve(H, g)
    return th


@pytest.mark.parametrize("part", [True, False])
def test_randint_bit_exact(part):
    from numpyro_b200 import _lib
    from oracle import ecs as oecs, hmc_util as ohu
    from oracle.threefry import PRNGKey

    lib, cfg = _lib.lib(), ohu.Config(partitionable=part)
    for seed, n, lo, hi in [(0, 0, 0, 581012), (1, 10, 0, 581012), (2, 1000, 0, 100), (3, 7, -5, 3),
                            (4, 33, 0, 1), (5, 257, 0, 2**31 - 1), (6, 64, 0, 65537)]:
        key = PRNGKey(seed)
        kd = _lib.DeviceArray.from_numpy(np.asarray(key, np.uint32))
        out = _lib.DeviceArray((max(n, 1),), np.int32)
        _lib.check(lib.b200mc_random_randint(kd.ptr, n, lo, hi, int(part), out.ptr, None))
        want = oecs.randint(cfg, key, (n,) if n else (), lo, hi)
        np.testing.assert_array_equal(out.numpy().reshape(np.shape(want)), want)


def _kernels(X, y, m, degree, ref, num_blocks, part=True, **kw):
    from numpyro_b200 import models
    from numpyro_b200.infer import HMCECS, NUTS
    from oracle import ecs as oecs, hmc_util as ohu

    proxy = None if degree == 0 else HMCECS.taylor_proxy({"coefs": ref}, degree)
    inner = NUTS(models.logistic_regression, **kw)
    inner.partitionable = part
    gk = HMCECS(inner, num_blocks=num_blocks, proxy=proxy)
    fam = oecs.LogisticECS(X, y, m, degree, ref=ref)
    ok = oecs.HMCECSKernel(fam, num_blocks=num_blocks, cfg=ohu.Config(partitionable=part), **kw)
    return gk, ok


@pytest.mark.parametrize("part", [True, False])
def test_init_subsample_and_keys_match_oracle(part):
    from oracle.threefry import PRNGKey

    X, y, th = _problem(n=5000, D=5)
    m = 150
    gk, ok = _kernels(X, y, m, 2, th.astype(np.float32), 5, part)
    gs = gk.init(PRNGKey(4), 10, None, (X, y), {"subsample_size": m})
    os_ = ok.init(PRNGKey(4), 10)
    np.testing.assert_array_equal(gs.z["N"], os_.z["N"])
    np.testing.assert_array_equal(gs.rng_key, os_.rng_key)
    np.testing.assert_array_equal(gs.hmc_state.rng_key, os_.hmc_state.rng_key)
    np.testing.assert_allclose(gs.hmc_state.z["coefs"], os_.hmc_state.z, rtol=1e-6)  # same init draw
    assert len(set(gs.z["N"].tolist())) == m


@pytest.mark.parametrize("shape", ["small", "covtype"])
@pytest.mark.parametrize("degree", [0, 1, 2])
def test_tier2_subsampled_potential(degree, shape):
    """U and grad on the chain's subsample vs the float64 oracle.  The proxy variants subtract
    per-row Taylor expansions from per-row log-likelihoods (O(1) quantities, differences of
    1e-3..1e-5) and scale by n: float32 round-off of ~6e-8 per row becomes ~4e-8 n in U -- the
    reference's float32 arithmetic has the same structure."""
    from oracle import ecs as oecs, hmc_util as ohu
    from oracle.threefry import PRNGKey

    if shape == "small":
        X, y, th = _problem()
        m = 400
    else:
        rng = np.random.default_rng(0)
        n, D = 581012, 55
        X = rng.standard_normal((n, D)).astype(np.float32)
        X[:, -1] = 1.0
        th = rng.standard_normal(D) * 0.5
        y = (rng.random(n) < 1 / (1 + np.exp(-X @ th))).astype(np.uint8)
        m = 1000
    n, D = X.shape
    ref = (th + 0.01 * np.random.default_rng(5).standard_normal(D)).astype(np.float32)
    gk, _ = _kernels(X, y, m, degree, ref, 10)
    st = gk.init(PRNGKey(2), 10, None, (X, y), {"subsample_size": m})
    fam = oecs.LogisticECS(X, y, m, degree, ref=ref.astype(np.float64), dtype=np.float64)
    fam.set_subsample(st.z["N"])
    z = (th[None] + 0.03 * np.random.default_rng(6).standard_normal((3, D))).astype(np.float32)
    for i in range(3):
        U, g = gk.inner_kernel.potential_and_grad(z[i:i + 1])  # one chain: its subsample
        Uo, go = fam.potential_and_grad(z[i].astype(np.float64))
        tol_u = 1e-5 * abs(Uo) + (4e-8 * n if degree else 0.0)
        assert abs(U[0] - Uo) <= tol_u, (U[0], Uo, tol_u)
        scale = np.abs(go).max()
        np.testing.assert_allclose(g[0], go, rtol=2e-4, atol=(2e-5 if degree == 0 else 5e-4) * scale)


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("degree,part", [(0, True), (2, True), (2, False)])
def test_tier3_runs_decision_for_decision(degree, part, fused):
    """Fixed step size, no adaptation: 20 HMCECS transitions from the same key -- subsample
    proposals, Metropolis decisions, wrapper keys and tree sizes equal the oracle's."""
    from oracle.threefry import PRNGKey

    X, y, th = _problem(n=4000, D=6, seed=3)
    m = 200
    kw = dict(step_size=0.02, adapt_step_size=False, adapt_mass_matrix=False, max_tree_depth=5)
    gk, ok = _kernels(X, y, m, degree, th.astype(np.float32), 4, part, **kw)
    gk._fused_step = fused  # the one-warp step machine of the resident loop / the stand-alone kernels
    gs = gk.init(PRNGKey(7), 0, None, (X, y), {"subsample_size": m})
    os_ = ok.init(PRNGKey(7), 0)
    same_idx = same_steps = 0
    T = 20
    for t in range(T):
        gs = gk.sample(gs, (X, y), {"subsample_size": m})
        os_ = ok.sample(os_)
        np.testing.assert_array_equal(gs.rng_key, os_.rng_key)
        same_idx += int(np.array_equal(gs.z["N"], os_.z["N"]))
        same_steps += int(gs.hmc_state.num_steps == os_.hmc_state.num_steps)
        if same_idx == t + 1 and same_steps == t + 1:
            np.testing.assert_allclose(gs.accept_prob, os_.accept_prob, rtol=2e-3, atol=2e-4)
            np.testing.assert_allclose(gs.hmc_state.z["coefs"], os_.hmc_state.z, rtol=5e-3, atol=5e-4)
    assert same_idx == T and same_steps == T, (same_idx, same_steps)


@pytest.mark.parametrize("m,nb", [(37, 5), (64, 64), (10, 1), (250, 7), (33, 40)])
def test_block_update_shapes_match_oracle(m, nb):
    """Ragged blocks (padding, clamped slice start, more blocks than indices): the subsample after
    every one of 6 transitions equals the oracle's (whose `update_block` equals the reference's
    `_update_block` on these shapes, tests/test_refecs.py)."""
    from oracle.threefry import PRNGKey

    X, y, th = _problem(n=3000, D=4, seed=8)
    kw = dict(step_size=0.05, adapt_step_size=False, adapt_mass_matrix=False, max_tree_depth=4)
    gk, ok = _kernels(X, y, m, 2, th.astype(np.float32), nb, True, **kw)
    gs = gk.init(PRNGKey(3), 0, None, (X, y), {"subsample_size": m})
    os_ = ok.init(PRNGKey(3), 0)
    np.testing.assert_array_equal(gs.z["N"], os_.z["N"])
    for _ in range(6):
        gs = gk.sample(gs, (X, y), {"subsample_size": m})
        os_ = ok.sample(os_)
        np.testing.assert_array_equal(gs.rng_key, os_.rng_key)
        np.testing.assert_array_equal(gs.z["N"], os_.z["N"])
        assert int(gs.hmc_state.num_steps) == int(os_.hmc_state.num_steps)


def test_device_resident_loop_equals_step_by_step():
    """``run_loop`` (one launch: Gibbs step + transition per iteration, adaptation included) gives
    the draws, accept probabilities, tree sizes and final state of ``sample()`` called in a loop."""
    from numpyro_b200.random import PRNGKey

    X, y, th = _problem(n=6000, D=7, seed=5)
    m, T, W = 250, 45, 25
    out = []
    for fused in (False, True):
        gk, _ = _kernels(X, y, m, 2, th.astype(np.float32), 5, True, max_tree_depth=6)
        st = gk.init(PRNGKey(11), W, None, (X, y), {"subsample_size": m})
        if fused:
            raw, st = gk.run_loop(st, T, 0, 1, T, ["z", "accept_prob", "hmc_state.num_steps"], (X, y),
                                  {"subsample_size": m})
            out.append((raw["z"][:, 0], raw["accept_prob"][:, 0], raw["hmc_state.num_steps"][:, 0], st))
        else:
            zs, ap, ns = [], [], []
            for _ in range(T):
                st = gk.sample(st, (X, y), {"subsample_size": m})
                zs.append(st.hmc_state.z["coefs"]); ap.append(st.accept_prob); ns.append(st.hmc_state.num_steps)
            out.append((np.stack(zs), np.asarray(ap), np.asarray(ns), st))
    (z0, a0, n0, s0), (z1, a1, n1, s1) = out
    np.testing.assert_array_equal(n0, n1)
    np.testing.assert_array_equal(z0, z1)
    np.testing.assert_array_equal(a0, a1)
    np.testing.assert_array_equal(s0.z["N"], s1.z["N"])
    np.testing.assert_array_equal(s0.rng_key, s1.rng_key)
    assert s0.hmc_state.adapt_state.step_size == s1.hmc_state.adapt_state.step_size


def test_tier4_posterior_matches_full_data_nuts():
    from numpyro_b200 import models
    from numpyro_b200.diagnostics import split_gelman_rubin
    from numpyro_b200.infer import HMCECS, MCMC, NUTS
    from numpyro_b200.random import PRNGKey

    X, y, th = _problem(n=20000, D=8, seed=1)
    ref = _mode(X, y).astype(np.float32)
    full = MCMC(NUTS(models.logistic_regression), num_warmup=400, num_samples=1500, num_chains=2)
    full.run(PRNGKey(0), X, y)
    a = full.get_samples(group_by_chain=True)["coefs"]
    kern = HMCECS(NUTS(models.logistic_regression), num_blocks=20, proxy=HMCECS.taylor_proxy({"coefs": ref}))
    ecs = MCMC(kern, num_warmup=400, num_samples=1500, num_chains=2, chain_method="vectorized")
    ecs.run(PRNGKey(1), X, y, subsample_size=500, extra_fields=("accept_prob", "hmc_state.num_steps"))
    b = ecs.get_samples(group_by_chain=True)["coefs"]
    assert set(ecs.get_samples()) == {"coefs"}  # the subsample indices are dropped (hmc_gibbs.py:535-544)
    assert b.shape == (2, 1500, 8)
    sd = a.reshape(-1, 8).std(0)
    assert np.all(np.abs(a.reshape(-1, 8).mean(0) - b.reshape(-1, 8).mean(0)) < 0.25 * sd)
    assert np.all(np.abs(b.reshape(-1, 8).std(0) / sd - 1) < 0.25)
    assert float(np.max(split_gelman_rubin(b.astype(np.float64)))) < 1.05
    acc = ecs.get_extra_fields()["accept_prob"]
    assert 0.3 < float(np.mean(acc)) <= 1.0
