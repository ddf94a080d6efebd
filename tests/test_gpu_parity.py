"""GPU parity tests: the CUDA path (through the C ABI / Python mirror) against the oracle.

Tier 1  Threefry bits / split / uniform: bit-exact.  normal: <= 2 ulp (libm log1p).
Tier 2  potential and gradient: rtol 1e-5 against the float64 oracle.
Tier 3  fixed step size and mass matrix, 50 transitions: identical tree sizes,
        divergences and (hence) accept decisions; positions to float32 round-off growth.
Tier 4  posterior moments within 4 MCSE, r_hat < 1.01 (tests/test_gpu_posterior.py).
"""
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


Y8 = np.array([28.0, 8.0, -3.0, 7.0, -1.0, 1.0, 18.0, 12.0])
S8 = np.array([15.0, 10.0, 16.0, 11.0, 9.0, 11.0, 10.0, 18.0])


def make_logistic(N, D, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(N, D)).astype(np.float32)
    if D > 1:
        X[:, -1] = 1.0
    w = rng.normal(size=D) * 0.5
    y = rng.random(N) < 1 / (1 + np.exp(-X @ w))
    return X, y


def make_sv(Tn, seed=21):
    rng = np.random.default_rng(seed)
    s = np.cumsum(rng.normal(size=Tn) * 0.02) - 4.5
    return (np.exp(s) * rng.standard_t(10, size=Tn)).astype(np.float32)


# ---------------------------------------------------------------- tier 1
def test_threefry_kat(nb):
    keys = [(0, 0), (0xFFFFFFFF, 0xFFFFFFFF), (0x13198A2E, 0x03707344)]
    ctrs = [(0, 0), (0xFFFFFFFF, 0xFFFFFFFF), (0x243F6A88, 0x85A308D3)]
    out = nb.random.threefry2x32(np.array(keys, np.uint32), np.array(ctrs, np.uint32))
    assert out.tolist() == [[0x6B200159, 0x99BA4EFE], [0x1CB996FC, 0xBB002BE7],
                            [0xC4923A9C, 0x483DF7A0]]
    rng = np.random.default_rng(3)
    k = rng.integers(0, 2**32, size=(1000, 2), dtype=np.uint32)
    c = rng.integers(0, 2**32, size=(1000, 2), dtype=np.uint32)
    o0, o1 = T.threefry2x32(k[:, 0], k[:, 1], c[:, 0], c[:, 1])
    assert np.array_equal(nb.random.threefry2x32(k, c), np.stack([o0, o1], -1))


@pytest.mark.parametrize("part", [False, True])
def test_random_bit_exact(nb, part):
    R = nb.random
    for seed in (0, 1, 42, 2**31 + 5, 2**40 + 3):
        key = R.PRNGKey(seed)
        assert np.array_equal(key, T.PRNGKey(seed))
        for num in (1, 2, 3, 4, 7, 16):
            assert np.array_equal(R.split(key, num, part), T.split(key, num, part))
        for n in (1, 2, 3, 10, 55, 3002):
            assert np.array_equal(R.bits(key, (n,), part), T.random_bits(key, (n,), part))
            assert np.array_equal(R.uniform(key, (n,), -2.0, 2.0, part),
                                  T.uniform(key, (n,), -2.0, 2.0, part))
            assert np.array_equal(R.uniform(key, (n,), partitionable=part),
                                  T.uniform(key, (n,), partitionable=part))
            a, b = R.normal(key, (n,), part), T.normal(key, (n,), part)
            assert_allclose(a, b, rtol=3e-7, atol=1e-7)
    batch = T.split(T.PRNGKey(9), 5, part)
    assert np.array_equal(R.split(batch, 3, part), np.stack([T.split(k, 3, part) for k in batch]))


def test_documented_jax_values(nb):
    R = nb.random
    assert R.split(R.PRNGKey(0), 2, False).tolist() == [[4146024105, 967050713],
                                                        [2718843009, 1272950319]]
    assert R.split(R.PRNGKey(0), 2, True).tolist() == [[1797259609, 2579123966],
                                                       [928981903, 3453687069]]
    assert_allclose(R.normal(R.PRNGKey(42), (), False), -0.18471177, atol=2e-8)
    assert_allclose(R.normal(R.PRNGKey(42), (), True), -0.028304616, atol=4e-9)


# ---------------------------------------------------------------- tier 2
def _kernel_for(nb, model, args, kwargs=None, chains=1, **kw):
    k = nb.infer.NUTS(model, **kw)
    keys = nb.random.PRNGKey(0) if chains == 1 else T.split(T.PRNGKey(0), chains, True)
    k.init(keys, 0, None, args, kwargs or {})
    return k


@pytest.mark.parametrize("N,D", [(1, 1), (31, 3), (32, 3), (1000, 8), (3000, 3), (4133, 55),
                                 (70001, 55), (5000, 100), (2048, 128)])
def test_logistic_potential_and_grad(nb, N, D):
    X, y = make_logistic(N, D)
    k = _kernel_for(nb, nb.models.logistic_regression, (X, y))
    fam = opot.Logistic(X, y, np.float64)
    rng = np.random.default_rng(1)
    zs = rng.normal(size=(3, D)) * np.array([0.1, 1.0, 3.0])[:, None]
    U, g = k.potential_and_grad(zs)
    for i in range(3):
        U0, g0 = fam.potential_and_grad(zs[i].astype(np.float32).astype(np.float64))
        assert_allclose(U[i], U0, rtol=1e-5)
        assert_allclose(g[i], g0, rtol=1e-5, atol=1e-5 * np.abs(g0).max())


def test_eight_schools_potential_and_grad(nb):
    k = _kernel_for(nb, nb.models.eight_schools_noncentered, (8, S8), {"y": Y8})
    fam = opot.EightSchools(S8, Y8, np.float64)
    zs = np.random.default_rng(2).normal(size=(64, 10)).astype(np.float32)
    U, g = k.potential_and_grad(zs)
    for i in range(64):
        U0, g0 = fam.potential_and_grad(zs[i].astype(np.float64))
        assert_allclose(U[i], U0, rtol=1e-5)
        assert_allclose(g[i], g0, rtol=1e-5, atol=1e-5 * np.abs(g0).max())


@pytest.mark.parametrize("Tn", [5, 30, 510, 3000])
def test_sv_potential_and_grad(nb, Tn):
    r = make_sv(Tn)
    k = _kernel_for(nb, nb.models.stochastic_volatility, (r,))
    fam = opot.StochasticVolatility(r, np.float64)
    rng = np.random.default_rng(3)
    for nu_u in (-1.0, 1.0, 3.0):
        z = np.concatenate([[nu_u], rng.normal(size=Tn) * 0.1 - 4.5, [-3.5]]).astype(np.float32)
        U, g = k.potential_and_grad(z[None])
        U0, g0 = fam.potential_and_grad(z.astype(np.float64))
        assert_allclose(U[0], U0, rtol=1e-5)
        assert_allclose(g[0], g0, rtol=2e-5, atol=2e-5 * np.abs(g0).max())


# ---------------------------------------------------------------- tier 3
def _tier3(nb, model, args, kwargs, fam32, n=50, part=True, algo="NUTS", chains=1, ztol=5e-3, **kw):
    cls = nb.infer.NUTS if algo == "NUTS" else nb.infer.HMC
    kern = cls(model, adapt_step_size=False, adapt_mass_matrix=False, **kw)
    kern.partitionable = part
    mc = nb.infer.MCMC(kern, num_warmup=0, num_samples=n, num_chains=chains,
                       chain_method="vectorized")
    key = T.PRNGKey(11)
    mc.run(key, *args, extra_fields=("num_steps", "accept_prob", "potential_energy", "energy",
                                     f"z.{fam32.trace_sites[0][0]}"), **kwargs)
    ex = mc.get_extra_fields(group_by_chain=True)
    res, states, _ = ohmc.run_mcmc(fam32, key, 0, n, num_chains=chains, cfg=Config(np.float32, part),
                                   algo=algo, adapt_step_size=False, adapt_mass_matrix=False, **kw)
    assert np.array_equal(ex["num_steps"], res["num_steps"])
    assert np.array_equal(ex["diverging"], res["diverging"])
    assert_allclose(ex["accept_prob"], res["accept_prob"], rtol=2e-3, atol=2e-4)
    assert_allclose(ex["potential_energy"], res["potential_energy"], rtol=1e-4, atol=1e-3)
    z_dev = kern._bound.ravel({k: v for k, v in kern._last_state.z.items()})
    z_ora = np.stack([s.z for s in states])
    assert_allclose(z_dev.reshape(z_ora.shape), z_ora, rtol=ztol, atol=ztol)
    return ex, res


@pytest.mark.parametrize("part", [True, False])
def test_tier3_logistic(nb, part):
    X, y = make_logistic(3000, 3, seed=5)
    ex, _ = _tier3(nb, nb.models.logistic_regression, (X, y), {}, opot.Logistic(X, y), part=part,
                   step_size=0.01)
    assert ex["num_steps"].max() > 3


def test_tier3_logistic_d55(nb):
    X, y = make_logistic(20000, 55, seed=6)
    _tier3(nb, nb.models.logistic_regression, (X, y), {}, opot.Logistic(X, y), step_size=0.01,
           n=30)


def test_tier3_logistic_divergent(nb):
    # test/infer/test_mcmc.py:488-511: a huge fixed step size makes every transition diverge
    X, y = make_logistic(3000, 3, seed=5)
    ex, _ = _tier3(nb, nb.models.logistic_regression, (X, y), {}, opot.Logistic(X, y), n=20,
                   step_size=10.0)
    assert ex["diverging"].all()


@pytest.mark.parametrize("part", [True, False])
def test_tier3_eight_schools(nb, part):
    _tier3(nb, nb.models.eight_schools_noncentered, (8, S8), {"y": Y8},
           opot.EightSchools(S8, Y8), part=part, chains=4, step_size=0.3)


def test_tier3_eight_schools_max_depth(nb):
    ex, _ = _tier3(nb, nb.models.eight_schools_noncentered, (8, S8), {"y": Y8},
                   opot.EightSchools(S8, Y8), chains=2, step_size=0.01, max_tree_depth=4)
    assert ex["num_steps"].max() == 15


def test_tier3_sv(nb):
    r = make_sv(100)
    _tier3(nb, nb.models.stochastic_volatility, (r,), {}, opot.StochasticVolatility(r), n=30,
           chains=2, step_size=0.01, max_tree_depth=6)


def test_tier3_hmc(nb):
    X, y = make_logistic(2000, 3, seed=7)
    _tier3(nb, nb.models.logistic_regression, (X, y), {}, opot.Logistic(X, y), algo="HMC",
           step_size=0.05, trajectory_length=0.5)
    _tier3(nb, nb.models.eight_schools_noncentered, (8, S8), {"y": Y8}, opot.EightSchools(S8, Y8),
           algo="HMC", chains=3, step_size=0.2, num_steps=7, trajectory_length=None)


# ---------------------------------------------------------------- init + warm-up adaptation
@pytest.mark.parametrize("part", [True, False])
def test_init_matches_oracle(nb, part):
    r = make_sv(40)
    for model, args, kwargs, fam in [
        (nb.models.eight_schools_noncentered, (8, S8), {"y": Y8}, opot.EightSchools(S8, Y8)),
        (nb.models.stochastic_volatility, (r,), {}, opot.StochasticVolatility(r)),
    ]:
        kern = nb.infer.NUTS(model)
        kern.partitionable = part
        keys = T.split(T.PRNGKey(3), 3, part)
        st = kern.init(keys, 100, None, args, kwargs)
        for c in range(3):
            ok = ohmc.HMCKernel(fam, cfg=Config(np.float32, part))
            so = ok.init(keys[c], 100)
            z = kern._bound.ravel({k: v[c] for k, v in st.z.items()})
            assert np.array_equal(z, so.z)  # uniform draws are exact
            assert np.array_equal(st.rng_key[c], so.rng_key)
            assert np.array_equal(st.adapt_state.rng_key[c], so.adapt_state.rng_key)
            assert_allclose(st.potential_energy[c], so.potential_energy, rtol=1e-5)
            assert_allclose(st.energy[c], so.energy, rtol=1e-5)


def test_warmup_adaptation_trace(nb):
    # dual averaging + Welford windows through a short warm-up, against the oracle
    fam = opot.EightSchools(S8, Y8)
    kern = nb.infer.NUTS(nb.models.eight_schools_noncentered)
    mc = nb.infer.MCMC(kern, num_warmup=60, num_samples=0)
    key = T.PRNGKey(5)
    mc.warmup(key, 8, S8, y=Y8, collect_warmup=True,
              extra_fields=("adapt_state.step_size", "num_steps"))
    ss = mc.get_extra_fields()["adapt_state.step_size"]
    res, st = ohmc.run_chain(ohmc.HMCKernel(fam, cfg=Config(np.float32, True)), key, 60, 0,
                             collect_warmup=True)
    m = 25  # trajectories are chaotic: compare the part before round-off has grown
    assert np.array_equal(mc.get_extra_fields()["num_steps"][:m], res["num_steps"][:m])
    assert_allclose(ss[:12], res["step_size"][:12], rtol=1e-4)
    assert_allclose(ss[:m], res["step_size"][:m], rtol=1e-2)
    last = mc.last_state
    assert int(last.adapt_state.window_idx) == st.adapt_state.window_idx == 3
    assert int(last.i) == 60
    imm = next(iter(last.adapt_state.inverse_mass_matrix.values()))
    assert imm.shape == (10,) and (imm > 0).all() and not np.allclose(imm, 1.0)


# ---------------------------------------------------------------- driver semantics
def test_thinning_and_collection(nb):
    fam = opot.EightSchools(S8, Y8)
    kw = dict(step_size=0.3, adapt_step_size=False, adapt_mass_matrix=False)
    key = T.PRNGKey(2)
    for thin, ns in ((3, 20), (4, 21)):
        mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.eight_schools_noncentered, **kw), num_warmup=5,
                           num_samples=ns, thinning=thin)
        mc.run(key, 8, S8, y=Y8, extra_fields=("num_steps",))
        res, _ = ohmc.run_chain(ohmc.HMCKernel(fam, cfg=Config(np.float32, True), **kw), key, 5, ns,
                                thinning=thin)
        assert mc.get_samples()["mu"].shape == (ns // thin,)
        assert np.array_equal(mc.get_extra_fields()["num_steps"], res["num_steps"])


def test_sample_by_sample_equals_fused_loop(nb):
    kw = dict(step_size=0.3)
    key = T.PRNGKey(4)
    k1 = nb.infer.NUTS(nb.models.eight_schools_noncentered, **kw)
    mc = nb.infer.MCMC(k1, num_warmup=30, num_samples=10)
    mc.run(key, 8, S8, y=Y8)
    k2 = nb.infer.NUTS(nb.models.eight_schools_noncentered, **kw)
    st = k2.init(key, 30, None, (8, S8), {"y": Y8})
    for _ in range(40):
        st = k2.sample(st, (8, S8), {"y": Y8})
    assert int(st.i) == 40
    for site in ("mu", "tau", "theta_base"):
        assert np.array_equal(st.z[site], mc.last_state.z[site])
    assert np.array_equal(st.rng_key, mc.last_state.rng_key)


def test_chain_consistency_and_resume(nb):
    # test/infer/test_mcmc.py:848-894: chain 0 of a 2-chain run == 1-chain run with split(key)[0]
    key = T.PRNGKey(8)
    mk = lambda c: nb.infer.MCMC(nb.infer.NUTS(nb.models.eight_schools_noncentered),  # noqa: E731
                                 num_warmup=40, num_samples=15, num_chains=c,
                                 chain_method="vectorized")
    m2, m1 = mk(2), mk(1)
    m2.run(key, 8, S8, y=Y8)
    m1.run(nb.random.split(key, 2)[0], 8, S8, y=Y8)
    assert np.array_equal(m2.get_samples(group_by_chain=True)["mu"][0],
                          m1.get_samples(group_by_chain=True)["mu"][0])
    # post_warmup_state: warmup() then run() == a single run() with the same keys
    m3 = mk(1)
    k0 = nb.random.split(key, 2)[0]
    m3.warmup(k0, 8, S8, y=Y8)
    assert int(m3.post_warmup_state.i) == 40
    m3.run(m3.post_warmup_state.rng_key, 8, S8, y=Y8)
    assert np.array_equal(m3.get_samples()["mu"], m1.get_samples()["mu"])
    assert m3.get_samples()["theta"].shape == (15, 8)


def test_init_params_and_extra_fields(nb):
    X, y = make_logistic(500, 3, seed=9)
    kern = nb.infer.NUTS(nb.models.logistic_regression, step_size=0.1, adapt_step_size=False,
                         adapt_mass_matrix=False)
    mc = nb.infer.MCMC(kern, num_warmup=0, num_samples=8, num_chains=2, chain_method="vectorized")
    z0 = {"coefs": np.zeros((2, 3), np.float32)}
    mc.run(T.PRNGKey(1), X, y, init_params=z0,
           extra_fields=("adapt_state.step_size", "num_steps", "~z.coefs"))
    assert mc.get_samples() == {}
    ex = mc.get_extra_fields(group_by_chain=True)
    assert ex["adapt_state.step_size"].shape == (2, 8) and ex["diverging"].dtype == bool
    res = []
    for c, k in enumerate(T.split(T.PRNGKey(1), 2, True)):
        ok = ohmc.HMCKernel(opot.Logistic(X, y), cfg=Config(np.float32, True), step_size=0.1,
                            adapt_step_size=False, adapt_mass_matrix=False)
        r, _ = ohmc.run_chain(ok, k, 0, 8, init_params=np.zeros(3, np.float32))
        res.append(r["num_steps"])
    assert np.array_equal(ex["num_steps"], np.stack(res))
    with pytest.raises(ValueError):
        mc.run(T.PRNGKey(1), X, y, init_params={"coefs": np.zeros((3, 3), np.float32)})
