"""GPU parity: dense mass matrix (hmc_util.py:133-239, 439-515, 710-746, 1183-1223),
find_heuristic_step_size (hmc_util.py:314-384) and the horseshoe family
(examples/horseshoe_regression.py:37-78) against the oracle, through the C ABI."""
import numpy as np
from numpy.testing import assert_allclose
import pytest

pytestmark = pytest.mark.gpu

from oracle import hmc as ohmc  # noqa: E402
from oracle import potentials as opot  # noqa: E402
from oracle import threefry as T  # noqa: E402
from oracle.hmc_util import Config  # noqa: E402

Y8 = np.array([28.0, 8.0, -3.0, 7.0, -1.0, 1.0, 18.0, 12.0])
S8 = np.array([15.0, 10.0, 16.0, 11.0, 9.0, 11.0, 10.0, 18.0])


@pytest.fixture(scope="module")
def nb():
    import __graft_entry__ as g

    return g.build()


def make_logistic(N, D, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(N, D)).astype(np.float32)
    X[:, -1] = 1.0
    y = rng.random(N) < 1 / (1 + np.exp(-X @ (rng.normal(size=D) * 0.5)))
    return X, y


def make_horseshoe(N, D, likelihood, seed=0):
    # examples/horseshoe_regression.py:105-125
    rng = np.random.RandomState(seed)
    X = rng.randn(N, D).astype(np.float32)
    W = np.array([2.0, -1.0, 0.50])
    Y = X[:, :3] @ W
    Y = Y - Y.mean()
    if likelihood == "normal":
        Y = Y + 0.05 * rng.randn(N)
    else:
        Y = rng.binomial(1, 1 / (1 + np.exp(-Y))).astype(np.float64)
    return X, Y.astype(np.float32)


def spd(D, seed):
    rng = np.random.default_rng(seed)
    A = rng.normal(size=(D, D)) * 0.3
    return (A @ A.T + np.eye(D) * 0.5).astype(np.float32)


def _run_pair(nb, model, args, kwargs, fam, n, chains, nw=0, collect_warmup=False, **kw):
    kern = nb.infer.NUTS(model, **kw)
    mc = nb.infer.MCMC(kern, num_warmup=nw, num_samples=n, num_chains=chains,
                       chain_method="vectorized")
    key = T.PRNGKey(11)
    fields = ("num_steps", "accept_prob", "potential_energy", "adapt_state.step_size")
    if collect_warmup:
        mc.warmup(key, *args, extra_fields=fields, collect_warmup=True, **kwargs)
    else:
        mc.run(key, *args, extra_fields=fields, **kwargs)
    ex = mc.get_extra_fields(group_by_chain=True)
    keys = T.split(key, chains, True) if chains > 1 else [key]
    res = []
    for k in keys:
        ok = ohmc.HMCKernel(fam, cfg=Config(np.float32, True), **kw)
        r, st = ohmc.run_chain(ok, k, nw, 0 if collect_warmup else n, collect_warmup=collect_warmup)
        res.append((r, st))
    return kern, mc, ex, res


# ---------------------------------------------------------------- dense mass, fixed parameters
@pytest.mark.parametrize("case", ["eight_schools", "logistic"])
def test_tier3_dense_fixed(nb, case):
    if case == "eight_schools":
        model, args, kwargs, fam, D, eps = (nb.models.eight_schools_noncentered, (8, S8),
                                            {"y": Y8}, opot.EightSchools(S8, Y8), 10, 0.2)
    else:
        X, y = make_logistic(3000, 5, seed=5)
        model, args, kwargs, fam, D, eps = (nb.models.logistic_regression, (X, y), {},
                                            opot.Logistic(X, y), 5, 0.03)
    M = spd(D, 3)
    if case == "logistic":
        M = M * 0.2
    kern, mc, ex, res = _run_pair(nb, model, args, kwargs, fam, 40, 2, step_size=eps,
                                  adapt_step_size=False, adapt_mass_matrix=False,
                                  dense_mass=True, inverse_mass_matrix=M)
    ns = np.stack([r["num_steps"] for r, _ in res])
    assert np.array_equal(ex["num_steps"], ns)
    assert ns.max() > 3
    assert_allclose(ex["accept_prob"], np.stack([r["accept_prob"] for r, _ in res]), rtol=5e-3,
                    atol=5e-4)
    st = mc.last_state.adapt_state
    block = next(iter(st.inverse_mass_matrix))
    so = res[0][1].adapt_state
    assert st.inverse_mass_matrix[block].shape == (2, D, D)
    assert_allclose(st.inverse_mass_matrix[block][0], so.inverse_mass_matrix, rtol=1e-6)
    assert_allclose(st.mass_matrix_sqrt[block][0], so.mass_matrix_sqrt, rtol=2e-4, atol=2e-5)
    assert_allclose(st.mass_matrix_sqrt_inv[block][0], so.mass_matrix_sqrt_inv, rtol=2e-4, atol=2e-5)


def test_dense_warmup_adaptation(nb):
    fam = opot.EightSchools(S8, Y8)
    kern, mc, ex, res = _run_pair(nb, nb.models.eight_schools_noncentered, (8, S8), {"y": Y8}, fam,
                                  0, 1, nw=120, collect_warmup=True, dense_mass=True)
    r, st = res[0]
    m = 20  # chaotic afterwards
    assert np.array_equal(ex["num_steps"][0][:m], r["num_steps"][:m])
    assert_allclose(ex["adapt_state.step_size"][0][:12], r["step_size"][:12], rtol=1e-4)
    last = mc.last_state
    imm = next(iter(last.adapt_state.inverse_mass_matrix.values()))
    msq = next(iter(last.adapt_state.mass_matrix_sqrt.values()))
    msqi = next(iter(last.adapt_state.mass_matrix_sqrt_inv.values()))
    assert imm.shape == (10, 10) and not np.allclose(imm, np.eye(10))
    assert int(last.adapt_state.window_idx) == st.adapt_state.window_idx
    # the factors belong to the adapted covariance: M^{1/2} M^{1/2}^T = (M^-1)^-1
    assert_allclose(msq @ msq.T @ imm, np.eye(10), atol=2e-3)
    assert_allclose(msqi @ msq, np.eye(10), atol=1e-3)
    assert np.allclose(np.triu(msq, 1), 0) and np.allclose(np.triu(msqi, 1), 0)


def test_dense_posterior_eight_schools(nb):
    mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.eight_schools_noncentered, dense_mass=True),
                       num_warmup=400, num_samples=400, num_chains=16, chain_method="vectorized")
    mc.run(T.PRNGKey(0), 8, S8, y=Y8)
    s = mc.get_samples(group_by_chain=True)
    assert nb.diagnostics.split_gelman_rubin(np.asarray(s["mu"], np.float64)) < 1.02
    assert abs(s["mu"].mean() - 4.3) < 0.6 and abs(s["tau"].mean() - 3.6) < 0.8


# ---------------------------------------------------------------- step-size heuristic
@pytest.mark.parametrize("dense", [False, True])
def test_find_heuristic_step_size(nb, dense):
    X, y = make_logistic(2000, 4, seed=2)
    for model, args, kwargs, fam in [
        (nb.models.eight_schools_noncentered, (8, S8), {"y": Y8}, opot.EightSchools(S8, Y8)),
        (nb.models.logistic_regression, (X, y), {}, opot.Logistic(X, y)),
    ]:
        kw = dict(find_heuristic_step_size=True, dense_mass=dense)
        kern = nb.infer.NUTS(model, **kw)
        keys = T.split(T.PRNGKey(4), 3, True)
        st = kern.init(keys, 100, None, args, kwargs)
        for c in range(3):
            so = ohmc.HMCKernel(fam, cfg=Config(np.float32, True), **kw).init(keys[c], 100)
            assert st.adapt_state.step_size[c] == so.adapt_state.step_size, (c, dense)
            assert np.array_equal(st.rng_key[c], so.rng_key)
            assert_allclose(st.energy[c], so.energy, rtol=1e-5)
    # through a warm-up with window ends (search re-run at every middle window end)
    kern, mc, ex, res = _run_pair(nb, nb.models.eight_schools_noncentered, (8, S8), {"y": Y8},
                                  opot.EightSchools(S8, Y8), 0, 1, nw=110, collect_warmup=True,
                                  find_heuristic_step_size=True, dense_mass=dense)
    r, st = res[0]
    assert np.array_equal(ex["num_steps"][0][:15], r["num_steps"][:15])
    assert int(mc.last_state.i) == 110


# ---------------------------------------------------------------- horseshoe family
@pytest.mark.parametrize("likelihood", ["bernoulli", "normal"])
@pytest.mark.parametrize("N,D", [(50, 3), (200, 20), (300, 300)])
def test_horseshoe_potential_and_grad(nb, likelihood, N, D):
    X, Y = make_horseshoe(N, D, likelihood)
    model = nb.models.horseshoe_bernoulli if likelihood == "bernoulli" else nb.models.horseshoe_normal
    k = nb.infer.NUTS(model)
    k.init(nb.random.PRNGKey(0), 0, None, (X, Y), {})
    fam = opot.Horseshoe(X, Y, likelihood, np.float64)
    assert k._bound.dim == fam.dim
    rng = np.random.default_rng(1)
    zs = (rng.normal(size=(4, fam.dim)) * 0.5).astype(np.float32)
    U, g = k.potential_and_grad(zs)
    for i in range(4):
        U0, g0 = fam.potential_and_grad(zs[i].astype(np.float64))
        assert_allclose(U[i], U0, rtol=2e-5)
        assert_allclose(g[i], g0, rtol=1e-4, atol=1e-4 * np.abs(g0).max())


@pytest.mark.parametrize("likelihood", ["bernoulli", "normal"])
def test_tier3_horseshoe(nb, likelihood):
    X, Y = make_horseshoe(100, 5, likelihood)
    model = nb.models.horseshoe_bernoulli if likelihood == "bernoulli" else nb.models.horseshoe_normal
    fam = opot.Horseshoe(X, Y, likelihood)
    kern, mc, ex, res = _run_pair(nb, model, (X, Y), {}, fam, 30, 2, step_size=0.02,
                                  adapt_step_size=False, adapt_mass_matrix=False, max_tree_depth=6)
    assert np.array_equal(ex["num_steps"], np.stack([r["num_steps"] for r, _ in res]))
    assert_allclose(ex["potential_energy"], np.stack([r["potential_energy"] for r, _ in res]),
                    rtol=5e-4, atol=2e-3)
    s = mc.get_samples()
    assert s["betas"].shape == (60, 5) and (s["lambdas"] > 0).all()


def test_horseshoe_init_and_short_run(nb):
    # trace order (init draws) differs from the flat sorted order for the normal variant
    X, Y = make_horseshoe(120, 6, "normal")
    fam = opot.Horseshoe(X, Y, "normal")
    kern = nb.infer.NUTS(nb.models.horseshoe_normal)
    keys = T.split(T.PRNGKey(3), 2, True)
    st = kern.init(keys, 50, None, (X, Y), {})
    for c in range(2):
        so = ohmc.HMCKernel(fam, cfg=Config(np.float32, True)).init(keys[c], 50)
        z = kern._bound.ravel({k: v[c] for k, v in st.z.items()})
        assert np.array_equal(z, so.z)
        assert_allclose(st.potential_energy[c], so.potential_energy, rtol=1e-5)
    mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.horseshoe_normal, dense_mass=True), num_warmup=150,
                       num_samples=100, num_chains=4, chain_method="vectorized")
    mc.run(T.PRNGKey(1), X, Y)
    b = mc.get_samples()["betas"].mean(0)
    assert_allclose(b[:3], [2.0, -1.0, 0.5], atol=0.15)
