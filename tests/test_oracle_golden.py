"""Pin the oracle against every golden vector / known answer the reference's own
tests hold for the hot path, plus the Random123 and JAX-documentation
known answers for Threefry.  CPU only."""
import numpy as np
from numpy.testing import assert_allclose
import pytest

from oracle import diagnostics as D
from oracle import hmc_util as hu
from oracle import threefry as T

CFG = hu.Config(np.float32, partitionable=False)


# ---------------------------------------------------------------- Threefry
@pytest.mark.parametrize(
    "key, ctr, expected",
    [  # Random123 kat_vectors, threefry2x32 20 rounds (also in jax/tests/random_test.py)
        ((0x0, 0x0), (0x0, 0x0), (0x6B200159, 0x99BA4EFE)),
        ((0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF, 0xFFFFFFFF), (0x1CB996FC, 0xBB002BE7)),
        ((0x13198A2E, 0x03707344), (0x243F6A88, 0x85A308D3), (0xC4923A9C, 0x483DF7A0)),
    ],
)
def test_threefry_kat(key, ctr, expected):
    o0, o1 = T.threefry2x32(key[0], key[1], ctr[0], ctr[1])
    assert (int(o0), int(o1)) == expected


def test_jax_documented_values_legacy_layout():
    # values printed in JAX's README / "JAX 101" docs (jax_threefry_partitionable=False)
    assert T.split(T.PRNGKey(0)).tolist() == [[4146024105, 967050713], [2718843009, 1272950319]]
    assert T.split(T.PRNGKey(42)).tolist() == [[2465931498, 3679230171], [255383827, 267815257]]
    assert_allclose(T.normal(T.PRNGKey(42)), -0.18471177, rtol=0, atol=1e-8)
    assert_allclose(T.normal(T.PRNGKey(0), (3,)), [1.8160863, -0.48262316, 0.33988908],
                    rtol=0, atol=1.5e-7)
    assert_allclose(T.uniform(T.PRNGKey(0)), 0.41845703, rtol=0, atol=1e-8)


def test_jax_documented_values_partitionable_layout():
    # values printed in JAX >= 0.5 docs (jax_threefry_partitionable=True)
    assert T.split(T.PRNGKey(0), partitionable=True).tolist() == [
        [1797259609, 2579123966], [928981903, 3453687069]]
    assert_allclose(T.normal(T.PRNGKey(42), partitionable=True), -0.028304616, rtol=0, atol=1e-9)


def test_random_layout_invariants():
    key = T.PRNGKey(7)
    for part in (False, True):
        # scalar and (1,) draws agree; odd sizes are a prefix-consistent function of size
        assert T.random_bits(key, (), part) == T.random_bits(key, (1,), part)[0]
        u = T.uniform(key, (55,), -2, 2, part)
        assert u.dtype == np.float32 and (u >= -2).all() and (u < 2).all()
        n = T.normal(key, (3002,), part)
        assert np.isfinite(n).all() and abs(float(n.mean())) < 0.1
    # erf_inv against scipy
    from scipy.special import erfinv

    x = np.linspace(-0.9999, 0.9999, 20001).astype(np.float32)
    assert_allclose(T.erf_inv_f32(x), erfinv(x.astype(np.float64)), rtol=1e-5, atol=1e-7)
    assert np.isinf(T.erf_inv_f32(np.float32(1.0)))


# ---------------------------------------------------------------- hmc_util golden vectors
@pytest.mark.parametrize(
    "num_steps, expected",
    [  # test/infer/test_hmc_util.py:278-292
        (18, [(0, 17)]),
        (50, [(0, 6), (7, 44), (45, 49)]),
        (100, [(0, 14), (15, 89), (90, 99)]),
        (150, [(0, 74), (75, 99), (100, 149)]),
        (200, [(0, 74), (75, 99), (100, 149), (150, 199)]),
        (280, [(0, 74), (75, 99), (100, 229), (230, 279)]),
    ],
)
def test_build_adaptation_schedule(num_steps, expected):
    assert hu.build_adaptation_schedule(num_steps) == [hu.AdaptWindow(i, j) for i, j in expected]


@pytest.mark.parametrize(
    "leaf_idx, ckpt_idxs",  # test/infer/test_hmc_util.py:381-386
    [(0, (1, 0)), (6, (3, 2)), (7, (0, 2)), (13, (2, 2)), (15, (0, 3))],
)
def test_leaf_idx_to_ckpt_idx(leaf_idx, ckpt_idxs):
    assert hu.leaf_idx_to_ckpt_idxs(leaf_idx) == ckpt_idxs


@pytest.mark.parametrize(
    "ckpt_idxs, expected_turning",  # test/infer/test_hmc_util.py:389-403
    [((3, 2), False), ((3, 3), True), ((0, 0), False), ((0, 1), True), ((1, 3), True)],
)
def test_is_iterative_turning(ckpt_idxs, expected_turning):
    imm = np.ones(1, np.float32)
    r = np.array([1.0], np.float32)
    r_sum = np.array([3.0], np.float32)
    r_ckpts = np.array([1.0, 2.0, 3.0, -2.0], np.float32)[:, None]
    r_sum_ckpts = np.array([2.0, 4.0, 4.0, -1.0], np.float32)[:, None]
    assert hu.is_iterative_turning(CFG, imm, r, r_sum, r_ckpts, r_sum_ckpts, *ckpt_idxs) \
        == expected_turning


def test_dual_averaging():
    # test/infer/test_hmc_util.py:36-52: minimise (x+1)^2 with gamma=0.5
    da_init, da_update = hu.dual_averaging(CFG, gamma=0.5)
    state = da_init()
    for _ in range(10):
        g = 2 * (state[0] + 1)
        state = da_update(g, state)
    assert_allclose(state[1], -1.0, atol=1e-3)


@pytest.mark.parametrize("diagonal", [True, False])
@pytest.mark.parametrize("regularize", [True, False])
def test_welford_covariance(diagonal, regularize):
    # test/infer/test_hmc_util.py:55-89
    np.random.seed(0)
    loc = np.random.randn(3)
    a = np.random.randn(3, 3)
    target_cov = a @ a.T
    x = np.random.multivariate_normal(loc, target_cov, size=(2000,)).astype(np.float32)
    wc_init, wc_update, wc_final = hu.welford_covariance(CFG, diagonal=diagonal)
    state = wc_init(3)
    for i in range(2000):
        state = wc_update(x[i], state)
    cov, cov_inv_sqrt, _ = wc_final(state, regularize=regularize)
    if diagonal:
        assert_allclose(cov, np.diagonal(target_cov), rtol=0.06)
        assert_allclose(cov_inv_sqrt, np.sqrt(1 / np.diagonal(target_cov)), rtol=0.06)
    else:
        assert_allclose(cov, target_cov, rtol=0.06)
        assert_allclose(cov_inv_sqrt, np.linalg.cholesky(np.linalg.inv(cov)), rtol=0.06, atol=1e-4)


VV_CASES = [  # test/infer/test_hmc_util.py:121-191
    ("harmonic", 0.01, 100, [0.0], [1.0], [np.sin(1.0)], [np.cos(1.0)], 1e-4,
     lambda q: (0.5 * q[0] ** 2, q.copy())),
    ("circular", 0.01, 628, [1.0, 0.0], [0.0, 1.0], [1.0, 0.0], [0.0, 1.0], 5e-3,
     lambda q: (-1.0 / np.sqrt(q @ q), q / (q @ q) ** 1.5)),
    ("quartic", 0.1, 1810, [0.02], [0.0], [-0.02], [0.0], 1e-4,
     lambda q: (0.25 * q[0] ** 4, q ** 3)),
]


@pytest.mark.parametrize("case", VV_CASES, ids=[c[0] for c in VV_CASES])
def test_velocity_verlet(case):
    _, eps, n, q_i, p_i, q_f, p_f, prec, pot = case
    cfg = hu.Config(np.float32)
    fn = lambda q: tuple(np.float32(v) for v in pot(q))  # noqa: E731
    m_inv = np.ones(len(q_i), np.float32)
    vv_init, vv_update = hu.velocity_verlet(cfg, fn)

    def run(q, p):
        st = vv_init(np.array(q, np.float32), np.array(p, np.float32))
        for _ in range(n):
            st = vv_update(eps, m_inv, st)
        return st.z, st.r

    q, p = run(q_i, p_i)
    assert_allclose(q, q_f, atol=prec)
    assert_allclose(p, p_f, atol=prec)
    e0 = 0.5 * np.sum(np.square(p_i)) + pot(np.array(q_i))[0]
    e1 = 0.5 * np.sum(np.square(p)) + pot(q)[0]
    assert_allclose(e0, e1, atol=1e-5)
    q_back, _ = run(q, -p)
    assert_allclose(q_back, q_i, atol=1e-4)


def test_warmup_adapter():
    # test/infer/test_hmc_util.py:304-378
    def find_reasonable_step_size(step_size, m_inv, z, rng_key):
        return step_size * 4 if step_size < 1 else step_size / 4

    num_steps = 150
    schedule = hu.build_adaptation_schedule(num_steps)
    wa_init, wa_update = hu.warmup_adapter(CFG, num_steps, find_reasonable_step_size)
    z = np.ones(3, np.float32)
    st = wa_init((z, None, None, None), T.PRNGKey(0), 1.0)
    assert st.step_size == find_reasonable_step_size(1.0, None, None, None)
    assert_allclose(st.inverse_mass_matrix, np.ones(3))
    assert st.window_idx == 0

    w = schedule[0]
    step_size = st.step_size
    for t in range(w.start, w.end + 1):
        st = wa_update(t, 0.7 + 0.1 * t / (w.end - w.start), (z,), st)
    assert st.window_idx == 1
    assert st.step_size < step_size
    assert_allclose(st.inverse_mass_matrix, np.ones(3))

    w = schedule[1]
    wl = w.end - w.start
    step_size = st.step_size
    for t in range(w.start, w.end + 1):
        st = wa_update(t, 0.8 + 0.1 * (t - w.start) / wl, (2 * z,), st)
    assert st.window_idx == 2
    assert st.step_size > step_size
    reg = 1e-3 * (5 / (w.end + 1 - w.start + 5))
    assert_allclose(st.inverse_mass_matrix, np.full(3, reg), atol=1e-7)
    imm = st.inverse_mass_matrix

    w = schedule[2]
    step_size = st.step_size
    for t in range(w.start, w.end + 1):
        st = wa_update(t, 0.8, (t * z,), st)
    assert st.window_idx == 3
    assert_allclose(st.step_size, step_size * 10, atol=1e-6)
    assert_allclose(st.inverse_mass_matrix, imm)


@pytest.mark.parametrize("step_size", [0.01, 1.0, 100.0])
@pytest.mark.parametrize("partitionable", [False, True])
def test_build_tree(step_size, partitionable):
    # test/infer/test_hmc_util.py:406-442
    cfg = hu.Config(np.float32, partitionable)
    fn = lambda q: (np.float32(0.5 * q[0] ** 2), q.copy())  # noqa: E731
    vv_init, vv_update = hu.velocity_verlet(cfg, fn)
    st = vv_init(np.array([0.0], np.float32), np.array([1.0], np.float32))
    tree = hu.build_tree(cfg, vv_update, st, np.ones(1, np.float32), step_size, T.PRNGKey(0))
    assert tree.num_proposals >= 2 ** (tree.depth - 1)
    assert tree.sum_accept_probs <= tree.num_proposals
    if tree.depth < 10:
        assert tree.turning or tree.diverging
    if step_size > 10:
        assert tree.diverging and tree.num_proposals == 1
    if step_size < 0.1:
        assert tree.num_proposals > 10


# ---------------------------------------------------------------- diagnostics known answers
def test_autocorrelation():
    # test/test_diagnostics.py:60-69
    x = np.arange(10.0)
    expected = np.array([1, 0.78, 0.52, 0.21, -0.13, -0.52, -0.94, -1.4, -1.91, -2.45])
    assert_allclose(D.autocorrelation(x, bias=False), expected, atol=0.01)
    assert_allclose(D.autocorrelation(x, bias=True), expected * np.arange(10, 0.0, -1) / 10,
                    atol=0.01)


def test_autocovariance():
    # test/test_diagnostics.py:79-85
    expected = np.array([8.25, 6.42, 4.25, 1.75, -1.08, -4.25, -7.75, -11.58, -15.75, -20.25])
    assert_allclose(D.autocovariance(np.arange(10.0), bias=False), expected, atol=0.01)


def test_gelman_rubin():
    # test/test_diagnostics.py:88-95
    x = np.empty((2, 10))
    x[0, :] = np.arange(10.0)
    x[1, :] = np.arange(10.0) + 1
    assert_allclose(D.gelman_rubin(x), 0.98, atol=0.01)
    y = np.random.default_rng(0).normal(size=(2, 10))
    assert_allclose(D.gelman_rubin(y.reshape(2, 2, 5).reshape(4, 5)), D.split_gelman_rubin(y))


def test_effective_sample_size():
    # test/test_diagnostics.py:105-107
    x = np.arange(1000.0).reshape(100, 10)
    assert_allclose(D.effective_sample_size(x, bias=False), 52.64, atol=0.01)


@pytest.mark.parametrize("target", [433, 124, 25, 300, 1, 3, 7])
def test_fft_next_fast_len(target):
    from scipy.fftpack import next_fast_len

    assert D._fft_next_fast_len(target) == next_fast_len(target)


def test_bulk_ess_iid():
    x = np.random.default_rng(1).normal(size=(4, 1000, 2))
    ess = D.bulk_ess(x)
    assert ess.shape == (2,) and (ess > 3000).all() and (ess < 5000).all()
