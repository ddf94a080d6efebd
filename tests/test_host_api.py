"""CPU-only checks of the host side: the C-ABI library loads and exports every symbol the
header declares, host-only entry points work, and the kernel classes keep the reference's
argument checking (hmc.py:643-658, mcmc.py:342-354).  No compute calls."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def nb():
    import __graft_entry__ as g

    return g.build()


def test_library_exports_every_declared_symbol(nb):
    from numpyro_b200 import _lib

    header = open(os.path.join(ROOT, "include", "b200mc.h")).read()
    declared = set(re.findall(r"\b(b200mc_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    lib = _lib.lib()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in b200mc.h but not exported"
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    m = re.search(r"#define\s+B200MC_ABI_VERSION\s+(\d+)", header)
    assert lib.b200mc_version() == int(m.group(1)) == _lib.ABI_VERSION
    # the ctypes mirror of b200mc_config starts with struct_size and matches the C layout
    assert _lib.Config._fields_[0][0] == "struct_size"
    assert C.sizeof(_lib.ChainScalars) == _lib.SCALARS_DTYPE.itemsize


@pytest.mark.parametrize(
    "num_steps, expected",  # reference golden vectors: test/infer/test_hmc_util.py:278-292
    [(18, [(0, 17)]), (50, [(0, 6), (7, 44), (45, 49)]), (100, [(0, 14), (15, 89), (90, 99)]),
     (150, [(0, 74), (75, 99), (100, 149)]), (200, [(0, 74), (75, 99), (100, 149), (150, 199)]),
     (280, [(0, 74), (75, 99), (100, 229), (230, 279)])],
)
def test_adaptation_schedule_through_c_abi(nb, num_steps, expected):
    from numpyro_b200 import _lib

    s, e = (C.c_int * 40)(), (C.c_int * 40)()
    n = _lib.lib().b200mc_adaptation_schedule(num_steps, s, e, 40)
    assert [(s[i], e[i]) for i in range(n)] == expected


def test_schedule_matches_oracle_everywhere(nb):
    from numpyro_b200 import _lib
    from oracle.hmc_util import build_adaptation_schedule

    s, e = (C.c_int * 40)(), (C.c_int * 40)()
    for num_steps in list(range(1, 400)) + [1000, 5000, 100000]:
        n = _lib.lib().b200mc_adaptation_schedule(num_steps, s, e, 40)
        assert [(s[i], e[i]) for i in range(n)] == [tuple(w) for w in
                                                    build_adaptation_schedule(num_steps)]


def test_constructor_contract(nb):
    from numpyro_b200 import models
    from numpyro_b200.infer import HMC, MCMC, NUTS

    with pytest.raises(ValueError, match="Only one of `model` or `potential_fn`"):
        NUTS()
    with pytest.raises(ValueError, match="Only one of `model` or `potential_fn`"):
        NUTS(models.logistic_regression, potential_fn=lambda z: 0.0)
    with pytest.raises(ValueError, match="num_steps` or `trajectory_length"):
        HMC(models.logistic_regression, trajectory_length=None)
    with pytest.warns(UserWarning, match="step size can't be adapted"):
        HMC(models.logistic_regression, num_steps=3)
    # outside the registered families: raises, never falls back
    with pytest.raises(NotImplementedError):
        NUTS(lambda x: x)
    with pytest.raises(NotImplementedError):
        NUTS(potential_fn=lambda z: 0.0)
    with pytest.raises(ValueError, match="list of tuples"):
        NUTS(models.logistic_regression, dense_mass=["coefs"])
    assert NUTS(models.logistic_regression, dense_mass=[("coefs",)])._dense_mass == [("coefs",)]
    assert NUTS(models.logistic_regression, dense_mass=True, find_heuristic_step_size=True)._dense_mass
    k = NUTS(models.eight_schools_noncentered, step_size=1, max_tree_depth=(5, 10))
    assert k._step_size == 1.0 and isinstance(k._step_size, float)
    assert k.sample_field == "z" and k.default_fields == ("z", "diverging")
    assert k._algo == "NUTS" and k._max_tree_depth == (5, 10) and k.model is models.eight_schools_noncentered
    assert k.postprocess_fn((), {})("x") == "x"  # identity before init (hmc.py:801-804)
    assert not k.is_ensemble_kernel
    with pytest.raises(ValueError, match="thinning"):
        MCMC(k, num_warmup=1, num_samples=1, thinning=0)
    with pytest.raises(ValueError, match="Only supporting"):
        MCMC(k, num_warmup=1, num_samples=1, chain_method="foo")
    with pytest.raises(NotImplementedError):
        models.logistic_regression(np.zeros((2, 2)), np.zeros(2))


def test_structured_mass_matrix_layout(nb):
    """Block order of the reference's dicts (hmc_util.py:439-486): ``inverse_mass_matrix`` =
    user entries, dense blocks, remaining sites; momentum keys (``mass_matrix_sqrt``) = dense
    blocks, user's other blocks, remaining sites; a block keeps its OWN site order."""
    from numpyro_b200 import models
    from numpyro_b200.infer import NUTS

    b = models.eight_schools_noncentered.bind(8, np.ones(8), y=np.ones(8))
    lay = lambda **kw: NUTS(models.eight_schools_noncentered, **kw)._mass_layout(b)  # noqa: E731
    blocks, keys, _ = lay()
    assert [n for n, _, _ in blocks] == [("mu", "tau", "theta_base")] and not blocks[0][2]
    blocks, keys, _ = lay(dense_mass=True)
    assert blocks[0][2] and np.array_equal(blocks[0][1], np.arange(10))
    blocks, keys, _ = lay(dense_mass=[("theta_base", "mu")])
    assert [n for n, _, _ in blocks] == [("theta_base", "mu"), ("tau",)]
    assert np.array_equal(blocks[0][1], [2, 3, 4, 5, 6, 7, 8, 9, 0]) and blocks[0][2]
    assert np.array_equal(blocks[1][1], [1]) and not blocks[1][2] and keys == [0, 1]
    blocks, keys, user = lay(dense_mass=[("tau", "theta_base")],
                             inverse_mass_matrix={("mu",): np.ones(1), ("tau", "theta_base"): np.eye(9)})
    assert [n for n, _, _ in blocks] == [("mu",), ("tau", "theta_base")] and keys == [1, 0]
    with pytest.raises(AssertionError, match="conflict of sites names"):
        lay(dense_mass=[("mu", "tau")], inverse_mass_matrix={("mu",): np.ones(1)})


def test_unconstrain_inverts_postprocess(nb):
    """``init_to_value`` takes CONSTRAINED values: log for the positive sites."""
    from numpyro_b200 import models

    b = models.horseshoe_normal.bind(np.ones((4, 3), np.float32), np.ones(4, np.float32))
    z = {k: np.random.default_rng(1).normal(size=shp).astype(np.float32)
         for k, (_, _, shp) in b.flat_sites.items()}
    c = b.postprocess(z)
    back = b.unconstrain(c)
    for k in z:
        np.testing.assert_allclose(back[k], z[k], rtol=1e-5, atol=1e-6)
    assert set(b.positive) == {"lambdas", "tau", "prec_obs"}


def test_no_device_raises_loudly(nb):
    from numpyro_b200 import _lib, models, random
    from numpyro_b200.infer import NUTS

    if _lib.lib().b200mc_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(_lib.B200Error, match="no CPU fallback"):
        random.split(random.PRNGKey(0))
    with pytest.raises(_lib.B200Error, match="no CPU fallback"):
        NUTS(models.eight_schools_noncentered).init(random.PRNGKey(0), 10, None,
                                                    (8, np.ones(8)), {"y": np.ones(8)})


def test_site_layout_follows_sorted_names(nb):
    from numpyro_b200 import models

    b = models.stochastic_volatility.bind(np.ones(5, np.float32))
    assert list(b.flat_sites) == ["nu", "s", "sigma"] and b.dim == 7
    assert [k for k, _ in b.trace_sites] == ["sigma", "s", "nu"]
    z = np.arange(14, dtype=np.float32).reshape(2, 7)
    u = b.unravel(z)
    assert u["nu"].shape == (2,) and u["s"].shape == (2, 5)
    assert np.array_equal(b.ravel(u), z)
    b8 = models.eight_schools_noncentered.bind(8, np.ones(8), y=np.ones(8))
    assert list(b8.flat_sites) == ["mu", "tau", "theta_base"] and b8.dim == 10
    out = b8.postprocess(b8.unravel(np.zeros((3, 10), np.float32)))
    assert out["theta"].shape == (3, 8) and np.allclose(out["tau"], 1.0)


def test_diagnostics_known_answers(nb):
    from numpyro_b200 import diagnostics as D

    x = np.arange(1000.0).reshape(100, 10)  # test/test_diagnostics.py:105-107
    np.testing.assert_allclose(D.effective_sample_size(x, bias=False), 52.64, atol=0.01)
    s = D.summary({"a": np.random.default_rng(0).normal(size=(2, 100, 3))})
    assert s["a"]["r_hat"].shape == (3,)
    D.print_summary({"a": np.random.default_rng(0).normal(size=(2, 100, 3))})


def test_kernel_pickles_after_binding_a_model(nb):
    # hmc.py:818-822: the kernel (and so MCMC) pickles; what init() built is dropped
    import pickle

    from numpyro_b200 import models
    from numpyro_b200.infer import MCMC, NUTS

    for model in (models.logistic_regression, models.horseshoe_normal):
        k = NUTS(model, step_size=0.5, dense_mass=False)
        k._bound = model.bind(np.zeros((4, 3), np.float32), np.zeros(4, np.float32))
        k._potential_fn_gen = k._bound
        k2 = pickle.loads(pickle.dumps(k))
        assert k2._model is model and k2._bound is None and k2._handle is None
        assert k2._step_size == 0.5
        m2 = pickle.loads(pickle.dumps(MCMC(k, num_warmup=3, num_samples=4)))
        assert m2.num_samples == 4 and m2.sampler._model is model


def _pickle_gibbs_fn(rng_key, gibbs_sites, hmc_sites):
    return {"mu": np.float32(0)}


def test_hmcgibbs_and_hmcecs_pickle(nb):
    # hmc_gibbs.py:187-190: the wrappers pickle too; the inner kernel comes back seeing the model
    # with the Gibbs sites conditioned
    import pickle

    from numpyro_b200 import models
    from numpyro_b200.infer import HMCECS, HMCGibbs, NUTS

    k = HMCGibbs(NUTS(models.eight_schools_noncentered, max_tree_depth=6), _pickle_gibbs_fn, ["mu"])
    k._prototype_trace = [{"mu": np.float32(1)}]
    k2 = pickle.loads(pickle.dumps(k))
    assert k2._prototype_trace is None and k2._model0 is models.eight_schools_noncentered
    view = k2.inner_kernel._model.bind(8, np.ones(8, np.float32), y=np.zeros(8, np.float32))
    assert view.dim == 9 and list(view.cond_index) == [0] and "mu" not in view.flat_sites
    e2 = pickle.loads(pickle.dumps(HMCECS(NUTS(models.logistic_regression), num_blocks=5)))
    assert e2.inner_kernel._model is models.logistic_regression and e2._num_blocks == 5


def test_conditioned_view_of_a_bound_model():
    """``BoundModel.condition`` (the registered-family form of ``hmc_gibbs.py:30-34``)."""
    import pytest

    from numpyro_b200 import models

    full = models.eight_schools_noncentered.bind(8, np.ones(8, np.float32), y=np.zeros(8, np.float32))
    v = full.condition(["tau"])
    assert v.dim == 9 and list(v.cond_index) == [1] and v.positive == ()
    assert [k for k, _ in v.trace_sites] == ["mu", "theta_base"]
    flat, off = v.cond_unconstrained({"tau": np.float32(2.0)})
    np.testing.assert_allclose(flat, [np.log(2.0)], rtol=1e-6)
    np.testing.assert_allclose(off, np.log(2.0), rtol=1e-6)  # log|J| of ExpTransform at the value
    v2 = full.condition(["mu", "theta_base"])
    assert v2.dim == 1 and list(v2.cond_index) == [0] + list(range(2, 10)) and v2.positive == ("tau",)
    with pytest.raises(ValueError):
        full.condition(["nope"])
    with pytest.raises(ValueError):
        full.condition(["mu", "tau", "theta_base"])
    with pytest.raises(NotImplementedError):
        models.logistic_regression.bind(np.zeros((4, 3), np.float32), np.zeros(4)).condition(["coefs"])


def test_nuts_alone_refuses_a_subsampled_plate():
    """primitives.py:444-449: without HMCECS nothing owns the subsample indices."""
    import numpy as np
    import pytest
    from numpyro_b200 import models
    from numpyro_b200.infer import HMCECS, HMCGibbs, NUTS

    X, y = np.zeros((300, 3), np.float32), np.zeros(300, np.uint8)
    with pytest.raises(ValueError, match="do not support subsampling"):
        NUTS(models.logistic_regression).init(np.array([0, 0], np.uint32), 5, None, (X, y),
                                              {"subsample_size": 50})
    with pytest.raises(ValueError, match="inner_kernel must be a HMC or NUTS sampler"):
        HMCGibbs(object(), lambda **k: {}, [])  # hmc_gibbs.py:89-90
    with pytest.raises(ValueError, match="gibbs_fn must be a callable"):
        HMCGibbs(NUTS(models.logistic_regression), None, [])
    with pytest.raises(AssertionError, match="Cannot detect any subsample statements"):
        HMCECS(NUTS(models.logistic_regression)).init(np.array([0, 0], np.uint32), 5, None, (X, y), {})
