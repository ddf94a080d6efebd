"""HMCECS oracle (``oracle/ecs.py``) against the reference's own functions, executed from
/root/reference over the NumPy stand-in when the fixtures were made
(``tests/golden/make_refecs.py`` -> ``refecs.npz``): ``_update_block``, ``_subsample_fn``,
``perturbed_method``.  Plus self-consistency of the closed-form potential (finite differences,
proxy quality) and of the sampler (posterior of a small problem)."""
import os

import numpy as np
import pytest

from oracle import ecs as oecs
from oracle import hmc_util as ohu
from oracle.potentials import Logistic
from oracle.threefry import PRNGKey

FIX = os.path.join(os.path.dirname(__file__), "golden", "refecs.npz")


@pytest.fixture(scope="module")
def fx():
    return np.load(FIX)


def _cases(fx, prefix):
    return sorted({"/".join(k.split("/")[:3]) for k in fx.files if k.startswith(prefix)})


def test_update_block_matches_reference(fx):
    tags = _cases(fx, "update_block/")
    assert len(tags) == 10
    for tag in tags:
        part = bool(int(tag.split("/")[1]))
        m, nb, size = (int(v) for v in tag.split("/")[2].split("_"))
        cfg = ohu.Config(partitionable=part)
        rk, idx, pad, new_idx, start = oecs.update_block(cfg, fx[tag + "/key"], nb, fx[tag + "/idx0"], (size, m))
        np.testing.assert_array_equal(rk, fx[tag + "/rng_key"])
        np.testing.assert_array_equal(idx, fx[tag + "/idx"])
        np.testing.assert_array_equal(new_idx, fx[tag + "/new_idx"])
        assert (start, pad) == (int(fx[tag + "/start"]), int(fx[tag + "/pad"]))
        changed = np.nonzero(idx != fx[tag + "/idx0"])[0]
        assert changed.size == 0 or (changed.min() >= start and changed.max() < start + new_idx.size)


def test_subsample_fn_matches_reference(fx):
    tags = _cases(fx, "subsample_fn/")
    assert len(tags) == 4
    for tag in tags:
        part = bool(int(tag.split("/")[1]))
        size, m = (int(v) for v in tag.split("/")[2].split("_"))
        idx = oecs.subsample_fn(ohu.Config(partitionable=part), size, m, fx[tag + "/key"])
        np.testing.assert_array_equal(idx, fx[tag + "/idx"])
        assert len(set(idx.tolist())) == m and idx.min() >= 0 and idx.max() < size


def test_perturbed_estimator_matches_reference(fx):
    X, y, th = fx["perturbed/X"], fx["perturbed/y"], fx["perturbed/theta"]
    n, m = (int(v) for v in fx["perturbed/n_m"])
    bce, _ = oecs._bce((X @ th).astype(np.float32), y)
    est, _, _ = oecs.perturbed_loglik(-bce.astype(np.float32), fx["perturbed/proxy_sub"],
                                      fx["perturbed/proxy_all"], n, m)
    np.testing.assert_allclose(est, fx["perturbed/value"], rtol=2e-6)


def _problem(n=6000, D=5, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, D)).astype(np.float32)
    th = rng.standard_normal(D) * 0.5
    y = (rng.random(n) < 1 / (1 + np.exp(-X @ th))).astype(np.uint8)
    return X, y, th


def test_randint_range_and_layouts():
    for part in (True, False):
        cfg = ohu.Config(partitionable=part)
        v = oecs.randint(cfg, PRNGKey(9), (4000,), 3, 17)
        assert v.min() == 3 and v.max() == 16 and v.dtype == np.int32
        assert abs(np.bincount(v - 3).min() / (4000 / 14) - 1) < 0.35
        assert oecs.randint(cfg, PRNGKey(9), (), 5, 5) == 5  # empty range -> minval


@pytest.mark.parametrize("degree", [0, 1, 2])
def test_ecs_potential_gradient_and_proxy(degree):
    X, y, th = _problem()
    m = 300
    fam = oecs.LogisticECS(X, y, m, degree, ref=th + 0.02, dtype=np.float64)
    fam.set_subsample(oecs.prototype_subsample(ohu.Config(), PRNGKey(3), X.shape[0], m))
    z = th + 0.05 * np.random.default_rng(1).standard_normal(th.size)
    U, g = fam.potential_and_grad(z)
    eye = np.eye(th.size)
    gn = np.array([(fam.potential_and_grad(z + 1e-6 * e)[0] - fam.potential_and_grad(z - 1e-6 * e)[0]) / 2e-6
                   for e in eye])
    np.testing.assert_allclose(g, gn, rtol=1e-6, atol=1e-5)
    full = Logistic(X, y, np.float64).potential_and_grad(z)[0]
    # the estimator targets the full-data potential: tight with the second-order proxy
    assert abs(U - full) < {0: 0.12 * full, 1: 5.0, 2: 0.5}[degree]


def test_hmcecs_oracle_recovers_posterior():
    X, y, th = _problem(n=4000, D=4, seed=2)
    fam = oecs.LogisticECS(X, y, 200, 2, ref=th.astype(np.float32))
    k = oecs.HMCECSKernel(fam, num_blocks=10, max_tree_depth=6)
    res, st = oecs.run_chain(k, PRNGKey(1), 120, 200)
    full, _, _ = __import__("oracle.hmc", fromlist=["run_mcmc"]).run_mcmc(
        Logistic(X, y), PRNGKey(2), 120, 200, max_tree_depth=6)
    sd = full["z"][0].std(0)
    assert np.all(np.abs(res["z"].mean(0) - full["z"][0].mean(0)) < 0.5 * sd + 0.02)
    assert 0.5 < res["accept_prob"].mean() <= 1.0
    assert st.z["N"].shape == (200,) and st.z["N"].dtype == np.int32
