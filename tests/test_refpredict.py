"""``log_likelihood`` / ``Predictive`` (numpyro/infer/util.py:803-1148) against the reference's own
outputs (``tests/golden/refpredict.npz``, produced by ``make_refpredict.py`` from /root/reference):
CPU -- the oracle restatement; GPU -- ``numpyro_b200.infer.log_likelihood`` / ``Predictive`` (CUDA
forward kernel through the C ABI)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import refpin_cases as rc  # noqa: E402

G = np.load(os.path.join(HERE, "golden", "refpredict.npz"))
DATA = rc.make_data()
KEY = np.array([0, 3], np.uint32)
CASES = {
    "es": ("eight_schools", {"sigma": rc.S8, "y": rc.Y8}, "obs"),
    "logit": ("logistic", {"X": DATA["logistic_X"], "y": DATA["logistic_y"]}, "obs"),
    "hsb": ("horseshoe", {"X": DATA["hs_X"], "y": DATA["hs_Yb"]}, "Y"),
    "hsn": ("horseshoe", {"X": DATA["hs_X"], "y": DATA["hs_Yn"]}, "Y"),
    "sv": ("sv", {"y": DATA["sv_returns"]}, "r"),
}


def _draws(key):
    return {k.split("/")[-1]: G[k] for k in G.files if k.startswith(key + "/draws/")}


@pytest.mark.parametrize("key", list(CASES))
def test_oracle_log_likelihood_matches_reference(key):
    from oracle import predict as P

    fam, data, site = CASES[key]
    ll = P.log_likelihood(fam, data, _draws(key))
    ref = G[f"{key}/loglik/{site}"]
    np.testing.assert_allclose(ll, ref, rtol=3e-5, atol=1e-6)


@pytest.mark.parametrize("part", [True, False])
@pytest.mark.parametrize("key", ["es", "logit", "hsb", "hsn"])
def test_oracle_predictive_matches_reference(key, part):
    from oracle import predict as P

    fam, data, site = CASES[key]
    out, det = P.predictive(fam, data, _draws(key), KEY, part)
    tag = "part" if part else "legacy"
    ref = G[f"{key}/predictive/{tag}/{site}"]
    if out.dtype == np.int32:
        assert np.array_equal(out, ref)
    else:
        np.testing.assert_allclose(out, ref, rtol=1e-6, atol=1e-6)
    for k, v in det.items():
        np.testing.assert_allclose(v, G[f"{key}/predictive/{tag}/{k}"], rtol=1e-6)


# ---------------------------------------------------------------- CUDA
def _bind(nb, key):
    m = nb.models
    return {"es": (m.eight_schools_noncentered, (8, rc.S8), {"y": rc.Y8}),
            "logit": (m.logistic_regression, (DATA["logistic_X"], DATA["logistic_y"]), {}),
            "hsb": (m.horseshoe_bernoulli, (DATA["hs_X"], DATA["hs_Yb"]), {}),
            "hsn": (m.horseshoe_normal, (DATA["hs_X"], DATA["hs_Yn"]), {}),
            "sv": (m.stochastic_volatility, (DATA["sv_returns"],), {})}[key]


@pytest.fixture(scope="module")
def nb():
    import __graft_entry__ as g

    return g.build()


@pytest.mark.gpu
@pytest.mark.parametrize("key", list(CASES))
def test_cuda_log_likelihood_matches_reference(nb, key):
    model, args, kwargs = _bind(nb, key)
    site = CASES[key][2]
    ll = nb.infer.log_likelihood(model, _draws(key), *args, **kwargs)
    assert list(ll) == [site]
    np.testing.assert_allclose(ll[site], G[f"{key}/loglik/{site}"], rtol=3e-5, atol=2e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("part", [True, False])
@pytest.mark.parametrize("key", ["es", "logit", "hsb", "hsn"])
def test_cuda_predictive_matches_reference(nb, key, part):
    model, args, kwargs = _bind(nb, key)
    site = CASES[key][2]
    tag = "part" if part else "legacy"
    # the observations are dropped (None), as Predictive requires (infer/util.py:905-907)
    if key == "es":
        pargs, pkw = args, {}
    else:
        pargs, pkw = args[:-1] + (None,), {}
    pred = nb.infer.Predictive(model, _draws(key), partitionable=part)(KEY, *pargs, **pkw)
    ref = G[f"{key}/predictive/{tag}/{site}"]
    assert set(pred) == {k.split("/")[-1] for k in G.files if k.startswith(f"{key}/predictive/{tag}/")}
    if ref.dtype == np.int32:
        assert pred[site].dtype == np.int32
        # a uniform draw within an ulp of expit(logit) may flip: none expected at this size
        assert (pred[site] != ref).sum() <= 1
    else:
        np.testing.assert_allclose(pred[site], ref, rtol=2e-6, atol=2e-6)
    for k in pred:
        if k != site:
            np.testing.assert_allclose(pred[k], G[f"{key}/predictive/{tag}/{k}"], rtol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("c", [1.0, 0.3, 0.0])
def test_cuda_log_likelihood_locscale_eight_schools(nb, c):
    """The centred eight schools / LocScaleReparam(c) give the reference's log-likelihood of the
    non-centred fixture at the same theta (the observation site is the same Normal(theta, sigma))."""
    from numpyro_b200.infer.reparam import LocScaleReparam

    d = _draws("es")
    mu, tau = np.asarray(d["mu"], np.float32), np.asarray(d["tau"], np.float32)
    theta = mu[:, None] + tau[:, None] * np.asarray(d["theta_base"], np.float32)
    if c == 1.0:
        model, draws = nb.models.eight_schools, {"mu": mu, "tau": tau, "theta": theta}
    else:
        model = nb.handlers.reparam(nb.models.eight_schools, config={"theta": LocScaleReparam(c)})
        dec = (theta - mu[:, None]) / np.power(tau[:, None], np.float32(1 - c)) + np.float32(c) * mu[:, None]
        draws = {"mu": mu, "tau": tau, "theta_decentered": dec.astype(np.float32)}
    ll = nb.infer.log_likelihood(model, draws, 8, rc.S8, y=rc.Y8)
    np.testing.assert_allclose(ll["obs"], G["es/loglik/obs"], rtol=2e-5, atol=2e-5)
