"""The oracle's potentials against an INDEPENDENT statement of the same models.

The four registered families are written here a second time, straight from the model code
of the reference (cited below) with ``scipy.stats`` log-densities and the reference's
``biject_to`` rule (``infer/util.py:269-299``: positive support -> ExpTransform, plus
``log|J| = z``, ``transforms.py:563-571``) -- no code shared with ``oracle/potentials.py``.
``U`` must agree to float64 round-off and the oracle's closed-form gradient must agree with
central differences of this independent ``U``.  (The sampler logic is pinned separately to the
reference's own source, tests/test_refpin.py; this closes the loop for the potentials, whose
reference form is a traced program over the distributions library.)
"""
import os
import sys

import numpy as np
import pytest
from scipy import stats
from scipy.special import expit

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import refpin_cases as rc  # noqa: E402

DATA = rc.make_data()


def bernoulli_logits_logpmf(logits, y):
    p = expit(logits)
    return stats.bernoulli.logpmf(y.astype(int), p)


def U_logistic(z):
    """examples/covtype.py:66-71: coefs ~ Normal(0, 1); obs ~ Bernoulli(logits=X @ coefs)."""
    X, y = DATA["logistic_X"].astype(np.float64), DATA["logistic_y"]
    return -(stats.norm.logpdf(z, 0, 1).sum() + bernoulli_logits_logpmf(X @ z, y).sum())


def U_eight_schools(z):
    """README.md:101-110 (TransformReparam form): mu ~ N(0, 5); tau ~ HalfCauchy(5);
    theta_base ~ N(0, 1); y ~ N(mu + tau * theta_base, sigma).  Flat order mu, tau, theta_base."""
    mu, tau_u, b = z[0], z[1], z[2:]
    tau = np.exp(tau_u)
    lp = stats.norm.logpdf(mu, 0, 5) + stats.halfcauchy.logpdf(tau, scale=5) + tau_u
    lp += stats.norm.logpdf(b, 0, 1).sum()
    lp += stats.norm.logpdf(rc.Y8, mu + tau * b, rc.S8).sum()
    return -lp


def U_horseshoe(z, normal):
    """examples/horseshoe_regression.py:37-78.  Flat (sorted) order: lambdas[D], (prec_obs,)
    tau[1], unscaled_betas[D]."""
    X = DATA["hs_X"].astype(np.float64)
    D = X.shape[1]
    lam_u = z[:D]
    k = D
    if normal:
        p_u = z[k]
        k += 1
    tau_u = z[k]
    ub = z[k + 1:]
    lam, tau = np.exp(lam_u), np.exp(tau_u)
    lp = (stats.halfcauchy.logpdf(lam, scale=1) + lam_u).sum()
    lp += stats.halfcauchy.logpdf(tau, scale=1) + tau_u
    lp += stats.norm.logpdf(ub, 0, 1).sum()
    mean = X @ (tau * lam * ub)
    if normal:
        prec = np.exp(p_u)
        lp += stats.gamma.logpdf(prec, a=3.0, scale=1.0) + p_u
        lp += stats.norm.logpdf(DATA["hs_Yn"].astype(np.float64), mean, prec ** -0.5).sum()
    else:
        lp += bernoulli_logits_logpmf(mean, DATA["hs_Yb"]).sum()
    return -lp


def U_sv(z):
    """examples/stochastic_volatility.py:57-65: sigma ~ Exponential(50); nu ~ Exponential(0.1);
    s ~ GaussianRandomWalk(sigma, T); r ~ StudentT(nu, 0, exp(s)).  Flat order nu, s[T], sigma."""
    r = DATA["sv_returns"].astype(np.float64)
    T = r.shape[0]
    nu_u, s, sg_u = z[0], z[1:1 + T], z[1 + T]
    nu, sg = np.exp(nu_u), np.exp(sg_u)
    lp = stats.expon.logpdf(sg, scale=1 / 50.0) + sg_u
    lp += stats.expon.logpdf(nu, scale=1 / 0.1) + nu_u
    lp += stats.norm.logpdf(np.diff(np.concatenate([[0.0], s])), 0, sg).sum()
    lp += stats.t.logpdf(r, df=nu, loc=0.0, scale=np.exp(s)).sum()
    return -lp


CASES = {"logit": U_logistic, "es": U_eight_schools, "hsb": lambda z: U_horseshoe(z, False),
         "hsn": lambda z: U_horseshoe(z, True), "sv": U_sv}


def _oracle64(key):
    from oracle import potentials as opot

    if key == "logit":
        return opot.Logistic(DATA["logistic_X"], DATA["logistic_y"], np.float64)
    if key == "es":
        return opot.EightSchools(rc.S8, rc.Y8, np.float64)
    if key == "hsb":
        return opot.Horseshoe(DATA["hs_X"], DATA["hs_Yb"], "bernoulli", np.float64)
    if key == "hsn":
        return opot.Horseshoe(DATA["hs_X"], DATA["hs_Yn"], "normal", np.float64)
    return opot.StochasticVolatility(DATA["sv_returns"], np.float64)


@pytest.mark.parametrize("key", list(CASES))
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_oracle_potential_and_gradient_against_independent_model(key, seed):
    fam = _oracle64(key)
    z = rc.init_point(fam, seed).astype(np.float64)
    z += np.random.default_rng(seed).normal(size=z.shape) * 0.3
    U_ind = CASES[key]
    U, g = fam.potential_and_grad(z)
    assert np.isfinite(U)
    np.testing.assert_allclose(U, U_ind(z), rtol=1e-10, atol=1e-9)
    num = np.zeros_like(z)
    for i in range(z.size):
        h = 1e-5 * max(1.0, abs(z[i]))
        zp, zm = z.copy(), z.copy()
        zp[i] += h
        zm[i] -= h
        num[i] = (U_ind(zp) - U_ind(zm)) / (2 * h)
    np.testing.assert_allclose(g, num, rtol=2e-5, atol=2e-6 * max(1.0, np.abs(num).max()))
