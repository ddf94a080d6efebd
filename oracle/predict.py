"""Oracle restatement of ``log_likelihood`` and ``Predictive`` (``numpyro/infer/util.py:803-1148``)
for the registered model families.  TEST INFRASTRUCTURE (see ``oracle/__init__.py``).

Every family's observation site is ``obs ~ L(linear predictor, scale)``; the reference's
handlers reduce, for these models, to:

* ``log_likelihood``: ``site["fn"].log_prob(site["value"])`` at the observed site, per draw
  (``infer/util.py:1113-1121``) -- Bernoulli(logits) ``discrete.py:137-139`` +
  ``distributions/util.py:295-298``; Normal ``continuous.py:2203-2207``; StudentT ``:2375-2385``.
* ``Predictive``: ``rng_key -> split(key, S)`` (``:874-877``), per draw ``seed(model, key_s)``: the
  seed handler splits ``key, sub = split(key)`` only at sample statements that have no value yet
  (``handlers.py`` ``seed.process_message``: substituted sites return early), so in these models
  the observation site takes the sub-key of the FIRST split;
  Bernoulli draws ``uniform(sub, shape) < expit(logits)``, Normal draws ``loc + scale * normal``.
"""
import numpy as np
from scipy.special import gammaln

from . import threefry as T

F = np.float32


def linear_predictor(family, data, draws):
    """(loc [S, N], extra) per family; draws are CONSTRAINED posterior samples."""
    if family == "logistic":
        return (np.asarray(draws["coefs"], F) @ np.asarray(data["X"], F).T).astype(F), {}
    if family == "eight_schools":
        theta = (np.asarray(draws["mu"], F)[:, None]
                 + np.asarray(draws["tau"], F)[:, None] * np.asarray(draws["theta_base"], F)).astype(F)
        return theta, {"theta": theta, "scale": np.asarray(data["sigma"], F)[None, :]}
    if family == "horseshoe":
        betas = (np.asarray(draws["tau"], F) * np.asarray(draws["lambdas"], F)
                 * np.asarray(draws["unscaled_betas"], F)).astype(F)
        extra = {"betas": betas}
        if "prec_obs" in draws:
            extra["scale"] = (F(1.0) / np.sqrt(np.asarray(draws["prec_obs"], F)))[:, None]
        return (betas @ np.asarray(data["X"], F).T).astype(F), extra
    if family == "sv":
        return np.zeros_like(np.asarray(draws["s"], F)), {"scale": np.exp(np.asarray(draws["s"], F)),
                                                          "df": np.asarray(draws["nu"], F)[:, None]}
    raise ValueError(family)


def likelihood_of(family, draws):
    if family == "logistic" or (family == "horseshoe" and "prec_obs" not in draws):
        return "bernoulli_logits"
    return "student_t" if family == "sv" else "normal"


def log_likelihood(family, data, draws):
    loc, ex = linear_predictor(family, data, draws)
    y = np.asarray(data["y"], F)[None, :]
    lik = likelihood_of(family, draws)
    with np.errstate(all="ignore"):
        if lik == "bernoulli_logits":  # -binary_cross_entropy_with_logits
            return -(np.maximum(loc, F(0)) - loc * y + np.log1p(np.exp(-np.abs(loc)))).astype(F)
        if lik == "normal":
            s = ex["scale"]
            return (-((y - loc) ** 2) / (F(2) * s * s) - np.log(s) - F(0.9189385332046727)).astype(F)
        df, s = ex["df"], ex["scale"]
        yy = (y - loc) / s
        z = np.log(s) + F(0.5) * np.log(df) + F(0.5 * np.log(np.pi)) + gammaln(F(0.5) * df) - gammaln(F(0.5) * (df + F(1)))
        return (-F(0.5) * (df + F(1)) * np.log1p(yy * yy / df) - z).astype(F)


def predictive(family, data, draws, rng_key, partitionable=True):
    """Draws of the observation site (+ the deterministic sites), ``[S, N]``."""
    loc, ex = linear_predictor(family, data, draws)
    S, N = loc.shape
    lik = likelihood_of(family, draws)
    keys = T.split(rng_key, S, partitionable) if S > 1 else np.asarray(rng_key)[None]
    out = np.zeros((S, N), np.int32 if lik == "bernoulli_logits" else F)
    for s in range(S):
        _, sub = T.split(keys[s], 2, partitionable)
        if lik == "bernoulli_logits":
            with np.errstate(all="ignore"):
                p = (F(1) / (F(1) + np.exp(-loc[s]))).astype(F)
            out[s] = T.uniform(sub, (N,), partitionable=partitionable) < p
        else:
            sc = np.broadcast_to(ex["scale"], (S, N))[s]
            out[s] = loc[s] + sc * T.normal(sub, (N,), partitionable)
    det = {k: v for k, v in ex.items() if k in ("theta", "betas")}
    return out, det
