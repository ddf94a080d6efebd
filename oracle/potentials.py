"""Oracle: closed-form potential energy U(z) and gradient for the four registered
model families.  TEST INFRASTRUCTURE (see ``oracle/__init__.py``).

The reference derives U by tracing a model through its effect handlers
(``numpyro/infer/util.py:302-327`` ``potential_energy`` -> ``:109-124``
``log_density``), substituting ``value = t(z)`` and adding ``log|det J|`` for
every constrained site (``:269-299`` ``_unconstrain_reparam``), and gets the
gradient by reverse-mode AD (``hmc_util.py:242-252``).  Here each family's U is
written out from the distributions' ``log_prob`` formulas and differentiated by
hand; ``tests/test_oracle_potentials.py`` checks the log-probs against scipy and
the gradients against central finite differences in float64.

Flattening follows ``ravel_pytree`` on the site dict = sorted site names
(SURVEY.md section 8a "Flattening rule").
"""
import math

import numpy as np

LOG_SQRT_2PI = 0.5 * math.log(2.0 * math.pi)


def _lgamma(x):
    from scipy.special import gammaln

    return gammaln(x)


def _digamma(x):
    from scipy.special import digamma

    return digamma(x)


class Family:
    """A registered potential family: flat dimension, site table, U and grad."""

    name = ""
    #: (site, size) in model TRACE order (init draws happen in this order,
    #: ``numpyro/infer/util.py:425-432``)
    trace_sites = ()

    def __init__(self, dtype=np.float32):
        self.dtype = np.dtype(dtype)

    @property
    def dim(self):
        return sum(s for _, s in self.trace_sites)

    @property
    def flat_sites(self):
        """(site, size, offset) in flat (sorted-name) order."""
        out, off = [], 0
        for k, s in sorted(self.trace_sites):
            out.append((k, s, off))
            off += s
        return out

    def offsets(self):
        return {k: (off, s) for k, s, off in self.flat_sites}

    def unravel(self, z):
        return {k: z[..., off:off + s] for k, s, off in self.flat_sites}

    def potential_and_grad(self, z):
        raise NotImplementedError

    def constrain(self, z):
        """postprocess_fn (``infer/util.py:176-230,583-609``): constrained sites +
        deterministic sites."""
        raise NotImplementedError


# --------------------------------------------------------------------------
# Bayesian logistic regression: examples/covtype.py:66-71
# --------------------------------------------------------------------------
class Logistic(Family):
    """coefs ~ Normal(0, 1)[D]; obs ~ Bernoulli(logits = X @ coefs).

    Normal.log_prob: ``continuous.py:2203-2207``; Bernoulli-logits log_prob =
    ``-binary_cross_entropy_with_logits`` (``discrete.py:137-139``,
    ``distributions/util.py:295-298``).
    """

    name = "logistic"

    def __init__(self, X, y, dtype=np.float32):
        super().__init__(dtype)
        self.X = np.ascontiguousarray(X, dtype=self.dtype)
        self.y = np.ascontiguousarray(y).astype(self.dtype)
        self.trace_sites = (("coefs", self.X.shape[1]),)

    def potential_and_grad(self, z):
        f = self.dtype.type
        z = np.asarray(z, self.dtype)
        logits = self.X @ z
        # -log N(z; 0, 1) = 0.5 z^2 + log sqrt(2 pi)
        prior = (f(0.5) * z * z + f(LOG_SQRT_2PI)).sum(dtype=self.dtype)
        e = np.exp(-np.abs(logits))
        bce = np.maximum(logits, f(0)) + np.log1p(e) - logits * self.y
        U = prior + bce.sum(dtype=self.dtype)
        sig = np.where(logits >= 0, f(1) / (f(1) + e), e / (f(1) + e)).astype(self.dtype)
        grad = z + self.X.T @ (sig - self.y)
        return self.dtype.type(U), grad.astype(self.dtype)

    def constrain(self, z):
        return {"coefs": np.asarray(z)}


# --------------------------------------------------------------------------
# Eight schools, non-centred: README.md:101-110 with TransformReparam
# (reparam.py:159-185): latent sites mu, tau, theta_base; theta deterministic.
# --------------------------------------------------------------------------
class EightSchools(Family):
    """z = [mu, log tau, theta_base[J]].

    mu ~ Normal(0, 5); tau ~ HalfCauchy(5) (``continuous.py:238-244,721-723``)
    through ExpTransform (``transforms.py:563-571``: log|J| = z);
    theta_base ~ Normal(0, 1); obs ~ Normal(mu + tau * theta_base, sigma).
    """

    name = "eight_schools"

    def __init__(self, sigma, y, dtype=np.float32):
        super().__init__(dtype)
        self.sigma = np.asarray(sigma, self.dtype)
        self.y = np.asarray(y, self.dtype)
        self.J = self.y.shape[0]
        self.trace_sites = (("mu", 1), ("tau", 1), ("theta_base", self.J))

    def potential_and_grad(self, z):
        f = self.dtype.type
        z = np.asarray(z, self.dtype)
        mu, tu, b = z[0], z[1], z[2:]
        tau = np.exp(tu)
        # log N(mu; 0, 5)
        lp = -f(0.5) * (mu / f(5)) ** 2 - np.log(f(math.sqrt(2 * math.pi)) * f(5))
        # log HalfCauchy(tau; 5) + jacobian
        q = tau / f(5)
        lp = lp + (f(math.log(2.0)) - f(math.log(math.pi)) - np.log(f(5)) - np.log1p(q * q)) + tu
        # theta_base
        lp = lp + (-f(0.5) * b * b - f(LOG_SQRT_2PI)).sum(dtype=self.dtype)
        # obs
        theta = mu + tau * b
        resid = (self.y - theta) / self.sigma
        lp = lp + (-f(0.5) * resid * resid - np.log(f(math.sqrt(2 * math.pi)) * self.sigma)).sum(
            dtype=self.dtype
        )
        # gradient of U = -lp
        dtheta = -(resid / self.sigma)  # dU/dtheta_j
        g = np.empty_like(z)
        g[0] = mu / f(25) + dtheta.sum(dtype=self.dtype)
        g[1] = (f(2) * q * q / (f(1) + q * q)) - f(1) + tau * (dtheta * b).sum(dtype=self.dtype)
        g[2:] = b + tau * dtheta
        return self.dtype.type(-lp), g

    def constrain(self, z):
        z = np.asarray(z)
        mu, tau, b = z[..., 0], np.exp(z[..., 1]), z[..., 2:]
        return {"mu": mu, "tau": tau, "theta_base": b,
                "theta": mu[..., None] + tau[..., None] * b}


# --------------------------------------------------------------------------
# Horseshoe regression: examples/horseshoe_regression.py:37-78
# --------------------------------------------------------------------------
class EightSchoolsLocScale(Family):
    """Eight schools with ``reparam(config={"theta": LocScaleReparam(centered)})``
    (``numpyro/infer/reparam.py:64-145``) around the centred model ``theta ~ Normal(mu, tau)``.

    ``centered`` = 1: the centred model itself, z = [mu, log tau, theta[J]].  Otherwise the latent
    is ``theta_decentered ~ Normal(centered * mu, tau ** centered)`` (``Normal(0, 1)`` for 0,
    ``reparam.py:129-134``) and ``theta = mu + tau ** (1 - centered) * (theta_decentered -
    centered * mu)`` is deterministic (``:141-142``); z = [mu, log tau, theta_decentered[J]].
    """

    name = "eight_schools_locscale"

    def __init__(self, sigma, y, centered=0.0, dtype=np.float32):
        super().__init__(dtype)
        self.sigma = np.asarray(sigma, self.dtype)
        self.y = np.asarray(y, self.dtype)
        self.J = self.y.shape[0]
        self.centered = float(centered)
        if not 0.0 <= self.centered <= 1.0:
            raise ValueError("`centered` argument does not satisfy `0 <= centered <= 1`.")
        self.site = "theta" if self.centered == 1.0 else "theta_decentered"
        self.trace_sites = (("mu", 1), ("tau", 1), (self.site, self.J))

    def potential_and_grad(self, z):
        f = self.dtype.type
        z = np.asarray(z, self.dtype)
        c = f(self.centered)
        mu, tu, t = z[0], z[1], z[2:]
        tau = np.exp(tu)
        lp = -f(0.5) * (mu / f(5)) ** 2 - np.log(f(math.sqrt(2 * math.pi)) * f(5))
        q = tau / f(5)
        lp = lp + (f(math.log(2.0)) - f(math.log(math.pi)) - np.log(f(5)) - np.log1p(q * q)) + tu
        if self.centered == 0.0:
            loc, s = f(0), f(1)
        else:
            loc, s = mu * c, np.power(tau, c)
        a = (t - loc) / s
        lp = lp + (-f(0.5) * a * a - np.log(s) - f(LOG_SQRT_2PI)).sum(dtype=self.dtype)
        w = np.power(tau, f(1) - c)
        delta = t - c * mu
        theta = mu + w * delta
        resid = (self.y - theta) / self.sigma
        lp = lp + (-f(0.5) * resid * resid - np.log(f(math.sqrt(2 * math.pi)) * self.sigma)).sum(
            dtype=self.dtype)
        e = resid / self.sigma  # d lp / d theta_j
        g = np.empty_like(z)
        dmu = -mu / f(25) + (a * c / s).sum(dtype=self.dtype) + (e * (f(1) - w * c)).sum(dtype=self.dtype)
        dtu = -(f(2) * q * q / (f(1) + q * q)) + f(1)
        dtu = dtu + (c * a * a - c).sum(dtype=self.dtype) + ((f(1) - c) * w * (e * delta)).sum(dtype=self.dtype)
        g[0] = -dmu
        g[1] = -dtu
        g[2:] = -(-a / s + e * w)
        return self.dtype.type(-lp), g

    def constrain(self, z):
        z = np.asarray(z)
        c = self.centered
        mu, tau, t = z[..., 0], np.exp(z[..., 1]), z[..., 2:]
        out = {"mu": mu, "tau": tau, self.site: t}
        if c != 1.0:
            out["theta"] = mu[..., None] + tau[..., None] ** (1 - c) * (t - c * mu[..., None])
        return out


# --------------------------------------------------------------------------
# GLMs on a dense design matrix
# --------------------------------------------------------------------------
class PoissonGLM(Family):
    """coefs ~ Normal(0, 1)[D]; obs ~ Poisson(rate = exp(X @ coefs)).

    ``Poisson.log_prob`` (``discrete.py:743``): ``log(rate) * y - gammaln(y + 1) - rate``."""

    name = "poisson_regression"

    def __init__(self, X, y, dtype=np.float32):
        super().__init__(dtype)
        self.X = np.ascontiguousarray(X, dtype=self.dtype)
        self.y = np.ascontiguousarray(y).astype(self.dtype)
        self.trace_sites = (("coefs", self.X.shape[1]),)
        self._lg = _lgamma(self.y.astype(np.float64) + 1.0)

    def potential_and_grad(self, z):
        f = self.dtype.type
        z = np.asarray(z, self.dtype)
        prior = (f(0.5) * z * z + f(LOG_SQRT_2PI)).sum(dtype=self.dtype)
        rate = np.exp(self.X @ z)
        lp = (np.log(rate) * self.y - self._lg.astype(self.dtype) - rate).sum(dtype=self.dtype)
        grad = z + self.X.T @ (rate - self.y)
        return f(prior - lp), grad.astype(self.dtype)

    def constrain(self, z):
        return {"coefs": np.asarray(z)}


class NormalGLM(Family):
    """coefs ~ Normal(0, 1)[D]; prec_obs ~ Gamma(3, 1) through ExpTransform;
    obs ~ Normal(X @ coefs, prec_obs ** -0.5) (the observation model of
    ``examples/horseshoe_regression.py:60-66``).  z = [coefs, log prec_obs]."""

    name = "linear_regression"

    def __init__(self, X, y, dtype=np.float32):
        super().__init__(dtype)
        self.X = np.ascontiguousarray(X, dtype=self.dtype)
        self.y = np.ascontiguousarray(y).astype(self.dtype)
        self.trace_sites = (("coefs", self.X.shape[1]), ("prec_obs", 1))

    def potential_and_grad(self, z):
        f = self.dtype.type
        z = np.asarray(z, self.dtype)
        D, N = self.X.shape[1], self.X.shape[0]
        b, pu = z[:D], z[D]
        prec = np.exp(pu)
        lp = (-f(0.5) * b * b - f(LOG_SQRT_2PI)).sum(dtype=self.dtype)
        # Gamma(3, 1).log_prob(prec) = 2 log prec - prec - lgamma(3); + log|J| = pu
        lp = lp + (f(2) * pu - prec - f(math.log(2.0))) + pu
        r = self.y - self.X @ b
        ss = (r * r).sum(dtype=self.dtype)
        lp = lp + (-f(0.5) * prec * ss - f(N) * f(LOG_SQRT_2PI) + f(0.5) * f(N) * pu)
        g = np.empty_like(z)
        g[:D] = b + self.X.T @ (-(prec * r))
        g[D] = -(f(3) - prec) + f(0.5) * prec * ss - f(0.5) * f(N)
        return f(-lp), g.astype(self.dtype)

    def constrain(self, z):
        z = np.asarray(z)
        return {"coefs": z[..., :-1], "prec_obs": np.exp(z[..., -1])}


class BetaBernoulli(Family):
    """``p_latent ~ Beta(alpha, beta)[K]``; ``obs[n, K] ~ Bernoulli(p_latent)``
    (``test/infer/test_mcmc.py:239-244``).  z = logit(p) (``SigmoidTransform``, clipped expit
    ``transforms.py:59-61``, log|J| = -softplus(z) - softplus(-z) ``:960-961``); Beta through the
    Dirichlet log-density (``continuous.py:193-194,292-299``); Bernoulli(probs) with clamped
    probabilities (``discrete.py:90-93``, ``util.py:492-494``).  The likelihood enters through
    the column counts of ones."""

    name = "beta_bernoulli"

    def __init__(self, data, alpha=1.1, beta=1.1, dtype=np.float32):
        super().__init__(dtype)
        data = np.asarray(data)
        self.K = data.shape[1]
        self.n = self.dtype.type(data.shape[0])
        self.s = data.sum(0).astype(self.dtype)
        self.a = np.broadcast_to(np.asarray(alpha, self.dtype), (self.K,)).copy()
        self.b = np.broadcast_to(np.asarray(beta, self.dtype), (self.K,)).copy()
        self.norm = np.array([math.lgamma(a) + math.lgamma(b) - math.lgamma(a + b)
                              for a, b in zip(self.a.astype(float), self.b.astype(float))], self.dtype)
        self.trace_sites = (("p_latent", self.K),)

    def _p(self, z):
        fi = np.finfo(self.dtype)
        return np.clip(1 / (1 + np.exp(-z)), fi.tiny, self.dtype.type(1) - fi.eps).astype(self.dtype)

    def potential_and_grad(self, z):
        f = self.dtype.type
        z = np.asarray(z, self.dtype)
        p = self._p(z)
        q = f(1) - p
        lp = (np.log(p) * (self.a - f(1)) + np.log(q) * (self.b - f(1))) - self.norm
        lp = lp + (-np.logaddexp(z, f(0)) - np.logaddexp(-z, f(0)))
        lp = lp + self.s * np.log(p) + (self.n - self.s) * np.log1p(-p)
        dl = (self.a - f(1)) * q - (self.b - f(1)) * p + (f(1) - f(2) * p) + self.s * q - (self.n - self.s) * p
        return f(-lp.sum(dtype=self.dtype)), (-dl).astype(self.dtype)

    def constrain(self, z):
        return {"p_latent": self._p(np.asarray(z, self.dtype))}


class DirichletCategorical(Family):
    """``p_latent ~ Dirichlet(concentration[K])``; ``obs[n] ~ Categorical(p_latent)``
    (``test/infer/test_mcmc.py:283-288``).  z[K-1] through ``StickBreakingTransform``
    (``transforms.py:1090-1122``: x = z - log(K-1-i), sticks = clipped expit, log|J| = sum(log
    y[:-1] + log_sigmoid(x) - x)); Dirichlet ``continuous.py:292-299``; Categorical(probs) =
    clipped ``log(probs)`` at the observed class (``discrete.py:66-68,323-329``): class counts."""

    name = "dirichlet_categorical"

    def __init__(self, data, concentration=(1.0, 1.0, 1.0), dtype=np.float32):
        super().__init__(dtype)
        self.c = np.asarray(concentration, self.dtype)
        self.K = self.c.shape[0]
        self.counts = np.bincount(np.asarray(data).astype(np.int64), minlength=self.K).astype(self.dtype)
        cf = self.c.astype(float)
        self.norm = self.dtype.type(sum(math.lgamma(v) for v in cf) - math.lgamma(cf.sum()))
        self.offset = np.log(np.arange(self.K - 1, 0, -1).astype(self.dtype)).astype(self.dtype)
        self.trace_sites = (("p_latent", self.K - 1),)

    def _probs(self, z):
        f, fi = self.dtype.type, np.finfo(self.dtype)
        x = (np.asarray(z, self.dtype) - self.offset).astype(self.dtype)
        s = np.clip(1 / (1 + np.exp(-x)), fi.tiny, f(1) - fi.eps).astype(self.dtype)
        rem = np.concatenate([[f(1)], np.cumprod(f(1) - s)]).astype(self.dtype)
        return x, s, (np.concatenate([s, [f(1)]]) * rem).astype(self.dtype)

    def potential_and_grad(self, z):
        f = self.dtype.type
        x, s, p = self._probs(z)
        logp = np.log(p)
        lp = (logp * (self.c - f(1))).sum(dtype=self.dtype) - self.norm
        log_sig = -np.logaddexp(-x, f(0))
        lp = lp + (logp[:-1] + (log_sig - x)).sum(dtype=self.dtype)
        lp = lp + (self.counts * np.maximum(logp, np.finfo(self.dtype).min)).sum(dtype=self.dtype)
        w = self.c - f(1) + self.counts
        w[:-1] += f(1)
        tail = np.cumsum(w[::-1])[::-1]  # tail[k] = sum_{j >= k} w_j
        dl = w[:-1] * (f(1) - s) - s * tail[1:] - s
        return f(-lp), (-dl).astype(self.dtype)

    def constrain(self, z):
        z = np.asarray(z, self.dtype)
        if z.ndim == 1:
            return {"p_latent": self._probs(z)[2]}
        return {"p_latent": np.stack([self._probs(r)[2] for r in z.reshape(-1, self.K - 1)]).reshape(
            z.shape[:-1] + (self.K,))}


class Horseshoe(Family):
    """z = [lambdas_u[D], (prec_obs_u,) tau_u, unscaled_betas[D]] (sorted names).

    lambdas ~ HalfCauchy(1)[D], tau ~ HalfCauchy(1)[1] (exp transform + log|J|),
    unscaled_betas ~ Normal(0,1)[D]; betas = tau*lambdas*unscaled_betas;
    likelihood "bernoulli": Y ~ Bernoulli(logits = X betas) (``:78``) or
    "normal": prec_obs ~ Gamma(3,1) (``continuous.py:515-525``),
    Y ~ Normal(X betas, prec_obs**-0.5) (``:54-58``).
    """

    name = "horseshoe"

    def __init__(self, X, Y, likelihood="bernoulli", dtype=np.float32):
        super().__init__(dtype)
        assert likelihood in ("bernoulli", "normal")
        self.likelihood = likelihood
        self.X = np.ascontiguousarray(X, dtype=self.dtype)
        self.Y = np.asarray(Y).astype(self.dtype)
        D = self.X.shape[1]
        self.D = D
        if likelihood == "normal":
            self.trace_sites = (("lambdas", D), ("tau", 1), ("unscaled_betas", D), ("prec_obs", 1))
        else:
            self.trace_sites = (("lambdas", D), ("tau", 1), ("unscaled_betas", D))

    def potential_and_grad(self, z):
        f = self.dtype.type
        z = np.asarray(z, self.dtype)
        off = self.offsets()
        lu = z[off["lambdas"][0]:off["lambdas"][0] + self.D]
        tu = z[off["tau"][0]]
        ub = z[off["unscaled_betas"][0]:off["unscaled_betas"][0] + self.D]
        lam, tau = np.exp(lu), np.exp(tu)
        c_hc = f(math.log(2.0) - math.log(math.pi))
        lp = (c_hc - np.log1p(lam * lam) + lu).sum(dtype=self.dtype)
        lp = lp + (c_hc - np.log1p(tau * tau) + tu)
        lp = lp + (-f(0.5) * ub * ub - f(LOG_SQRT_2PI)).sum(dtype=self.dtype)
        beta = tau * lam * ub
        m = self.X @ beta
        g = np.zeros_like(z)
        if self.likelihood == "bernoulli":
            e = np.exp(-np.abs(m))
            lp = lp - (np.maximum(m, f(0)) + np.log1p(e) - m * self.Y).sum(dtype=self.dtype)
            sig = np.where(m >= 0, f(1) / (f(1) + e), e / (f(1) + e)).astype(self.dtype)
            dm = sig - self.Y  # dU/dm
        else:
            pu = z[off["prec_obs"][0]]
            prec = np.exp(pu)
            # Gamma(3, 1): (a-1) log x - x - lgamma(a) ; + jacobian pu
            lp = lp + (f(2) * pu - prec - f(math.lgamma(3.0))) + pu
            resid = self.Y - m
            n = f(self.Y.shape[0])
            # log N(Y; m, prec^-1/2) = -0.5 prec r^2 - log(sqrt(2pi)) + 0.5 log prec
            ss = (resid * resid).sum(dtype=self.dtype)
            lp = lp + (-f(0.5) * prec * ss - n * f(LOG_SQRT_2PI) + f(0.5) * n * pu)
            dm = -prec * resid
            g[off["prec_obs"][0]] = -(f(3) - prec) + f(0.5) * prec * ss - f(0.5) * n
        dbeta = self.X.T @ dm  # dU/dbeta
        g[off["lambdas"][0]:off["lambdas"][0] + self.D] = (
            f(2) * lam * lam / (f(1) + lam * lam) - f(1) + dbeta * beta
        )
        g[off["tau"][0]] = f(2) * tau * tau / (f(1) + tau * tau) - f(1) + (dbeta * beta).sum(
            dtype=self.dtype
        )
        g[off["unscaled_betas"][0]:off["unscaled_betas"][0] + self.D] = ub + dbeta * tau * lam
        return self.dtype.type(-lp), g

    def constrain(self, z):
        z = np.asarray(z)
        off = self.offsets()
        out = {}
        for k, (o, s) in off.items():
            v = z[..., o:o + s]
            out[k] = v if k == "unscaled_betas" else np.exp(v)
        if "prec_obs" in out:
            out["prec_obs"] = out["prec_obs"][..., 0]
        out["betas"] = out["tau"] * out["lambdas"] * out["unscaled_betas"]
        return out


# --------------------------------------------------------------------------
# Stochastic volatility: examples/stochastic_volatility.py:57-65
# --------------------------------------------------------------------------
class StochasticVolatility(Family):
    """z = [nu_u, s[T], sigma_u] (sorted names: nu, s, sigma).

    sigma ~ Exponential(50) (``continuous.py:473-475``), s ~
    GaussianRandomWalk(sigma, T) (``:684-689``; support real_vector -> no
    transform, ``infer/util.py:287-291``), nu ~ Exponential(0.1),
    r ~ StudentT(nu, 0, exp(s)) (``:2375-2385``).
    """

    name = "stochastic_volatility"

    def __init__(self, returns, dtype=np.float32):
        super().__init__(dtype)
        self.r = np.asarray(returns, self.dtype)
        self.T = self.r.shape[0]
        self.trace_sites = (("sigma", 1), ("s", self.T), ("nu", 1))

    def potential_and_grad(self, z):
        f = self.dtype.type
        z = np.asarray(z, self.dtype)
        T = self.T
        nu_u, s, sg_u = z[0], z[1:1 + T], z[1 + T]
        nu, sg = np.exp(nu_u), np.exp(sg_u)
        lp = (f(math.log(50.0)) - f(50) * sg) + sg_u
        lp = lp + (f(math.log(0.1)) - f(0.1) * nu) + nu_u
        # GRW: N(s0; 0, sg) * prod N(s_t; s_{t-1}, sg)
        d = np.empty_like(s)
        d[0] = s[0]
        d[1:] = s[1:] - s[:-1]
        ds = d / sg
        lp = lp + (-f(0.5) * ds * ds).sum(dtype=self.dtype) - f(T) * (f(LOG_SQRT_2PI) + sg_u)
        # StudentT
        y = self.r * np.exp(-s)
        y2n = y * y / nu
        zt = f(0.5) * nu_u + f(0.5 * math.log(math.pi)) + f(_lgamma(0.5 * float(nu))) - f(
            _lgamma(0.5 * (float(nu) + 1.0))
        )
        l1p = np.log1p(y2n)
        lp = lp + (-f(0.5) * (nu + f(1)) * l1p - s).sum(dtype=self.dtype) - f(T) * zt
        g = np.empty_like(z)
        # d/ds_t of -GRW
        gs = ds / sg  # dU/dd_t
        gs_full = gs.copy()
        gs_full[:-1] -= gs[1:]
        # d/ds_t of -StudentT: 1 - (nu+1) * y2n/(1+y2n)
        w = y2n / (f(1) + y2n)
        g[1:1 + T] = gs_full + (f(1) - (nu + f(1)) * w)
        # d/dsigma_u
        g[1 + T] = f(50) * sg - f(1) - (ds * ds).sum(dtype=self.dtype) + f(T)
        # d/dnu_u: nu * d/dnu
        dig = f(_digamma(0.5 * float(nu))) - f(_digamma(0.5 * (float(nu) + 1.0)))
        dnu = (f(0.5) * l1p - f(0.5) * (nu + f(1)) * w / nu).sum(dtype=self.dtype) + f(T) * (
            f(0.5) / nu + f(0.5) * dig
        )
        g[0] = f(0.1) * nu - f(1) + nu * dnu
        return self.dtype.type(-lp), g

    def constrain(self, z):
        z = np.asarray(z)
        T = self.T
        return {"nu": np.exp(z[..., 0]), "s": z[..., 1:1 + T], "sigma": np.exp(z[..., 1 + T])}
