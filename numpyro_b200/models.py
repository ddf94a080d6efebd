"""Registry of potential families.

The reference turns an arbitrary Python model into a potential by tracing it through
its effect handlers (``numpyro/infer/util.py:632-800``).  Here generic tracing is replaced
by closed-form potentials for the registered families; a model outside the registry
raises ``NotImplementedError`` -- there is no fallback.

Each registered model is a callable marker with the same *signature* as the reference's
example model, so user code reads ``NUTS(models.logistic_regression)`` /
``mcmc.run(key, X, y)``.
"""
from collections import OrderedDict

import numpy as np

LOGISTIC, EIGHT_SCHOOLS, HORSESHOE, SV, LOGISTIC_ECS, EIGHT_SCHOOLS_LOCSCALE, BETA_BERNOULLI = 0, 1, 2, 3, 4, 5, 6
DIRICHLET_CATEGORICAL = 7
GLM = 8


class BoundModel:
    """A family bound to its data: what ``initialize_model`` returns in the reference
    (``ModelInfo``: site layout, postprocess_fn), minus the potential, which lives in CUDA."""

    def __init__(self, family, name, N, ncols, data0, data1, trace_sites, constrain,
                 likelihood=0, positive=()):
        self.family, self.name, self.N, self.ncols = family, name, N, ncols
        self.likelihood = likelihood
        self.data0, self.data1 = data0, data1
        self.trace_sites = tuple(trace_sites)  # (site, shape) in model trace order
        self._constrain = constrain
        self.positive = tuple(positive)  # sites with `positive` support (ExpTransform)
        off, self.flat_sites = 0, OrderedDict()
        for k, shape in sorted(self.trace_sites):  # ravel_pytree: sorted site names
            size = int(np.prod(shape)) if len(shape) else 1
            self.flat_sites[k] = (off, size, tuple(shape))
            off += size
        self.dim = off

    def unravel(self, zflat):
        """flat ``[..., D]`` -> dict of sites (batch dims kept)."""
        zflat = np.asarray(zflat)
        batch = zflat.shape[:-1]
        return {k: zflat[..., o:o + s].reshape(batch + shp) for k, (o, s, shp) in
                self.flat_sites.items()}

    def ravel(self, zdict):
        parts = []
        batch = None
        for k, (o, s, shp) in self.flat_sites.items():
            v = np.asarray(zdict[k], np.float32)
            b = v.shape[:v.ndim - len(shp)]
            batch = b if batch is None else batch
            parts.append(v.reshape(b + (s,)))
        return np.concatenate(parts, axis=-1)

    def unconstrain(self, values):
        """Constrained site values -> unconstrained (``transforms.py:563-571`` inverse: log for
        the ``positive`` sites, identity for real ones), as ``init_to_value`` needs."""
        out = {}
        for k, (o, size, shp) in self.flat_sites.items():
            v = np.asarray(values[k], np.float32)
            if k in self.positive:
                out[k] = np.log(v)
            elif k in getattr(self, "unit_interval", ()):  # SigmoidTransform inverse: logit
                out[k] = np.log(v) - np.log1p(-v)
            else:
                out[k] = v
        return out

    def postprocess(self, zdict):
        """``postprocess_fn``: unconstrained dict -> constrained + deterministic sites
        (``infer/util.py:176-230,583-609``)."""
        return self._constrain(zdict)

    def condition(self, names):
        """The model with the latent sites ``names`` conditioned on values that are supplied at
        run time (``hmc_gibbs.py:30-34``: ``condition`` + ``substitute`` of the Gibbs sites): a view
        whose site layout, dimension and trace order cover the remaining sites only; ``full`` is
        the unconditioned model and ``cond_index`` the conditioned coordinates of ITS flat vector
        (``b200mc_config.cond_index_host``)."""
        import copy

        names = list(names)
        unknown = [k for k in names if k not in self.flat_sites]
        if unknown:
            raise ValueError(f"cannot condition on {unknown}: not latent sites of {self.name}")
        if self.family in (LOGISTIC, LOGISTIC_ECS):
            raise NotImplementedError("conditioned sites are not available for the streaming logistic kernels")
        b = copy.copy(self)
        b.full = self
        b.cond_sites = OrderedDict((k, v) for k, v in self.flat_sites.items() if k in names)
        b.cond_index = np.ascontiguousarray(np.concatenate(
            [np.arange(o, o + n) for o, n, _ in b.cond_sites.values()]), np.int32)
        b.trace_sites = tuple((k, shp) for k, shp in self.trace_sites if k not in names)
        if not b.trace_sites:
            raise ValueError("every latent site is conditioned: nothing left for HMC")
        b.positive = tuple(k for k in self.positive if k not in names)
        if hasattr(self, "unit_interval"):
            b.unit_interval = tuple(k for k in self.unit_interval if k not in names)
        off, b.flat_sites = 0, OrderedDict()
        for k, shape in sorted(b.trace_sites):
            size = int(np.prod(shape)) if len(shape) else 1
            b.flat_sites[k] = (off, size, tuple(shape))
            off += size
        b.dim = off
        return b

    def cond_unconstrained(self, values):
        """Constrained values of the conditioned sites -> (flat unconstrained ``[..., cond_count]``,
        potential offset ``[...]``).  The offset is the conditioned sites' log|Jacobian|: the
        device potential of the full model carries ``-log|J|`` for every site, the reference's
        conditioned density does not for a site that is given in constrained space."""
        u = self.full.unconstrain({**{k: np.zeros(shp, np.float32) + 0.5 for k, (_, _, shp) in
                                      self.full.flat_sites.items()}, **values})
        parts, off = [], 0.0
        for k, (_, n, shp) in self.cond_sites.items():
            v = np.asarray(u[k], np.float32)
            batch = v.shape[:v.ndim - len(shp)]
            parts.append(v.reshape(batch + (n,)))
            if k in self.full.positive:        # ExpTransform: log|J| = z
                off = off + parts[-1].sum(-1)
            elif k in getattr(self.full, "unit_interval", ()):  # SigmoidTransform: log|J| = -softplus(-z) - softplus(z)
                off = off - (np.logaddexp(0, -parts[-1]) + np.logaddexp(0, parts[-1])).sum(-1)
            elif self.full.family == DIRICHLET_CATEGORICAL:
                raise NotImplementedError("a simplex site cannot be conditioned")
        flat = np.concatenate(parts, axis=-1)
        return flat, np.zeros(flat.shape[:-1], np.float32) + np.asarray(off, np.float32)


class RegisteredModel:
    def __init__(self, name, binder, doc):
        self.__name__ = name
        self.__doc__ = doc
        self._binder = binder

    def bind(self, *args, **kwargs):
        return self._binder(*args, **kwargs)

    def __call__(self, *args, **kwargs):
        raise NotImplementedError(
            f"{self.__name__} is a registered potential family, not a traced program: "
            "pass it to numpyro_b200.infer.NUTS/HMC")


def _bind_logistic(data, labels, subsample_size=None):
    X = np.ascontiguousarray(data, np.float32)
    y = np.ascontiguousarray(labels).astype(np.uint8, copy=False)
    if X.ndim != 2 or y.shape != (X.shape[0],):
        raise ValueError("logistic_regression expects data [N, D] and labels [N]")
    if subsample_size is not None and subsample_size != X.shape[0]:
        # plate("N", N, subsample_size=m) (examples/covtype.py:69): the potential is evaluated on a
        # subsample of the rows that HMCECS owns (hmc_gibbs.py:502-665)
        bound = BoundModel(LOGISTIC_ECS, "logistic", X.shape[0], X.shape[1], X, y,
                           [("coefs", (X.shape[1],))], lambda z: dict(z))
        bound.subsample_size = int(subsample_size)
        bound.plate_name = "N"
        return bound
    return BoundModel(LOGISTIC, "logistic", X.shape[0], X.shape[1], X, y,
                      [("coefs", (X.shape[1],))], lambda z: dict(z))


def _bind_eight_schools(J, sigma, y=None):
    if y is None:
        raise NotImplementedError("eight_schools without observations (prior sampling) is not registered")
    sigma = np.ascontiguousarray(sigma, np.float32)
    y = np.ascontiguousarray(y, np.float32)
    assert sigma.shape == (J,) and y.shape == (J,)

    def constrain(z):
        mu, tau = z["mu"], np.exp(z["tau"])
        out = {"mu": mu, "tau": tau, "theta_base": z["theta_base"]}
        out["theta"] = mu[..., None] + tau[..., None] * z["theta_base"]
        return out

    bound = BoundModel(EIGHT_SCHOOLS, "eight_schools", J, 0, y, sigma,
                       [("mu", ()), ("tau", ()), ("theta_base", (J,))], constrain,
                       positive=("tau",))
    # priors with a registered sampler (HMCGibbs' prototype trace draws its Gibbs sites from them)
    bound.prior = {"mu": ("normal", 0.0, 5.0), "theta_base": ("normal", 0.0, 1.0)}
    return bound


def _bind_eight_schools_locscale(centered):
    """The centred model ``theta ~ Normal(mu, tau)`` (README.md:60-68) under
    ``LocScaleReparam(centered)`` (``numpyro/infer/reparam.py:64-145``); ``centered`` = 1 is the
    centred model itself (``:119-120``)."""
    centered = float(centered)
    if not 0.0 <= centered <= 1.0:  # reparam.py:95-99
        raise ValueError("`centered` argument does not satisfy `0 <= centered <= 1`.")
    site = "theta" if centered == 1.0 else "theta_decentered"

    def bind(J, sigma, y=None):
        if y is None:
            raise NotImplementedError("eight_schools without observations (prior sampling) is not registered")
        sigma_ = np.ascontiguousarray(sigma, np.float32)
        y_ = np.ascontiguousarray(y, np.float32)
        assert sigma_.shape == (J,) and y_.shape == (J,)

        def constrain(z):
            mu, tau = z["mu"], np.exp(z["tau"])
            out = {"mu": mu, "tau": tau, site: z[site]}
            if centered != 1.0:  # reparam.py:141-145: the reparametrised site becomes deterministic
                delta = z[site] - np.float32(centered) * mu[..., None]
                out["theta"] = mu[..., None] + np.power(tau[..., None], np.float32(1.0 - centered)) * delta
            return out

        bound = BoundModel(EIGHT_SCHOOLS_LOCSCALE, "eight_schools_locscale", J, 0, y_, sigma_,
                           [("mu", ()), ("tau", ()), (site, (J,))], constrain, positive=("tau",))
        bound.centered = centered
        return bound

    return bind


def _bind_beta_bernoulli(data, alpha=1.1, beta=1.1):
    """``test/infer/test_mcmc.py:239-244``: ``p_latent ~ Beta(alpha, beta)``, one per column of
    ``data``; ``obs ~ Bernoulli(p_latent)``.  The potential sees the column counts of ones."""
    import math

    data = np.asarray(data)
    if data.ndim != 2:
        raise ValueError("beta_bernoulli expects data [n, K] of 0/1")
    n, K = data.shape
    a = np.broadcast_to(np.asarray(alpha, np.float32), (K,))
    b = np.broadcast_to(np.asarray(beta, np.float32), (K,))
    norm = [math.lgamma(float(x)) + math.lgamma(float(y)) - math.lgamma(float(x) + float(y)) for x, y in zip(a, b)]
    table = np.stack([np.full(K, n, np.float32), a, b, np.asarray(norm, np.float32)], axis=1)
    fi = np.finfo(np.float32)

    def constrain(z):  # SigmoidTransform with the clipped expit (transforms.py:59-61)
        zz = np.asarray(z["p_latent"], np.float32)
        return {"p_latent": np.clip(1 / (1 + np.exp(-zz)), fi.tiny, np.float32(1) - fi.eps).astype(np.float32)}

    bound = BoundModel(BETA_BERNOULLI, "beta_bernoulli", K, 0, data.sum(0).astype(np.float32),
                       np.ascontiguousarray(table, np.float32), [("p_latent", (K,))], constrain)
    bound.unit_interval = ("p_latent",)
    return bound


def _stick_breaking(z):
    """``StickBreakingTransform.__call__`` (``transforms.py:1094-1108``), float32."""
    z = np.asarray(z, np.float32)
    K1 = z.shape[-1]
    fi = np.finfo(np.float32)
    x = z - np.log(np.arange(K1, 0, -1, dtype=np.float32))
    s = np.clip(1 / (1 + np.exp(-x)), fi.tiny, np.float32(1) - fi.eps).astype(np.float32)
    rem = np.cumprod(1 - s, axis=-1)
    ones = np.ones(z.shape[:-1] + (1,), np.float32)
    return (np.concatenate([s, ones], -1) * np.concatenate([ones, rem], -1)).astype(np.float32)


def _bind_dirichlet_categorical(data, concentration=(1.0, 1.0, 1.0)):
    """``test/infer/test_mcmc.py:283-288``: ``p_latent ~ Dirichlet(concentration)``;
    ``obs ~ Categorical(p_latent)``.  The potential sees the class counts."""
    import math

    c = np.asarray(concentration, np.float32)
    K = c.shape[0]
    data = np.asarray(data)
    if data.ndim != 1 or data.min() < 0 or data.max() >= K:
        raise ValueError("dirichlet_categorical expects class labels [n] in 0..K-1")
    counts = np.bincount(data.astype(np.int64), minlength=K).astype(np.float32)
    norm = sum(math.lgamma(float(v)) for v in c) - math.lgamma(float(c.sum()))
    table = np.concatenate([c, [np.float32(norm)]]).astype(np.float32)
    bound = BoundModel(DIRICHLET_CATEGORICAL, "dirichlet_categorical", K, 0, counts, table,
                       [("p_latent", (K - 1,))], lambda z: {"p_latent": _stick_breaking(z["p_latent"])})

    def unconstrain(values):  # StickBreakingTransform._inverse (transforms.py:1110-1115)
        y = np.asarray(values["p_latent"], np.float32)
        crop = y[..., :-1]
        rem = np.clip(1 - np.cumsum(crop, -1), np.finfo(np.float32).tiny, None)
        return {"p_latent": np.log(crop / rem) + np.log(np.arange(K - 1, 0, -1, dtype=np.float32))}

    bound.unconstrain = unconstrain
    return bound


def _bind_glm(likelihood):
    """GLMs on a dense design matrix: ``coefs ~ Normal(0, 1)[D]`` and
    ``obs ~ Poisson(exp(data @ coefs))`` (likelihood 0) or ``prec_obs ~ Gamma(3, 1)``,
    ``obs ~ Normal(data @ coefs, prec_obs ** -0.5)`` (likelihood 1)."""
    def bind(data, obs):
        import math

        X = np.ascontiguousarray(data, np.float32)
        y = np.ascontiguousarray(obs, np.float32)
        if X.ndim != 2 or y.shape != (X.shape[0],):
            raise ValueError("expects data [N, D] and observations [N]")
        N, D = X.shape
        const = 0.0
        if likelihood == 0:
            if (y < 0).any() or (y != np.round(y)).any():
                raise ValueError("poisson_regression expects non-negative integer counts")
            const = sum(math.lgamma(float(v) + 1.0) for v in y)  # discrete.py:743, constant in z
        table = np.concatenate([y, [np.float32(const)]]).astype(np.float32)
        sites = [("coefs", (D,))] + ([("prec_obs", ())] if likelihood else [])

        def constrain(z):
            out = {"coefs": z["coefs"]}
            if likelihood:
                out["prec_obs"] = np.exp(z["prec_obs"])
            return out

        return BoundModel(GLM, "poisson_regression" if likelihood == 0 else "linear_regression", N, D, X,
                          table, sites, constrain, likelihood=likelihood,
                          positive=("prec_obs",) if likelihood else ())
    return bind


def _bind_sv(returns):
    r = np.ascontiguousarray(returns, np.float32)
    T = r.shape[0]

    def constrain(z):
        return {"nu": np.exp(z["nu"]), "s": z["s"], "sigma": np.exp(z["sigma"])}

    return BoundModel(SV, "stochastic_volatility", T, 0, r, None,
                      [("sigma", ()), ("s", (T,)), ("nu", ())], constrain,
                      positive=("sigma", "nu"))


def _bind_horseshoe(likelihood):
    def bind(X, Y):
        X = np.ascontiguousarray(X, np.float32)
        Y = np.ascontiguousarray(Y).astype(np.float32)
        if X.ndim != 2 or Y.shape != (X.shape[0],):
            raise ValueError("horseshoe regression expects X [N, D] and Y [N]")
        Dh = X.shape[1]
        sites = [("lambdas", (Dh,)), ("tau", (1,)), ("unscaled_betas", (Dh,))]
        if likelihood == 1:
            sites.append(("prec_obs", ()))

        def constrain(z):
            out = {"lambdas": np.exp(z["lambdas"]), "tau": np.exp(z["tau"]),
                   "unscaled_betas": z["unscaled_betas"]}
            if likelihood == 1:
                out["prec_obs"] = np.exp(z["prec_obs"])
            out["betas"] = out["tau"] * out["lambdas"] * out["unscaled_betas"]
            return out

        return BoundModel(HORSESHOE, "horseshoe", X.shape[0], Dh, X, Y, sites, constrain,
                          likelihood=likelihood,
                          positive=("lambdas", "tau") + (("prec_obs",) if likelihood == 1 else ()))

    return bind


logistic_regression = RegisteredModel(
    "logistic_regression", _bind_logistic,
    "coefs ~ Normal(0, 1)[D]; obs ~ Bernoulli(logits=data @ coefs)  (examples/covtype.py:66-71)")
eight_schools_noncentered = RegisteredModel(
    "eight_schools_noncentered", _bind_eight_schools,
    "Non-centred eight schools with TransformReparam (README.md:101-110)")
eight_schools = RegisteredModel(
    "eight_schools", _bind_eight_schools_locscale(1.0),
    "Centred eight schools: theta ~ Normal(mu, tau) (README.md:60-68)")
beta_bernoulli = RegisteredModel(
    "beta_bernoulli", _bind_beta_bernoulli,
    "p_latent ~ Beta(1.1, 1.1)[K]; obs ~ Bernoulli(p_latent)  (test/infer/test_mcmc.py:239-244)")
dirichlet_categorical = RegisteredModel(
    "dirichlet_categorical", _bind_dirichlet_categorical,
    "p_latent ~ Dirichlet(1, 1, 1); obs ~ Categorical(p_latent)  (test/infer/test_mcmc.py:283-288)")
stochastic_volatility = RegisteredModel(
    "stochastic_volatility", _bind_sv,
    "Random-walk stochastic volatility (examples/stochastic_volatility.py:57-65)")

horseshoe_bernoulli = RegisteredModel(
    "horseshoe_bernoulli", _bind_horseshoe(0),
    "Horseshoe regression, Bernoulli likelihood (examples/horseshoe_regression.py:62-78)")
horseshoe_normal = RegisteredModel(
    "horseshoe_normal", _bind_horseshoe(1),
    "Horseshoe regression, Normal likelihood (examples/horseshoe_regression.py:37-58)")

poisson_regression = RegisteredModel(
    "poisson_regression", _bind_glm(0),
    "coefs ~ Normal(0, 1)[D]; obs ~ Poisson(exp(data @ coefs))")
linear_regression = RegisteredModel(
    "linear_regression", _bind_glm(1),
    "coefs ~ Normal(0, 1)[D]; prec_obs ~ Gamma(3, 1); obs ~ Normal(data @ coefs, prec_obs ** -0.5)")

REGISTRY = {m.__name__: m for m in (logistic_regression, eight_schools_noncentered, eight_schools,
                                    beta_bernoulli, dirichlet_categorical, poisson_regression,
                                    linear_regression,
                                    stochastic_volatility, horseshoe_bernoulli,
                                    horseshoe_normal)}


def reparam(model, config):
    """``numpyro.handlers.reparam(model, config=...)`` for the registered families: the centred
    ``eight_schools`` with ``{"theta": LocScaleReparam(centered)}`` (``reparam.py:64-145``) or
    ``{"theta": TransformReparam()}`` is NOT a rewrite of a traced program here -- it selects the
    closed-form potential of the reparametrised model."""
    from .infer.reparam import LocScaleReparam

    model = resolve(model)
    if model is not eight_schools or set(config) != {"theta"} or not isinstance(config["theta"], LocScaleReparam):
        raise NotImplementedError(
            "reparam is registered for models.eight_schools with config={'theta': LocScaleReparam(c)}")
    c = config["theta"].centered
    name = f"eight_schools[theta=LocScaleReparam({c!r})]"
    if name not in REGISTRY:
        REGISTRY[name] = RegisteredModel(name, _bind_eight_schools_locscale(c),
                                         "eight schools under LocScaleReparam (reparam.py:64-145)")
    return REGISTRY[name]


def lookup(name):
    """Registered model by name (how kernels pickle their model); reparametrised entries are
    re-created on demand in a fresh process."""
    if name not in REGISTRY and name.startswith("eight_schools[theta=LocScaleReparam("):
        from .infer.reparam import LocScaleReparam

        c = float(name[len("eight_schools[theta=LocScaleReparam("):-2])
        reparam(eight_schools, {"theta": LocScaleReparam(c)})
    return REGISTRY[name]


def resolve(model):
    if isinstance(model, RegisteredModel):
        return model
    raise NotImplementedError(
        f"model {getattr(model, '__name__', model)!r} is outside the registered potential families "
        f"{sorted(REGISTRY)}; numpyro_b200 has no generic tracing fallback")
