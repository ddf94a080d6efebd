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

LOGISTIC, EIGHT_SCHOOLS, HORSESHOE, SV, LOGISTIC_ECS, EIGHT_SCHOOLS_LOCSCALE = 0, 1, 2, 3, 4, 5


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
            out[k] = np.log(v) if k in self.positive else v
        return out

    def postprocess(self, zdict):
        """``postprocess_fn``: unconstrained dict -> constrained + deterministic sites
        (``infer/util.py:176-230,583-609``)."""
        return self._constrain(zdict)


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

    return BoundModel(EIGHT_SCHOOLS, "eight_schools", J, 0, y, sigma,
                      [("mu", ()), ("tau", ()), ("theta_base", (J,))], constrain,
                      positive=("tau",))


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
stochastic_volatility = RegisteredModel(
    "stochastic_volatility", _bind_sv,
    "Random-walk stochastic volatility (examples/stochastic_volatility.py:57-65)")

horseshoe_bernoulli = RegisteredModel(
    "horseshoe_bernoulli", _bind_horseshoe(0),
    "Horseshoe regression, Bernoulli likelihood (examples/horseshoe_regression.py:62-78)")
horseshoe_normal = RegisteredModel(
    "horseshoe_normal", _bind_horseshoe(1),
    "Horseshoe regression, Normal likelihood (examples/horseshoe_regression.py:37-58)")

REGISTRY = {m.__name__: m for m in (logistic_regression, eight_schools_noncentered, eight_schools,
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
