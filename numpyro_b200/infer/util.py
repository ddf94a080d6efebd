"""``log_likelihood`` and ``Predictive`` with the surface of ``numpyro/infer/util.py:888-1148``
for the registered model families: the observation site is evaluated (or sampled) over ALL
posterior draws by one CUDA launch (``b200mc_obs_forward``, ``csrc/obs_kernel.cuh``); the host
only lays out the per-draw coefficients.  A model outside the registry raises."""
import warnings

import numpy as np

from .. import _lib, models
from .. import random as b2random

_BERNOULLI, _NORMAL, _STUDENT_T = 0, 1, 2


def _site_plan(model, draws, args, kwargs):
    """-> dict(site, X, B_or_loc [S, .], scale, scale_kind, df, y, likelihood, N, deterministic)."""
    name = models.resolve(model).__name__
    f32 = lambda a: np.ascontiguousarray(a, np.float32)  # noqa: E731
    S = None
    for v in draws.values():
        S = np.shape(v)[0]
        break
    if name == "logistic_regression":
        X = f32(args[0])
        y = None if len(args) < 2 or args[1] is None else f32(args[1])
        return dict(site="obs", X=X, B=f32(draws["coefs"]).reshape(S, X.shape[1]), scale=None, scale_kind=0,
                    df=None, y=y, likelihood=_BERNOULLI, N=X.shape[0], det={})
    if name in ("horseshoe_bernoulli", "horseshoe_normal"):
        X = f32(args[0])
        y = None if len(args) < 2 or args[1] is None else f32(args[1])
        betas = f32(np.asarray(draws["tau"], np.float32).reshape(S, 1)
                    * np.asarray(draws["lambdas"], np.float32) * np.asarray(draws["unscaled_betas"], np.float32))
        normal = name == "horseshoe_normal"
        scale = f32(1.0 / np.sqrt(np.asarray(draws["prec_obs"], np.float32))).reshape(S) if normal else None
        return dict(site="Y", X=X, B=betas, scale=scale, scale_kind=1 if normal else 0, df=None, y=y,
                    likelihood=_NORMAL if normal else _BERNOULLI, N=X.shape[0], det={"betas": betas})
    if name == "linear_regression":  # the Normal GLM: the horseshoe example's observation model on coefs
        X = f32(args[0])
        y = None if len(args) < 2 or args[1] is None else f32(args[1])
        scale = f32(1.0 / np.sqrt(np.asarray(draws["prec_obs"], np.float32))).reshape(S)
        return dict(site="obs", X=X, B=f32(draws["coefs"]).reshape(S, X.shape[1]), scale=scale, scale_kind=1,
                    df=None, y=y, likelihood=_NORMAL, N=X.shape[0], det={})
    if name == "eight_schools_noncentered":
        sigma = f32(args[1])
        y = kwargs.get("y", args[2] if len(args) > 2 else None)
        theta = f32(np.asarray(draws["mu"], np.float32).reshape(S, 1)
                    + np.asarray(draws["tau"], np.float32).reshape(S, 1) * np.asarray(draws["theta_base"], np.float32))
        return dict(site="obs", X=None, B=theta, scale=sigma, scale_kind=2, df=None,
                    y=None if y is None else f32(y), likelihood=_NORMAL, N=sigma.shape[0], det={"theta": theta})
    if name == "eight_schools" or name.startswith("eight_schools[theta=LocScaleReparam("):
        # the centred model, possibly under LocScaleReparam(c) (reparam.py:64-145): theta is the
        # latent itself (c = 1) or theta = mu + tau^(1-c) (theta_decentered - c mu)
        sigma = f32(args[1])
        y = kwargs.get("y", args[2] if len(args) > 2 else None)
        if "theta" in draws:
            theta = f32(draws["theta"]).reshape(S, sigma.shape[0])
        else:
            c = np.float32(name[len("eight_schools[theta=LocScaleReparam("):-2])
            mu = np.asarray(draws["mu"], np.float32).reshape(S, 1)
            tau = np.asarray(draws["tau"], np.float32).reshape(S, 1)
            theta = f32(mu + np.power(tau, np.float32(1) - c)
                        * (np.asarray(draws["theta_decentered"], np.float32) - c * mu))
        return dict(site="obs", X=None, B=theta, scale=sigma, scale_kind=2, df=None,
                    y=None if y is None else f32(y), likelihood=_NORMAL, N=sigma.shape[0],
                    det={} if name == "eight_schools" else {"theta": theta})
    if name == "stochastic_volatility":
        r = f32(args[0])
        s = f32(draws["s"]).reshape(S, r.shape[0])
        return dict(site="r", X=None, B=np.zeros_like(s), scale=f32(np.exp(s)), scale_kind=3,
                    df=f32(draws["nu"]).reshape(S), y=r, likelihood=_STUDENT_T, N=r.shape[0], det={})
    raise NotImplementedError(name)


def _forward(plan, keys, partitionable):
    lib = _lib.lib()
    _lib.require_device()
    S, N = plan["B"].shape[0], plan["N"]
    dev = lambda a: None if a is None else _lib.DeviceArray.from_numpy(a)  # noqa: E731
    X, B, sc, df, y = dev(plan["X"]), dev(plan["B"]), dev(plan["scale"]), dev(plan["df"]), dev(plan["y"])
    kd = dev(None if keys is None else np.ascontiguousarray(keys, np.uint32))
    sampled_int = keys is not None and plan["likelihood"] == _BERNOULLI
    out = _lib.DeviceArray((S, N), np.int32 if sampled_int else np.float32)
    ptr = lambda d: None if d is None else d.ptr  # noqa: E731
    _lib.check(lib.b200mc_obs_forward(ptr(X), N, 0 if plan["X"] is None else plan["X"].shape[1], B.ptr, S,
                                      ptr(sc), plan["scale_kind"], ptr(df), ptr(y), ptr(kd),
                                      plan["likelihood"], int(partitionable), out.ptr, None))
    return out.numpy()


def _batch(posterior_samples, batch_ndims):
    """Flatten the leading ``batch_ndims`` axes of every site (infer/util.py:1124-1139)."""
    batch_shape = None
    flat = {}
    for name, v in posterior_samples.items():
        v = np.asarray(v)
        if batch_shape is not None and v.shape[:batch_ndims] != batch_shape:
            raise ValueError(f"Batch shapes at site {name} and {proto} should be the same, but got "
                             f"{v.shape[:batch_ndims]} and {batch_shape}")
        proto, batch_shape = name, v.shape[:batch_ndims]
        flat[name] = v.reshape((-1,) + v.shape[batch_ndims:])
    return flat, batch_shape


def log_likelihood(model, posterior_samples, *args, parallel=False, batch_ndims=1, **kwargs):
    """``numpyro/infer/util.py:1094-1148``: dict of log likelihoods at the observation sites."""
    flat, batch_shape = _batch(posterior_samples, batch_ndims)
    plan = _site_plan(model, flat, args, kwargs)
    if plan["y"] is None:
        return {}
    ll = _forward(plan, None, True)
    return {plan["site"]: ll.reshape(batch_shape + ll.shape[1:])}


class Predictive(object):
    """``numpyro/infer/util.py:888-1091`` for the registered families: the posterior draws are
    substituted, the observation site (given as ``None``) is sampled.  ``guide`` / ``params`` /
    ``infer_discrete`` belong to SVI and enumeration, outside the registered hot path."""

    def __init__(self, model, posterior_samples=None, *, guide=None, params=None, num_samples=None,
                 return_sites=None, infer_discrete=False, parallel=False, batch_ndims=None,
                 exclude_deterministic=True, partitionable=None):
        if guide is not None or params or infer_discrete:
            raise NotImplementedError("Predictive with a guide / params / infer_discrete is outside the "
                                      "registered hot path")
        if posterior_samples is None and num_samples is None:
            raise ValueError("Either posterior_samples or num_samples must be specified.")
        if not posterior_samples:
            raise NotImplementedError("prior predictive sampling (no posterior_samples) is not registered")
        batch_ndims = 1 if batch_ndims is None else batch_ndims
        self.model = models.resolve(model)
        self.posterior_samples, self._batch_shape = _batch(posterior_samples, batch_ndims)
        batch_size = int(np.prod(self._batch_shape))
        if num_samples is not None and num_samples != batch_size:
            warnings.warn("Sample's batch dimension size {} is different from the provided {} num_samples "
                          "argument. Defaulting to {}.".format(batch_size, num_samples, batch_size),
                          UserWarning, stacklevel=2)
        self.num_samples = batch_size
        if return_sites is not None:
            assert isinstance(return_sites, (list, tuple, set))
        self.return_sites = return_sites
        self.batch_ndims = batch_ndims
        self.exclude_deterministic = exclude_deterministic
        self.partitionable = partitionable

    def __call__(self, rng_key, *args, **kwargs):
        part = b2random.DEFAULT_PARTITIONABLE if self.partitionable is None else self.partitionable
        plan = _site_plan(self.model, self.posterior_samples, args, kwargs)
        if plan["y"] is not None:
            # the site is observed: nothing to sample (infer/util.py:905-907); deterministic sites only
            out = {}
        else:
            S = self.num_samples
            rng_key = np.asarray(rng_key, np.uint32)
            keys = b2random.split(rng_key, S, part) if S > 1 else rng_key[None]  # infer/util.py:874-877
            # seed(model, key_s): the observation site is the first sample statement without a value,
            # it takes the sub-key of the first split (handlers.py seed.process_message)
            site_keys = b2random.split(keys, 2, part)[:, 1]
            out = {plan["site"]: _forward(plan, site_keys, part)}
        out.update(plan["det"])
        if self.return_sites is not None:
            out = {k: v for k, v in {**out, **self.posterior_samples}.items() if k in set(self.return_sites)}
        return {k: np.asarray(v).reshape(self._batch_shape + np.shape(v)[1:]) for k, v in out.items()}
