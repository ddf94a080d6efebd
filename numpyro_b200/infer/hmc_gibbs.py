"""``HMCGibbs`` / ``HMCECS`` (``numpyro/infer/hmc_gibbs.py:38-190,502-688``) around the CUDA
NUTS/HMC kernels.

``HMCECS`` -- HMC with energy-conserving subsampling, the default algorithm of the reference's
covtype example (``examples/covtype.py:154-165``) -- is built for the registered logistic model
with a subsampled plate (``models.logistic_regression(data, labels, subsample_size=m)``): the
potential on the current subsample (with or without the Taylor control variate of
``contrib/ecs_proxies.py``), the block update of the subsample and its Metropolis test all run
on the device (``csrc/ecs_kernel.cuh``, ``csrc/models.cuh``); this module holds the reference's
key handling and state threading.

``HMCGibbs`` with a user ``gibbs_fn``: the Gibbs sites are conditioned sites of the registered
family (``models.BoundModel.condition``; on the device the family's potential is evaluated on the
full vector with those coordinates held, ``csrc/models.cuh`` ``model_potential``), the HMC/NUTS
transition over the remaining sites is the CUDA one, and ``gibbs_fn`` is the caller's Python
function on host arrays -- one host round trip per transition, as a Python ``gibbs_fn`` implies.
"""
import copy
import ctypes as C
from collections import namedtuple

import numpy as np

from .. import _lib, models
from .. import random as b2random
from .hmc import HMC, MCMCKernel

HMCGibbsState = namedtuple("HMCGibbsState", "z, hmc_state, rng_key")
HMCECSState = namedtuple("HMCECSState", "z, hmc_state, rng_key, gibbs_state, accept_prob")

TaylorProxy = namedtuple("TaylorProxy", "reference_params, degree")


def taylor_proxy(reference_params, degree=2):
    """``contrib/ecs_proxies.py:94-104``: Taylor control variate around ``reference_params``
    (constrained values per site).  ``degree`` 2 is the second-order proxy; any other value gives
    the first-order one, as in the reference (``:160-165``)."""
    return TaylorProxy(dict(reference_params), 2 if degree == 2 else 1)


def _identity(x):
    return x


class _ConditionedModel:
    """``_wrap_model`` (``hmc_gibbs.py:30-34``) for a registered family: binding gives the view
    with the Gibbs sites conditioned."""

    def __init__(self, model, sites):
        self._model, self._sites = model, list(sites)
        self.__name__ = getattr(model, "__name__", "model")

    def bind(self, *args, **kwargs):
        return self._model.bind(*args, **kwargs).condition(self._sites)


class HMCGibbs(MCMCKernel):
    """``hmc_gibbs.py:38-190``."""

    sample_field = "z"

    def __init__(self, inner_kernel, gibbs_fn, gibbs_sites):
        if not isinstance(inner_kernel, HMC):
            raise ValueError("inner_kernel must be a HMC or NUTS sampler.")
        if not callable(gibbs_fn):
            raise ValueError("gibbs_fn must be a callable")
        assert inner_kernel.model is not None, (
            "HMCGibbs does not support models specified via a potential function.")
        self.inner_kernel = copy.copy(inner_kernel)
        # hmc_gibbs.py:99-100: the inner kernel sees the model with the Gibbs sites conditioned
        self._model0 = inner_kernel._model
        if gibbs_fn is not _identity:
            self.inner_kernel._model = _ConditionedModel(inner_kernel._model, gibbs_sites)
        self._gibbs_sites = gibbs_sites
        self._gibbs_fn = gibbs_fn
        self._prototype_trace = None
        self._sample_fn = None

    @property
    def model(self):
        return self.inner_kernel._model

    @property
    def default_fields(self):
        return ("z",)

    def get_diagnostics_str(self, state):
        state = state.hmc_state
        return "{} steps of size {:.2e}. acc. prob={:.2f}".format(
            state.num_steps, state.adapt_state.step_size, state.mean_accept_prob)

    def postprocess_fn(self, args, kwargs):
        """``hmc_gibbs.py:109-119``: the HMC sites through the inner kernel's ``postprocess_fn``
        with the Gibbs sites substituted; the Gibbs sites as they are (constrained)."""
        full = self._model0.bind(*args, **(kwargs or {}))
        view = full.condition(self._gibbs_sites)

        def fn(z):
            hmc_sites = {k: v for k, v in z.items() if k not in self._gibbs_sites}
            gibbs_sites = {k: v for k, v in z.items() if k in self._gibbs_sites}
            flat, _ = view.cond_unconstrained(gibbs_sites)
            off, un = 0, {}
            for k, (_, n, shp) in view.cond_sites.items():
                un[k] = flat[..., off:off + n].reshape(flat.shape[:-1] + shp)
                off += n
            out = full.postprocess({**hmc_sites, **un})
            hmc_sites = {k: v for k, v in out.items() if k not in self._gibbs_sites}
            return {**gibbs_sites, **hmc_sites}

        return fn

    # ------------------------------------------------------------------ conditioned values
    def _push_gibbs(self, gibbs_sites):
        """Constrained Gibbs-site values (``[...]`` or ``[C, ...]`` per site) -> the device."""
        inner = self.inner_kernel
        view, Cn = inner._bound, inner._num_chains
        flat, off = view.cond_unconstrained(gibbs_sites)
        flat = np.ascontiguousarray(np.broadcast_to(flat.reshape(-1, flat.shape[-1]), (Cn, flat.shape[-1])),
                                    np.float32)
        off = np.ascontiguousarray(np.broadcast_to(np.reshape(off, (-1,)), (Cn,)), np.float32)
        fd, od = _lib.DeviceArray.from_numpy(flat), _lib.DeviceArray.from_numpy(off)
        _lib.check(_lib.lib().b200mc_set_conditioned(inner._handle, fd.ptr, od.ptr, None))
        _lib.check(_lib.lib().b200mc_stream_synchronize(None))

    def _prior_sample(self, full, name, key, part):
        """``init_to_sample`` (``initialization.py:88-93``): ONE prior draw of the site, its median
        over that one draw.  Registered priors: ``Normal(loc, scale)`` (``continuous.py``
        ``Normal.sample``: ``loc + scale * normal(key, (1,) + shape)``)."""
        prior = getattr(full, "prior", {}).get(name)
        if prior is None or prior[0] != "normal":
            raise NotImplementedError(
                f"HMCGibbs: no registered prior sampler for the Gibbs site {name!r}; pass its initial "
                "value in `init_params`")
        _, loc, scale = prior
        shape = dict(full.trace_sites)[name]
        eps = b2random.normal(key, (1,) + tuple(shape), part)
        return (np.float32(loc) + np.float32(scale) * eps)[0].astype(np.float32)

    def init(self, rng_key, num_warmup, init_params, model_args, model_kwargs):
        """``hmc_gibbs.py:121-152``."""
        inner = self.inner_kernel
        part = inner._part()
        model_kwargs = {} if model_kwargs is None else dict(model_kwargs)
        rng_key = np.asarray(rng_key, np.uint32)
        vectorized = not b2random.is_prng_key(rng_key)
        keys = rng_key.reshape(-1, 2)
        full = self._model0.bind(*model_args, **model_kwargs)
        init_params = None if init_params is None else dict(init_params)
        first = self._prototype_trace is None
        proto = [dict() for _ in keys]
        out_keys, keys_z = [], []
        for c, key in enumerate(keys):
            if first:
                key, key_u = b2random.split(key, 2, part)
                k = key_u  # seed(model, key_u): one split per latent site, trace order
                for name, _ in full.trace_sites:
                    k, ks = b2random.split(k, 2, part)
                    if name in self._gibbs_sites and not (init_params and name in init_params):
                        proto[c][name] = self._prior_sample(full, name, ks, part)
            key, key_z = b2random.split(key, 2, part)
            out_keys.append(key)
            keys_z.append(key_z)
        if first:
            self._prototype_trace = proto
        gibbs_sites = {}
        for name, _ in full.trace_sites:
            if name not in self._gibbs_sites:
                continue
            if init_params and name in init_params:
                gibbs_sites[name] = np.asarray(init_params.pop(name), np.float32)
            else:
                v = np.stack([self._prototype_trace[min(c, len(self._prototype_trace) - 1)][name]
                              for c in range(len(keys))])
                gibbs_sites[name] = v if vectorized else v[0]
        inner._ecs = {"degree": 0, "ref": None,
                      "after_create": lambda kern: self._push_gibbs(gibbs_sites)}
        key_z = np.stack(keys_z) if vectorized else keys_z[0]
        hmc_state = inner.init(key_z, num_warmup, init_params or None, model_args=model_args,
                               model_kwargs=model_kwargs)
        self._sample_fn = True
        rk = np.stack(out_keys) if vectorized else out_keys[0]
        return HMCGibbsState({**gibbs_sites, **hmc_state.z}, hmc_state, rk)

    def sample(self, state, model_args, model_kwargs):
        """``hmc_gibbs.py:154-190``: new Gibbs-site values from ``gibbs_fn``, the HMC state's
        ``potential_energy`` / ``z_grad`` re-evaluated under them, one HMC/NUTS transition."""
        inner = self.inner_kernel
        part = inner._part()
        ks = b2random.split(state.rng_key, 2, part)
        rng_key, rng_gibbs = ks[..., 0, :], ks[..., 1, :]
        z_gibbs = {k: v for k, v in state.z.items() if k not in state.hmc_state.z}
        z_hmc = {k: v for k, v in state.z.items() if k in state.hmc_state.z}
        post = self.postprocess_fn(model_args, model_kwargs)({**z_gibbs, **z_hmc})
        z_hmc_c = {k: v for k, v in post.items() if k not in z_gibbs}
        z_gibbs = self._gibbs_fn(rng_key=rng_gibbs, gibbs_sites=z_gibbs, hmc_sites=z_hmc_c)
        z_gibbs = {k: np.asarray(v, np.float32) for k, v in z_gibbs.items()}
        self._push_gibbs(z_gibbs)
        view = inner._bound
        flat = view.ravel(state.hmc_state.z).reshape(inner._num_chains, view.dim)
        pe, grad = inner.potential_and_grad(flat)
        if not inner._vectorized:
            pe, grad = pe[0], grad[0]
        hmc_state = state.hmc_state._replace(z_grad=view.unravel(grad),
                                             potential_energy=np.asarray(pe, np.float32))
        hmc_state = inner.sample(hmc_state, model_args, model_kwargs)
        return HMCGibbsState({**z_gibbs, **hmc_state.z}, hmc_state, rng_key)

    def __getstate__(self):
        state = self.__dict__.copy()
        state["_prototype_trace"] = None
        state["_sample_fn"] = None
        state["_model0"] = self._model0.__name__  # registry entries hold closures: pickle by name
        return state

    def __setstate__(self, state):
        self.__dict__.update(state)
        if isinstance(self._model0, str):
            self._model0 = models.lookup(self._model0)
        if self._gibbs_fn is not _identity:  # the inner kernel came back with the plain model
            self.inner_kernel._model = _ConditionedModel(self._model0, self._gibbs_sites)


class HMCECS(HMCGibbs):
    """``hmc_gibbs.py:502-665``: ``HMCECS(NUTS(model), num_blocks=1, proxy=None)``."""

    def __init__(self, inner_kernel, *, num_blocks=1, proxy=None):
        super().__init__(inner_kernel, _identity, None)
        if proxy is not None and not isinstance(proxy, TaylorProxy):
            raise NotImplementedError("proxy must be None or HMCECS.taylor_proxy(reference_params)")
        self._num_blocks = int(num_blocks)
        self._proxy = proxy
        self._last = None
        # sample(): one iteration of the device-resident loop (b200mc_ecs_sample); False: the
        # stand-alone Gibbs kernel followed by one inner transition (b200mc_ecs_update + b200mc_sample)
        self._fused_step = True

    @staticmethod
    def taylor_proxy(reference_params, degree=2):
        """``hmc_gibbs.py:667-673``."""
        return taylor_proxy(reference_params, degree)

    def postprocess_fn(self, args, kwargs):
        def fn(z):  # hmc_gibbs.py:535-544: the subsample indices are dropped
            hmc_sites = {k: v for k, v in z.items() if k not in self._gibbs_sites}
            return self.inner_kernel.postprocess_fn(args, kwargs)(hmc_sites)

        return fn

    # ------------------------------------------------------------------ device plumbing
    def _get_subsample(self):
        inner = self.inner_kernel
        out = _lib.DeviceArray((inner._num_chains, inner._bound.subsample_size), np.int32)
        _lib.check(_lib.lib().b200mc_ecs_get_subsample(inner._handle, out.ptr, None))
        idx = out.numpy()
        return idx if inner._vectorized else idx[0]

    def _set_subsample(self, idx):
        inner = self.inner_kernel
        a = np.ascontiguousarray(idx, np.int32).reshape(inner._num_chains, inner._bound.subsample_size)
        d = _lib.DeviceArray.from_numpy(a)
        _lib.check(_lib.lib().b200mc_ecs_set_subsample(inner._handle, d.ptr, None))
        _lib.check(_lib.lib().b200mc_stream_synchronize(None))

    # ------------------------------------------------------------------ MCMCKernel interface
    def init(self, rng_key, num_warmup, init_params=None, model_args=(), model_kwargs=None):
        model_kwargs = {} if model_kwargs is None else dict(model_kwargs)
        inner = self.inner_kernel
        part = inner._part()
        rng_key = np.asarray(rng_key, np.uint32)
        bound = inner._model.bind(*model_args, **model_kwargs)
        # hmc_gibbs.py:553-561: the plates with size > subsample_size are the Gibbs sites
        self._gibbs_sites = [bound.plate_name] if bound.family == models.LOGISTIC_ECS else []
        assert self._gibbs_sites, "Cannot detect any subsample statements in the model."
        ks = b2random.split(rng_key, 2, part)  # hmc_gibbs.py:547
        rng_key, key_u = ks[..., 0, :], ks[..., 1, :]
        degree, ref = 0, None
        if self._proxy is not None:
            degree = self._proxy.degree
            ref = bound.unconstrain(self._proxy.reference_params)
            rng_key = b2random.split(rng_key, 2, part)[..., 0, :]  # :594, gibbs_init's key is unused
        ks = b2random.split(rng_key, 2, part)  # HMCGibbs.init, hmc_gibbs.py:135
        rng_key, key_z = ks[..., 0, :], ks[..., 1, :]

        def after_create(kernel):
            kd = _lib.DeviceArray.from_numpy(np.ascontiguousarray(key_u, np.uint32).reshape(-1, 2))
            _lib.check(_lib.lib().b200mc_ecs_init_subsample(kernel._handle, kd.ptr, None))
            if init_params is not None and bound.plate_name in init_params:  # hmc_gibbs.py:140-141
                self._set_subsample(init_params[bound.plate_name])

        inner._ecs = {"degree": degree, "ref": ref, "after_create": after_create}
        hmc_init = None
        if init_params is not None:
            hmc_init = {k: v for k, v in init_params.items() if k != bound.plate_name} or None
        hmc_state = inner.init(key_z, num_warmup, hmc_init, model_args, model_kwargs)
        z = {bound.plate_name: self._get_subsample(), **hmc_state.z}
        acc = np.zeros(rng_key.shape[:-1], np.float32)
        self._sample_fn = self.sample
        self._last = HMCECSState(z, hmc_state, np.ascontiguousarray(rng_key), (), acc)
        return self._last

    def sample(self, state, model_args, model_kwargs):
        """``hmc_gibbs.py:617-665``: block update of the subsample + Metropolis test
        (``b200mc_ecs_update``), then one transition of the inner kernel on that subsample."""
        inner, lib = self.inner_kernel, _lib.lib()
        if inner._handle is None:
            raise RuntimeError("call `init` before `sample`")
        name = self._gibbs_sites[0]
        if state is not self._last:
            inner._push_state(state.hmc_state)
            self._set_subsample(state.z[name])
        Cn = inner._num_chains
        keys = _lib.DeviceArray.from_numpy(np.ascontiguousarray(state.rng_key, np.uint32).reshape(Cn, 2))
        acc = _lib.DeviceArray((Cn,), np.float32)
        if self._fused_step:
            _lib.check(lib.b200mc_ecs_sample(inner._handle, keys.ptr, self._num_blocks, 1, 1, 1, None,
                                             acc.ptr, None, None))
        else:
            _lib.check(lib.b200mc_ecs_update(inner._handle, keys.ptr, self._num_blocks, acc.ptr, None))
            _lib.check(lib.b200mc_sample(inner._handle, 1, 1, 1, None, None))
        hmc_state = inner._to_state(*inner._snapshot())
        rng_key, ap = keys.numpy(), acc.numpy()
        if not inner._vectorized:
            rng_key, ap = rng_key[0], ap[0]
        z = {name: self._get_subsample(), **hmc_state.z}
        self._last = HMCECSState(z, hmc_state, rng_key, (), ap)
        return self._last

    # ------------------------------------------------------------------ fused loop (fori_collect)
    @property
    def _bound(self):
        return self.inner_kernel._bound

    def run_loop(self, state, num_iters, collect_start, thinning, n_keep, fields,
                 model_args=(), model_kwargs=None):
        """``num_iters`` iterations of { Gibbs step; inner transition } in ONE launch
        (``b200mc_ecs_sample``), collecting like ``fori_collect`` (``numpyro/util.py:277-407``).
        ``fields``: ``"z"`` (the HMC sites, flat), ``"accept_prob"`` (of the subsample update) and
        ``"hmc_state.<field>"``.  Returns ``(dict of [n_keep, C, ...] arrays, last state)``."""
        inner, lib = self.inner_kernel, _lib.lib()
        if inner._handle is None:
            raise RuntimeError("call `init` before `run_loop`")
        name = self._gibbs_sites[0]
        if state is not self._last:
            inner._push_state(state.hmc_state)
            self._set_subsample(state.z[name])
        Cn = inner._num_chains
        want_acc = "accept_prob" in fields
        inner_fields = []
        for f in fields:
            if f == "z":
                inner_fields.append("z")
            elif f.startswith("hmc_state."):
                inner_fields.append(f[len("hmc_state."):])
            elif f != "accept_prob":
                raise ValueError(f"HMCECS cannot collect field {f!r}")
        bufs, col = inner._collect_buffers(n_keep, inner_fields)
        keys = _lib.DeviceArray.from_numpy(np.ascontiguousarray(state.rng_key, np.uint32).reshape(Cn, 2))
        acc = _lib.DeviceArray((Cn,), np.float32).zero_()
        acc_all = _lib.DeviceArray((int(num_iters), Cn), np.float32).zero_() if want_acc else None
        _lib.check(lib.b200mc_ecs_sample(inner._handle, keys.ptr, self._num_blocks, int(num_iters),
                                         int(collect_start), int(thinning), C.byref(col), acc.ptr,
                                         acc_all.ptr if acc_all is not None else None, None))
        hmc_state = inner._to_state(*inner._snapshot())
        rng_key, ap = keys.numpy(), acc.numpy()
        if not inner._vectorized:
            rng_key, ap = rng_key[0], ap[0]
        z = {name: self._get_subsample(), **hmc_state.z}
        self._last = HMCECSState(z, hmc_state, rng_key, (), ap)
        out = {}
        for f in fields:
            if f == "z":
                out[f] = bufs["z"].numpy()
            elif f.startswith("hmc_state."):
                v = bufs[f[len("hmc_state."):]].numpy()
                out[f] = v.astype(bool) if f.endswith("diverging") else v
        if want_acc:  # fori_collect keeps the LAST iteration of every thinning block
            a = acc_all.numpy()
            rows = collect_start + thinning - 1 + thinning * np.arange(n_keep)
            out["accept_prob"] = a[rows[rows < num_iters]]
        return out, self._last
