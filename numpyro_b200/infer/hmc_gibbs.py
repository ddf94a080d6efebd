"""``HMCGibbs`` / ``HMCECS`` (``numpyro/infer/hmc_gibbs.py:38-190,502-688``) around the CUDA
NUTS/HMC kernels.

``HMCECS`` -- HMC with energy-conserving subsampling, the default algorithm of the reference's
covtype example (``examples/covtype.py:154-165``) -- is built for the registered logistic model
with a subsampled plate (``models.logistic_regression(data, labels, subsample_size=m)``): the
potential on the current subsample (with or without the Taylor control variate of
``contrib/ecs_proxies.py``), the block update of the subsample and its Metropolis test all run
on the device (``csrc/ecs_kernel.cuh``, ``csrc/models.cuh``); this module holds the reference's
key handling and state threading.

``HMCGibbs`` with a user ``gibbs_fn`` needs a potential that is a function of arbitrary
conditioned sites; the registered closed-form potentials have none, so ``HMCGibbs.init``
raises -- the class carries the shared surface (constructor checks, state, diagnostics string).
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
        def fn(z):
            hmc_sites = {k: v for k, v in z.items() if k not in self._gibbs_sites}
            gibbs_sites = {k: v for k, v in z.items() if k in self._gibbs_sites}
            hmc_sites = self.inner_kernel.postprocess_fn(args, kwargs)(hmc_sites)
            return {**gibbs_sites, **hmc_sites}

        return fn

    def init(self, rng_key, num_warmup, init_params, model_args, model_kwargs):
        raise NotImplementedError(
            "HMCGibbs with a user gibbs_fn needs a potential conditioned on arbitrary sites; the "
            "registered closed-form potentials have none (use HMCECS for subsampled plates)")

    def sample(self, state, model_args, model_kwargs):
        raise NotImplementedError("see HMCGibbs.init")

    def __getstate__(self):
        state = self.__dict__.copy()
        state["_prototype_trace"] = None
        state["_sample_fn"] = None
        return state


def _identity(x):
    return x


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
