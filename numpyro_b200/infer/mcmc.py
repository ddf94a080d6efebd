"""``MCMC`` driver with the surface of ``numpyro/infer/mcmc.py:224-801``.

It drives any ``MCMCKernel`` through ``init`` / ``sample``; a kernel that offers
``run_loop`` (the CUDA kernels do) gets the whole ``fori_collect`` range
(``numpyro/util.py:277-407``) in one launch, like the reference's single jitted
``fori_loop`` with ``progress_bar=False``.
"""
from operator import attrgetter
import warnings

import numpy as np

from .. import random as b2random
from ..diagnostics import print_summary


def _tree_map(fn, tree):
    if isinstance(tree, dict):
        return {k: _tree_map(fn, v) for k, v in tree.items()}
    if isinstance(tree, tuple) and hasattr(tree, "_fields"):
        return type(tree)(*[_tree_map(fn, v) for v in tree])
    if isinstance(tree, (tuple, list)):
        return type(tree)(_tree_map(fn, v) for v in tree)
    if tree is None:
        return None
    return fn(tree)


class MCMC(object):
    def __init__(self, sampler, *, num_warmup, num_samples, num_chains=1, thinning=1,
                 postprocess_fn=None, chain_method="parallel", progress_bar=True,
                 jit_model_args=False):
        self.sampler = sampler
        self._sample_field = sampler.sample_field
        self._default_fields = sampler.default_fields
        self.num_warmup = num_warmup
        self.num_samples = num_samples
        self.num_chains = num_chains
        if not isinstance(thinning, int) or thinning < 1:
            raise ValueError("thinning must be a positive integer")
        self.thinning = thinning
        self.postprocess_fn = postprocess_fn
        if not callable(chain_method) and chain_method not in ["parallel", "vectorized", "sequential"]:
            raise ValueError("Only supporting the following methods to draw chains:"
                             ' "sequential", "parallel", or "vectorized"')
        if callable(chain_method):
            raise NotImplementedError("a callable chain_method (jax transform) has no meaning here")
        # "vectorized": every chain in one batched launch on the current GPU (mcmc.py:709-710).
        # "parallel": the chains are spread over the local GPUs, one handle and one host thread
        # per device (the reference pmaps over devices, mcmc.py:707-708); with fewer devices
        # than chains the reference warns and runs sequentially (mcmc.py:355-366) -- here the
        # same warning is issued and every device takes a contiguous slice of the chains.
        # "sequential": one chain after the other (mcmc.py:704-705).
        # Per-chain results do not depend on the method (test_mcmc.py:848-894).
        if chain_method == "parallel" and num_chains > 1:
            ndev = self._local_device_count()
            if ndev < num_chains:
                warnings.warn(
                    "There are not enough devices to run parallel chains: expected {} but got {}."
                    " Chains will be drawn sequentially. If you are running MCMC in CPU,"
                    " consider using `numpyro.set_host_device_count({})` at the beginning"
                    " of your program. You can double-check how many devices are available in"
                    " your system using `jax.local_device_count()`.".format(num_chains, ndev, num_chains),
                    stacklevel=2)
        self.chain_method = chain_method
        self._device_samplers = None
        self.progress_bar = False  # the loop runs on device in one launch
        self._jit_model_args = jit_model_args
        self._states = None
        self._states_flat = None
        self._last_state = None
        self._warmup_state = None
        self._collection_params = {}
        self._set_collection_params()

    # ------------------------------------------------------------------ collection bookkeeping
    def _set_collection_params(self, lower=None, upper=None, collection_size=None, phase=None):
        self._collection_params["lower"] = self.num_warmup if lower is None else lower
        self._collection_params["upper"] = (self.num_warmup + self.num_samples if upper is None
                                            else upper)
        self._collection_params["collection_size"] = collection_size
        self._collection_params["phase"] = phase

    @property
    def post_warmup_state(self):
        return self._warmup_state

    @post_warmup_state.setter
    def post_warmup_state(self, state):
        self._warmup_state = state

    @property
    def last_state(self):
        return self._last_state

    @staticmethod
    def _local_device_count():
        try:
            from .. import _lib

            return max(1, int(_lib.lib().b200mc_device_count()))
        except Exception:  # library not built / no driver: argument checking still works
            return 1

    def _run_chains_sequential(self, rng_keys, init_state, init_params, args, kwargs, collect_fields,
                               remove_sites):
        """``chain_method="sequential"`` (mcmc.py:704-705): one chain after the other, each a
        one-chain batch; results stacked in chain order."""
        cut = lambda tree, c: _tree_map(lambda x: np.asarray(x)[c:c + 1], tree)  # noqa: E731
        outs = []
        for c in range(rng_keys.shape[0]):
            outs.append(self._run_chains(
                rng_keys[c:c + 1], None if init_state is None else cut(init_state, c),
                None if init_params is None else cut(init_params, c), args, kwargs, collect_fields,
                remove_sites))
        cat = lambda trees, axis: (  # noqa: E731
            {k: cat([t[k] for t in trees], axis) for k in trees[0]} if isinstance(trees[0], dict)
            else type(trees[0])(*[cat([t[i] for t in trees], axis) for i in range(len(trees[0]))])
            if isinstance(trees[0], tuple) and hasattr(trees[0], "_fields")
            else type(trees[0])(cat([t[i] for t in trees], axis) for i in range(len(trees[0])))
            if isinstance(trees[0], (tuple, list)) else None if trees[0] is None
            else np.concatenate([np.asarray(t) for t in trees], axis=axis))
        return cat([o[0] for o in outs], 1), cat([o[1] for o in outs], 0)

    # ------------------------------------------------------------------ chains over local GPUs
    def _run_chains_parallel(self, rng_keys, init_state, init_params, args, kwargs, collect_fields,
                             remove_sites, ndev):
        """``chain_method="parallel"`` with several local GPUs: device d runs the contiguous
        slice ``shard_bounds(C, d, ndev)`` of the chains with its own kernel object (handle) on
        its own host thread; collected fields and the last state are concatenated in chain
        order, so the result equals the one-GPU run chain for chain."""
        import pickle
        import threading

        from .. import _lib
        from ..distributed import shard_bounds

        C = rng_keys.shape[0]
        if self._device_samplers is None or len(self._device_samplers) != ndev:
            blob = pickle.dumps(self.sampler)  # __getstate__ drops everything init() built
            self._device_samplers = [pickle.loads(blob) for _ in range(ndev)]
        bounds = [shard_bounds(C, d, ndev) for d in range(ndev)]
        cut = lambda tree, lo, hi: _tree_map(lambda x: np.asarray(x)[lo:hi], tree)  # noqa: E731
        out, err = [None] * ndev, [None] * ndev

        def work(d):
            lo, hi = bounds[d]
            if hi == lo:
                return
            try:
                _lib.check(_lib.lib().b200mc_set_device(d))  # cudaSetDevice is per host thread
                out[d] = self._run_chains(
                    rng_keys[lo:hi], None if init_state is None else cut(init_state, lo, hi),
                    None if init_params is None else cut(init_params, lo, hi), args, kwargs,
                    collect_fields, remove_sites, sampler=self._device_samplers[d])
            except BaseException as e:  # noqa: BLE001 -- re-raised on the calling thread
                err[d] = e

        threads = [threading.Thread(target=work, args=(d,)) for d in range(ndev)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        _lib.check(_lib.lib().b200mc_set_device(0))
        for e in err:
            if e is not None:
                raise e
        got = [o for o in out if o is not None]

        def cat(trees, axis):
            t0 = trees[0]
            if isinstance(t0, dict):
                return {k: cat([t[k] for t in trees], axis) for k in t0}
            if isinstance(t0, tuple) and hasattr(t0, "_fields"):
                return type(t0)(*[cat([t[i] for t in trees], axis) for i in range(len(t0))])
            if isinstance(t0, (tuple, list)):
                return type(t0)(cat([t[i] for t in trees], axis) for i in range(len(t0)))
            if t0 is None:
                return None
            return np.concatenate([np.asarray(t) for t in trees], axis=axis)

        states = cat([g[0] for g in got], 1)      # [draw, chain, ...]
        last_state = cat([g[1] for g in got], 0)  # [chain, ...]
        return states, last_state

    # ------------------------------------------------------------------ one batch of chains
    def _run_chains(self, rng_key, init_state, init_params, args, kwargs, collect_fields,
                    remove_sites, sampler=None):
        sampler = self.sampler if sampler is None else sampler
        if init_state is None or getattr(sampler, "_sample_fn", None) is None:
            new_init_state = sampler.init(rng_key, self.num_warmup, init_params, model_args=args,
                                          model_kwargs=kwargs)
            init_state = new_init_state if init_state is None else init_state
        lower, upper = self._collection_params["lower"], self._collection_params["upper"]
        collection_size = self._collection_params["collection_size"]
        collection_size = (collection_size if collection_size is None
                           else collection_size // self.thinning)
        thinning = self.thinning
        n_keep = (upper - lower) // thinning if collection_size is None else collection_size
        assert n_keep >= (upper - lower) // thinning
        start_idx = lower + (upper - lower) % thinning  # util.py:330
        postprocess = (self.postprocess_fn if self.postprocess_fn is not None
                       else sampler.postprocess_fn(args, kwargs))

        device_fields = []
        for f in collect_fields:
            if f == self._sample_field or f.startswith(self._sample_field + "."):
                if "z" not in device_fields:
                    device_fields.append("z")
            else:
                device_fields.append(f)
        vectorized = not b2random.is_prng_key(rng_key)

        if hasattr(sampler, "run_loop"):
            raw, last_state = sampler.run_loop(init_state, upper, start_idx, thinning, n_keep,
                                               device_fields, model_args=args, model_kwargs=kwargs)
            if not vectorized:
                raw = {k: v[:, 0] for k, v in raw.items()}
            if "z" in raw:
                raw["z"] = sampler._bound.unravel(raw["z"])
        else:  # generic MCMCKernel: python loop over sample() (util.py:397-403)
            state, rows = init_state, {f: [None] * n_keep for f in device_fields}
            for i in range(upper):
                state = sampler.sample(state, args, kwargs)
                idx = (i - start_idx) // thinning
                if i >= start_idx and 0 <= idx < n_keep:
                    for f in device_fields:
                        rows[f][idx] = attrgetter(f)(state)
            last_state = state
            raw = {}
            for f, v in rows.items():
                if n_keep and isinstance(v[0], dict):
                    raw[f] = {k: np.stack([r[k] for r in v]) for k in v[0]}
                else:
                    raw[f] = np.asarray(v)

        states = {}
        for f in collect_fields:
            if f == self._sample_field:
                sites = postprocess(raw["z"]) if n_keep else raw["z"]
                if remove_sites:
                    sites = {k: v for k, v in sites.items() if k not in remove_sites}
                states[f] = sites
            elif f.startswith(self._sample_field + "."):
                states[f] = raw["z"][f[len(self._sample_field) + 1:]]
            else:
                states[f] = raw[f]
        return states, last_state

    # ------------------------------------------------------------------ public API
    def warmup(self, rng_key, *args, extra_fields=(), collect_warmup=False, init_params=None,
               **kwargs):
        """``mcmc.py:580-624``."""
        self._warmup_state = None
        if collect_warmup:
            self._set_collection_params(0, self.num_warmup, self.num_warmup, "warmup")
        else:
            self._set_collection_params(self.num_warmup, self.num_warmup, self.num_samples, "warmup")
        self.run(rng_key, *args, extra_fields=extra_fields, init_params=init_params, **kwargs)
        self._warmup_state = self._last_state

    def run(self, rng_key, *args, extra_fields=(), init_params=None, **kwargs):
        """``mcmc.py:626-720``."""
        self._args, self._kwargs = args, kwargs
        rng_key = np.asarray(rng_key, np.uint32)
        init_state = None
        if init_params is not None and self.num_chains > 1:
            leaves = list(init_params.values()) if isinstance(init_params, dict) else [init_params]
            if np.shape(leaves[0])[0] != self.num_chains:
                raise ValueError("`init_params` must have the same leading dimension"
                                 " as `num_chains`.")
        if self.num_chains > 1 and b2random.is_prng_key(rng_key):
            rng_key = b2random.split(rng_key, self.num_chains,
                                     getattr(self.sampler, "partitionable", None))
        if self._warmup_state is not None:
            self._set_collection_params(0, self.num_samples, self.num_samples, "sample")
            init_state = self._warmup_state._replace(rng_key=rng_key)
        assert isinstance(extra_fields, (tuple, list))
        collect_fields, remove_sites = {}, {}
        for field_name in ((self._sample_field,) + tuple(self._default_fields) + tuple(extra_fields)):
            if field_name.startswith(f"~{self._sample_field}."):
                remove_sites[field_name[len(self._sample_field) + 2:]] = None
            else:
                collect_fields[field_name] = None
        collect_fields, remove_sites = tuple(collect_fields), tuple(remove_sites)

        ndev = self._local_device_count() if self.chain_method == "parallel" else 1
        vector_keys = not b2random.is_prng_key(rng_key)
        if self.chain_method == "parallel" and vector_keys and ndev > 1 and hasattr(self.sampler, "run_loop"):
            states, last_state = self._run_chains_parallel(rng_key, init_state, init_params, args,
                                                           kwargs, collect_fields, remove_sites,
                                                           min(ndev, rng_key.shape[0]))
        elif self.chain_method == "sequential" and vector_keys:
            states, last_state = self._run_chains_sequential(rng_key, init_state, init_params, args,
                                                             kwargs, collect_fields, remove_sites)
        else:
            states, last_state = self._run_chains(rng_key, init_state, init_params, args, kwargs,
                                                  collect_fields, remove_sites)
        if self.num_chains == 1 and b2random.is_prng_key(rng_key):
            states = _tree_map(lambda x: x[np.newaxis, ...], states)
        else:
            # [draw, chain, ...] -> [chain, draw, ...]  (mcmc.py:713-715)
            states = _tree_map(lambda x: np.swapaxes(x, 0, 1), states)
        self._last_state = last_state
        self._states = states
        self._states_flat = None
        self._set_collection_params()

    def _get_states_flat(self):
        if self._states_flat is None:
            self._states_flat = _tree_map(
                lambda x: np.reshape(x, (x.shape[0] * x.shape[1],) + x.shape[2:]), self._states)
        return self._states_flat

    def get_samples(self, group_by_chain=False):
        return (self._states[self._sample_field] if group_by_chain
                else self._get_states_flat()[self._sample_field])

    def get_extra_fields(self, group_by_chain=False):
        states = self._states if group_by_chain else self._get_states_flat()
        return {k: v for k, v in states.items() if k != self._sample_field}

    def print_summary(self, prob=0.9, exclude_deterministic=True):
        sites = self._states[self._sample_field]
        if isinstance(sites, dict) and exclude_deterministic:
            state_sample_field = attrgetter(self._sample_field)(self._last_state)
            if isinstance(state_sample_field, dict):
                sites = {k: v for k, v in sites.items() if k in state_sample_field}
        print_summary(sites, prob=prob)
        extra_fields = self.get_extra_fields()
        if "diverging" in extra_fields:
            print("Number of divergences: {}".format(np.sum(extra_fields["diverging"])))

    def transfer_states_to_host(self):
        pass  # collected states already live on the host

    def __getstate__(self):
        return self.__dict__.copy()
