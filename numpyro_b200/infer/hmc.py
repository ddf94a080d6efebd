"""``HMC`` / ``NUTS`` kernels with the constructor and ``MCMCKernel`` surface of
``numpyro/infer/hmc.py:541-948``, running on the CUDA library.

The transition itself (``hmc.py:459-530`` and everything below it in
``hmc_util.py``) executes inside persistent CUDA kernels; this module only mirrors the
reference's plugin interface: argument checking, ``HMCState`` pytrees, key handling.
"""
from collections import namedtuple
import ctypes as C
import math
import warnings

import numpy as np

from .. import _lib, models
from .. import random as b2random

HMCState = namedtuple(  # hmc.py:31-48
    "HMCState",
    ["i", "z", "z_grad", "potential_energy", "energy", "r", "trajectory_length", "num_steps",
     "accept_prob", "mean_accept_prob", "diverging", "adapt_state", "rng_key"],
)
HMCAdaptState = namedtuple(  # hmc_util.py:18-30
    "HMCAdaptState",
    ["step_size", "inverse_mass_matrix", "mass_matrix_sqrt", "mass_matrix_sqrt_inv", "ss_state",
     "mm_state", "window_idx", "rng_key"],
)

_VEC = {"z": 0, "z_grad": 1, "inverse_mass_matrix": 2, "mass_matrix_sqrt": 3, "welford_mean": 4,
        "welford_m2": 5}
_MAT = {"inverse_mass_matrix": 6, "mass_matrix_sqrt": 7, "mass_matrix_sqrt_inv": 8, "welford_m2": 9}
EXTRA_FIELDS = ("potential_energy", "energy", "num_steps", "accept_prob", "mean_accept_prob",
                "diverging", "adapt_state.step_size")


def _args_signature(args, kwargs):
    """Cheap identity of model arguments: arrays by (address, shape, dtype, a few bytes)."""
    def one(a):
        if isinstance(a, np.ndarray):
            flat = a.reshape(-1)
            probe = flat[:: max(1, flat.size // 64)][:64].tobytes() if flat.size else b""
            return ("nd", a.__array_interface__["data"][0], a.shape, str(a.dtype), probe)
        try:
            hash(a)
            return ("h", a)
        except TypeError:
            return ("nd*", np.asarray(a).tobytes())
    return (tuple(one(a) for a in args), tuple((k, one(v)) for k, v in sorted(kwargs.items())))


def init_to_uniform(site=None, radius=2):
    """``numpyro/infer/initialization.py:95-129`` (marker: the draws happen on device)."""
    return ("uniform", float(radius))


def init_to_value(site=None, values={}):
    """``numpyro/infer/initialization.py`` ``init_to_value``: constrained values per site."""
    return ("value", dict(values))


class MCMCKernel:
    """The abstract plugin surface ``MCMC`` drives (``numpyro/infer/mcmc.py:32-158``)."""

    def postprocess_fn(self, model_args, model_kwargs):
        return lambda x: x

    def init(self, rng_key, num_warmup, init_params, model_args, model_kwargs):
        raise NotImplementedError

    def sample(self, state, model_args, model_kwargs):
        raise NotImplementedError

    @property
    def sample_field(self):
        raise NotImplementedError

    @property
    def default_fields(self):
        return (self.sample_field,)

    @property
    def is_ensemble_kernel(self):
        return False

    def get_diagnostics_str(self, state):
        return ""


class HMC(MCMCKernel):
    """Hamiltonian Monte Carlo with fixed trajectory length (``hmc.py:541-822``)."""

    def __init__(self, model=None, potential_fn=None, kinetic_fn=None, step_size=1.0,
                 inverse_mass_matrix=None, adapt_step_size=True, adapt_mass_matrix=True,
                 dense_mass=False, target_accept_prob=0.8, num_steps=None,
                 trajectory_length=2 * math.pi, init_strategy=init_to_uniform,
                 find_heuristic_step_size=False, forward_mode_differentiation=False,
                 regularize_mass_matrix=True):
        if not (model is None) ^ (potential_fn is None):
            raise ValueError("Only one of `model` or `potential_fn` must be specified.")
        if type(self) is HMC:
            if (num_steps is None) & (trajectory_length is None):
                raise ValueError(
                    "At least one of `num_steps` or `trajectory_length` must be specified.")
            if adapt_step_size & (num_steps is not None) & (trajectory_length is not None):
                warnings.warn("If both `num_steps` and `trajectory_length` are specified step size "
                              "can't be adapted", stacklevel=2)
        if potential_fn is not None:
            raise NotImplementedError(
                "an arbitrary `potential_fn` is outside the registered potential families; "
                "numpyro_b200 has no generic autodiff fallback")
        if kinetic_fn is not None:
            raise NotImplementedError("only the euclidean kinetic energy is implemented")
        if not isinstance(dense_mass, bool):  # structured form, hmc.py:605-616
            if not isinstance(dense_mass, (list, tuple)) or not all(
                    isinstance(b, tuple) and all(isinstance(k, str) for k in b) for b in dense_mass):
                raise ValueError("dense_mass must be a bool or a list of tuples of site names")
            dense_mass = [tuple(b) for b in dense_mass]
        self._model = models.resolve(model)
        self._potential_fn = None
        self._kinetic_fn = None
        self._num_steps = num_steps
        self._step_size = float(step_size) if isinstance(step_size, int) else step_size
        self._inverse_mass_matrix = inverse_mass_matrix
        self._adapt_step_size = adapt_step_size
        self._adapt_mass_matrix = adapt_mass_matrix
        self._dense_mass = dense_mass
        self._target_accept_prob = target_accept_prob
        self._trajectory_length = (float(trajectory_length) if isinstance(trajectory_length, int)
                                   else trajectory_length)
        self._algo = "HMC"
        self._max_tree_depth = 10
        self._init_strategy = init_strategy
        self._find_heuristic_step_size = find_heuristic_step_size
        self._forward_mode_differentiation = forward_mode_differentiation
        self._regularize_mass_matrix = regularize_mass_matrix
        self.partitionable = None  # None -> numpyro_b200.random.DEFAULT_PARTITIONABLE
        # set on first call to init
        self._init_fn = None
        self._potential_fn_gen = None
        self._postprocess_fn = None
        self._sample_fn = None
        self._handle = None
        self._bound = None
        self._data = None
        self._vectorized = False
        self._last_state = None
        self._row_shard = None  # (rank, nranks, allgather) -- see shard_rows
        self._bound_sig = None  # identity of the model args the device handle was built from
        self._ecs = None  # set by HMCECS: {"degree", "ref", "after_create"} (hmc_gibbs.py)

    # ------------------------------------------------------------------ properties
    @property
    def model(self):
        return self._model

    @property
    def sample_field(self):
        return "z"

    @property
    def default_fields(self):
        return ("z", "diverging")

    def get_diagnostics_str(self, state):
        return "{} steps of size {:.2e}. acc. prob={:.2f}".format(
            state.num_steps, state.adapt_state.step_size, state.mean_accept_prob)

    def postprocess_fn(self, args, kwargs):
        if self._postprocess_fn is None:
            return lambda x: x
        return self._postprocess_fn

    def shard_rows(self, rank, nranks, allgather):
        """Row-shard a tall-data (logistic) model over ``nranks`` GPUs, one process each: this
        process passes only ITS rows to ``init``/``MCMC.run``; the sampler state is replicated
        and the D+1 partial sums are exchanged per leapfrog over NVLink peer memory (the
        reference's analogue is GSPMD row sharding, ``mcmc.py:239-265``).
        ``allgather(bytes) -> list[bytes]`` (one entry per rank, rank order) is the host
        transport used once to exchange CUDA IPC handles, e.g.
        ``numpyro_b200.distributed.allgather_bytes``."""
        self._row_shard = None if nranks <= 1 else (int(rank), int(nranks), allgather)
        return self

    # ------------------------------------------------------------------ handle management
    def _destroy(self):
        """Release the handle AND this kernel's device copies of the model data now (they go
        back to the library's block cache), instead of whenever the object is collected: the
        bound ``_sample_fn`` makes a reference cycle, so without this the 128 MB design-matrix
        copy of a finished run stays allocated until a cyclic GC pass, and the next run pays a
        fresh ``cudaMalloc`` (seen as 15-20 ms spikes of the H2D phase end to end)."""
        if self._handle is not None:
            _lib.lib().b200mc_destroy(self._handle)
            self._handle = None
        if self._data is not None:
            for d in self._data:
                if d is not None:
                    d.free()
            self._data = None
        self._sample_fn = None

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    def __getstate__(self):  # hmc.py:818-822
        state = self.__dict__.copy()
        # everything that only exists after init() -- device handle and buffers, the model bound
        # to its data (holds closures), transports -- is dropped; init() rebuilds it
        for k in ("_sample_fn", "_init_fn", "_handle", "_data", "_last_state", "_postprocess_fn",
                  "_row_shard", "_bound", "_potential_fn_gen", "_bound_sig", "_ecs"):
            state[k] = None
        state["_model"] = self._model.__name__  # registry entries hold closures: pickle by name
        return state

    def __setstate__(self, state):
        self.__dict__.update(state)
        if isinstance(self._model, str):
            self._model = models.lookup(self._model)

    def _part(self):
        return int(b2random.DEFAULT_PARTITIONABLE if self.partitionable is None
                   else self.partitionable)

    def _mass_layout(self, bound):
        """Block structure of the mass matrix, ``hmc.py:760-769`` + ``hmc_util.py:439-486``.

        Returns ``(blocks, key_order, user)``: ``blocks`` = ``[(site_names, flat_idx, dense)]``
        in the insertion order of the reference's ``inverse_mass_matrix`` dict (the user's
        entries, then the dense blocks, then ONE diagonal block of the remaining sites in
        sorted order); ``key_order`` = block indices in the order of its ``mass_matrix_sqrt``
        dict, which is the order of the per-block momentum keys (``hmc.py:94-96``): dense
        blocks, the user's other blocks, the remaining sites; ``user`` = the user's matrices
        per block (or None)."""
        sites = sorted(k for k, _ in bound.trace_sites)
        dm = self._dense_mass
        dm = ([tuple(sites)] if dm else []) if isinstance(dm, bool) else list(dm)
        m = self._inverse_mass_matrix
        if m is None:
            user = {}
        elif not isinstance(m, dict):
            user = {tuple(sites): m}
        else:
            user = {tuple(k): v for k, v in m.items()}
        order = list(user)
        order += [b for b in dm if b not in user]
        used = [k for b in order for k in b]
        for k in used:
            if k not in sites:
                raise KeyError(k)
        remaining = tuple(sorted(set(sites) - set(used)))
        if remaining:
            order.append(remaining)
        if sorted(k for b in order for k in b) != sites:
            raise AssertionError(
                "There seems to be a conflict of sites names specified in the initial"
                " `inverse_mass_matrix` and in `dense_mass` argument.")
        blocks = []
        for names in order:
            idx = np.concatenate([np.arange(bound.flat_sites[k][0],
                                            bound.flat_sites[k][0] + bound.flat_sites[k][1])
                                  for k in names]).astype(np.int32)
            blocks.append((names, idx, names in dm))
        keys = [b for b in dm] + [b for b in user if b not in dm] + ([remaining] if remaining else [])
        key_order = [order.index(b) for b in keys]
        return blocks, key_order, user

    def _create(self, bound, num_chains, num_warmup):
        if bound.family == models.LOGISTIC_ECS and self._ecs is None:  # primitives.py:444-449
            raise ValueError("Missing random key to generate subsample indices."
                             " Algorithms like HMC/NUTS do not support subsampling."
                             " You might want to use SVI or HMCECS instead.")
        _lib.require_device()
        self._destroy()
        lib = _lib.lib()
        d0 = _lib.DeviceArray.from_numpy(bound.data0)
        d1 = _lib.DeviceArray.from_numpy(bound.data1) if bound.data1 is not None else None
        blocks, key_order, user = self._mass_layout(bound)
        D = bound.dim
        any_dense = any(dn for _, _, dn in blocks)
        if len(blocks) > 1 and key_order != list(range(len(blocks))):
            if self._adapt_mass_matrix:
                raise NotImplementedError(
                    "an `inverse_mass_matrix` dict that lists a diagonal block before a dense one "
                    "changes the reference's momentum-key order at the first window end "
                    "(hmc_util.py:205-215 vs :449-474); list the dense blocks first")
            blocks = [blocks[i] for i in key_order]
        self._blocks = blocks
        self._dense = any_dense
        imm = None
        if user:
            full = np.eye(D, dtype=np.float32) if any_dense else np.ones(D, np.float32)
            for names, idx, dn in blocks:
                if names not in user:
                    continue
                m = np.asarray(user[names], np.float32)
                d = idx.size
                if dn:
                    if m.shape == (d,):
                        m = np.diag(m).astype(np.float32)  # hmc_util.py:492-493
                    if m.shape != (d, d):
                        raise ValueError(f"inverse_mass_matrix of block {names} must have shape {(d, d)}")
                    full[np.ix_(idx, idx)] = m
                else:
                    if m.shape == (d, d):
                        m = np.diag(m).copy()  # hmc_util.py:509-510
                    if m.shape != (d,):
                        raise ValueError(f"inverse_mass_matrix of block {names} must have shape {(d,)}")
                    if any_dense:
                        full[idx, idx] = m
                    else:
                        full[idx] = m
            imm = _lib.DeviceArray.from_numpy(full)
        tables = None
        if len(blocks) > 1:
            tables = (np.ascontiguousarray(np.concatenate([idx for _, idx, _ in blocks]), np.int32),
                      np.array([idx.size for _, idx, _ in blocks], np.int32),
                      np.array([int(dn) for _, _, dn in blocks], np.int32))
        depth = self._max_tree_depth
        depth = depth if isinstance(depth, tuple) else (depth, depth)
        kind, arg = self._strategy()
        ecs_ref = None
        if bound.family == models.LOGISTIC_ECS and self._ecs["degree"]:
            ecs_ref = _lib.DeviceArray.from_numpy(bound.ravel(self._ecs["ref"]).reshape(-1), np.float32)
        cond = getattr(bound, "cond_index", None)
        cfg = _lib.Config(
            cond_count=0 if cond is None else int(cond.size),
            cond_index_host=None if cond is None else cond.ctypes.data,
            struct_size=C.sizeof(_lib.Config), family=bound.family, num_chains=num_chains, N=bound.N, ncols=bound.ncols,
            data0_dev=d0.ptr, data1_dev=d1.ptr if d1 is not None else None,
            algo=0 if self._algo == "NUTS" else 1, step_size=float(self._step_size),
            adapt_step_size=int(bool(self._adapt_step_size)),
            adapt_mass_matrix=int(bool(self._adapt_mass_matrix)),
            dense_mass=int(any_dense),
            find_heuristic_step_size=int(bool(self._find_heuristic_step_size)),
            likelihood=int(getattr(bound, "likelihood", 0)),
            nranks=self._row_shard[1] if self._row_shard else 1,
            rank=self._row_shard[0] if self._row_shard else 0,
            regularize_mass_matrix=int(bool(self._regularize_mass_matrix)),
            target_accept_prob=float(self._target_accept_prob),
            max_tree_depth_warmup=int(depth[0]), max_tree_depth_sample=int(depth[1]),
            num_warmup=int(num_warmup),
            has_trajectory_length=int(self._trajectory_length is not None),
            trajectory_length=float(self._trajectory_length or 0.0),
            num_steps=int(self._num_steps or 0), threefry_partitionable=self._part(),
            init_radius=float(arg) if kind == "uniform" else 2.0,
            inverse_mass_matrix_dev=imm.ptr if imm is not None else None,
            num_blocks=len(blocks),
            block_perm_host=tables[0].ctypes.data if tables else None,
            block_sizes_host=tables[1].ctypes.data if tables else None,
            block_dense_host=tables[2].ctypes.data if tables else None,
            subsample_size=int(getattr(bound, "subsample_size", 0)),
            ecs_degree=int(self._ecs["degree"]) if self._ecs else 0,
            ecs_ref_dev=ecs_ref.ptr if ecs_ref is not None else None,
            centered=float(getattr(bound, "centered", 0.0)))
        h = C.c_void_p()
        if self._row_shard and bound.family != models.LOGISTIC:
            raise NotImplementedError("row sharding exists for the tall-data logistic family only")
        _lib.check(lib.b200mc_create(C.byref(cfg), C.byref(h)))
        self._handle = h
        if self._row_shard:
            mine = C.create_string_buffer(128)
            _lib.check(lib.b200mc_comm_export(h, mine))
            table = self._row_shard[2](mine.raw)
            if len(table) != self._row_shard[1] or any(len(t) != 128 for t in table):
                raise _lib.B200Error("row sharding: allgather must return one 128-byte entry per rank")
            _lib.check(lib.b200mc_comm_import(h, C.create_string_buffer(b"".join(table), 128 * len(table))))
        self._data = (d0, d1, imm, ecs_ref)
        self._bound = bound
        self._num_chains = num_chains
        self._num_warmup = num_warmup
        self._postprocess_fn = bound.postprocess
        self._potential_fn_gen = bound

    def _strategy(self):
        s = self._init_strategy
        if callable(s):
            s = s()
        if not (isinstance(s, tuple) and s[0] in ("uniform", "value")):
            raise NotImplementedError("init_strategy must be init_to_uniform or init_to_value")
        return s

    # ------------------------------------------------------------------ state <-> device
    def _snapshot(self):
        lib, h = _lib.lib(), self._handle
        Cn, D = self._num_chains, self._bound.dim
        sc = np.zeros(Cn, _lib.SCALARS_DTYPE)
        _lib.check(lib.b200mc_get_scalars(h, sc.ctypes.data))
        tmp = _lib.DeviceArray((Cn, D), np.float32)
        vec = {}
        for name, which in _VEC.items():
            _lib.check(lib.b200mc_get_vector(h, which, tmp.ptr, None))
            vec[name] = tmp.numpy()
        if self._dense:
            mt = _lib.DeviceArray((Cn, D, D), np.float32)
            for name, which in _MAT.items():
                _lib.check(lib.b200mc_get_vector(h, which, mt.ptr, None))
                vec[name] = mt.numpy()
        return sc, vec

    def _to_state(self, sc, vec):
        b = self._bound
        sq = (lambda a: a) if self._vectorized else (lambda a: a[0])
        imm = vec["inverse_mass_matrix"]
        sqrt_inv = vec["mass_matrix_sqrt_inv"] if self._dense else np.sqrt(imm)

        def cut(a, idx, dn):
            """block of a per-chain [C, D] vector / [C, D, D] matrix (device layout: flat order)"""
            if a.ndim == 3:
                return a[:, idx[:, None], idx[None, :]] if dn else a[:, idx, idx]
            return a[:, idx]

        per_block = lambda a: {names: sq(cut(a, idx, dn)) for names, idx, dn in self._blocks}  # noqa: E731
        adapt = HMCAdaptState(
            step_size=sq(sc["step_size"].copy()),
            inverse_mass_matrix=per_block(imm),
            mass_matrix_sqrt=per_block(vec["mass_matrix_sqrt"]),
            mass_matrix_sqrt_inv=per_block(sqrt_inv),
            ss_state=(sq(sc["ss_x_t"].copy()), sq(sc["ss_x_avg"].copy()), sq(sc["ss_g_avg"].copy()),
                      sq(sc["ss_t"].copy()), sq(sc["ss_prox_center"].copy())),
            mm_state={names: (sq(vec["welford_mean"][:, idx]), sq(cut(vec["welford_m2"], idx, dn)),
                              sq(sc["mm_n"].copy())) for names, idx, dn in self._blocks},
            window_idx=sq(sc["window_idx"].copy()),
            rng_key=sq(sc["adapt_key"].copy()))
        tl = None
        if self._algo == "HMC" and self._trajectory_length is not None:
            tl = np.float32(self._trajectory_length)
        state = HMCState(
            i=sq(sc["i"].copy()), z=b.unravel(sq(vec["z"])), z_grad=b.unravel(sq(vec["z_grad"])),
            potential_energy=sq(sc["potential_energy"].copy()), energy=sq(sc["energy"].copy()),
            r=None, trajectory_length=tl, num_steps=sq(sc["num_steps"].copy()),
            accept_prob=sq(sc["accept_prob"].copy()),
            mean_accept_prob=sq(sc["mean_accept_prob"].copy()),
            diverging=sq(sc["diverging"].astype(bool)), adapt_state=adapt,
            rng_key=sq(sc["rng_key"].copy()))
        self._last_state = state
        self._last_grad_evals = sc["grad_evals"].copy()
        return state

    def _push_state(self, state):
        """Make the device hold ``state`` (no-op when it is the state we last returned)."""
        if state is self._last_state:
            return
        lib, h, b = _lib.lib(), self._handle, self._bound
        Cn = self._num_chains
        up = (lambda a: np.asarray(a)) if self._vectorized else (lambda a: np.asarray(a)[None])
        sc = np.zeros(Cn, _lib.SCALARS_DTYPE)
        a = state.adapt_state
        block = next(iter(a.inverse_mass_matrix))
        D = b.dim

        def join(d, mat):
            """per-block dict -> [C, D] vector or (``mat``) [C, D, D] matrix in flat order"""
            out = np.zeros((Cn, D, D) if mat else (Cn, D), np.float32)
            for names, idx, dn in self._blocks:
                v = up(d[names]).astype(np.float32)
                if mat and dn:
                    out[:, idx[:, None], idx[None, :]] = v.reshape(Cn, idx.size, idx.size)
                elif mat:
                    out[:, idx, idx] = v.reshape(Cn, idx.size)
                else:
                    out[:, idx] = v.reshape(Cn, idx.size)
            return out
        for name, val in (("i", state.i), ("potential_energy", state.potential_energy),
                          ("energy", state.energy), ("num_steps", state.num_steps),
                          ("accept_prob", state.accept_prob),
                          ("mean_accept_prob", state.mean_accept_prob),
                          ("diverging", np.asarray(state.diverging).astype(np.int32)),
                          ("rng_key", state.rng_key), ("step_size", a.step_size),
                          ("ss_x_t", a.ss_state[0]), ("ss_x_avg", a.ss_state[1]),
                          ("ss_g_avg", a.ss_state[2]), ("ss_t", a.ss_state[3]),
                          ("ss_prox_center", a.ss_state[4]), ("mm_n", a.mm_state[block][2]),
                          ("window_idx", a.window_idx), ("adapt_key", a.rng_key)):
            sc[name] = up(val)
        if getattr(self, "_last_grad_evals", None) is not None:
            sc["grad_evals"] = self._last_grad_evals
        _lib.check(lib.b200mc_set_scalars(h, sc.ctypes.data))
        dn_ = self._dense
        vecs = {"z": b.ravel(state.z), "z_grad": b.ravel(state.z_grad),
                "welford_mean": join({k: v[0] for k, v in a.mm_state.items()}, False)}
        mats = {}
        (mats if dn_ else vecs).update({
            "inverse_mass_matrix": join(a.inverse_mass_matrix, dn_),
            "mass_matrix_sqrt": join(a.mass_matrix_sqrt, dn_),
            "welford_m2": join({k: v[1] for k, v in a.mm_state.items()}, dn_)})
        if dn_:
            mats["mass_matrix_sqrt_inv"] = join(a.mass_matrix_sqrt_inv, True)
        for name, val in vecs.items():
            d = _lib.DeviceArray.from_numpy(up(val).reshape(Cn, b.dim), np.float32)
            _lib.check(lib.b200mc_set_vector(h, _VEC[name], d.ptr, None))
            _lib.check(lib.b200mc_stream_synchronize(None))
        for name, val in mats.items():
            d = _lib.DeviceArray.from_numpy(up(val).reshape(Cn, b.dim, b.dim), np.float32)
            _lib.check(lib.b200mc_set_vector(h, _MAT[name], d.ptr, None))
            _lib.check(lib.b200mc_stream_synchronize(None))
        self._last_state = state

    # ------------------------------------------------------------------ MCMCKernel interface
    def init(self, rng_key, num_warmup, init_params=None, model_args=(), model_kwargs={}):
        """``hmc.py:740-799``: ``rng_key`` is ``uint32[2]`` (one chain) or ``uint32[C, 2]``
        (vectorized chains)."""
        rng_key = np.asarray(rng_key, np.uint32)
        self._vectorized = not b2random.is_prng_key(rng_key)
        keys = rng_key.reshape(-1, 2)
        bound = self._model.bind(*model_args, **model_kwargs)
        self._create(bound, keys.shape[0], num_warmup)
        self._bound_sig = _args_signature(model_args, model_kwargs)
        kind, arg = self._strategy()
        if init_params is None and kind == "value":
            # initialization.py `init_to_value`: CONSTRAINED values per site.  The device init
            # search draws either every site or none; a partial dict is completed on the host
            # (device RNG, device potential) and handed over as the initial point.
            missing = [k for k, _ in bound.trace_sites if k not in arg]
            if missing:
                # the sites that are not listed fall back to init_to_uniform (initialization.py)
                k_init = b2random.split(keys, 2, self._part())[:, 1]  # hmc.py:745
                z = self._partial_init_values(bound, k_init, arg)
                init_params = bound.unravel(z if self._vectorized else z[0])
            else:
                init_params = bound.unconstrain(arg)
                if self._vectorized:
                    init_params = {k: np.broadcast_to(v, (keys.shape[0],) + np.shape(v)).copy()
                                   for k, v in init_params.items()}
        z0 = None
        if init_params is not None:
            z = bound.ravel(init_params)
            z = z.reshape(keys.shape[0], bound.dim) if self._vectorized else z.reshape(1, bound.dim)
            z0 = _lib.DeviceArray.from_numpy(z, np.float32)
        kd = _lib.DeviceArray.from_numpy(keys)
        if self._ecs is not None and self._ecs.get("after_create") is not None:
            self._ecs["after_create"](self)  # HMCECS: the initial subsample, before the init search
        _lib.check(_lib.lib().b200mc_init(self._handle, kd.ptr, z0.ptr if z0 is not None else None,
                                          None))
        sc, vec = self._snapshot()
        if (sc["status"] != 0).any():
            raise RuntimeError("Cannot find valid initial parameters. Please check your model again.")
        self._sample_fn = self._sample_one
        self._init_fn = True
        return self._to_state(sc, vec)

    def _partial_init_values(self, bound, init_keys, values, radius=2.0):
        """``find_valid_initial_params`` (``infer/util.py:386-424``, the branch that traces the seeded
        model) under ``init_to_value(values=...)`` with only SOME sites given: per attempt
        ``key, subkey = split(key)``; the seed handler then splits ``subkey`` once per latent site
        in trace order; a listed site takes its (constrained) value, any other one
        ``Uniform(-radius, radius)`` in unconstrained space pushed through the site's transform
        (``initialization.py`` ``init_to_uniform``); the constrained values go back through the
        inverse transforms; the point must have a finite potential and gradient (100 attempts).
        ``init_keys``: ``uint32[C, 2]`` (``rng_key_init_model`` of every chain).  Returns z[C, D]."""
        part = self._part()
        unit = getattr(bound, "unit_interval", ())
        for k, _ in bound.trace_sites:
            if k not in values and k not in bound.positive and k not in unit and \
                    bound.family == models.DIRICHLET_CATEGORICAL:
                raise NotImplementedError("init_to_value with a missing simplex site is not available")
        out = np.zeros((init_keys.shape[0], bound.dim), np.float32)
        fi = np.finfo(np.float32)
        for c in range(init_keys.shape[0]):
            key = np.asarray(init_keys[c], np.uint32)
            for _ in range(100):
                key, subkey = b2random.split(key, 2, part)
                k, zc = subkey, {}
                for name, shape in bound.trace_sites:
                    k, ks = b2random.split(k, 2, part)
                    if name in values:
                        v = np.asarray(values[name], np.float32)
                        if name in bound.positive:
                            v = np.log(v)
                        elif name in unit:
                            v = np.log(v) - np.log1p(-v)
                    else:
                        v = b2random.uniform(ks, shape, -radius, radius, part)
                        with np.errstate(all="ignore"):
                            if name in bound.positive:  # ExpTransform and back
                                v = np.log(np.exp(v))
                            elif name in unit:          # SigmoidTransform (clipped expit) and back
                                p = np.clip(1 / (1 + np.exp(-v)), fi.tiny, np.float32(1) - fi.eps).astype(np.float32)
                                v = np.log(p) - np.log1p(-p)
                    zc[name] = np.broadcast_to(np.asarray(v, np.float32), shape)
                z = bound.ravel(zc).astype(np.float32).reshape(1, bound.dim)
                U, g = self.potential_and_grad(z)
                if np.isfinite(U).all() and np.isfinite(g).all():
                    out[c] = z[0]
                    break
            else:
                raise RuntimeError("Cannot find valid initial parameters. Please check your model again.")
        return out

    def _check_args(self, model_args, model_kwargs):
        """The reference rebuilds the potential from the args on every call
        (``hmc.py:806-816`` ``potential_fn_gen(*model_args)``); here the data lives in the
        device handle, so new args after ``init`` are re-bound (new handle, same chain state)."""
        if self._bound_sig is None or (not model_args and not model_kwargs):
            return
        sig = _args_signature(model_args, model_kwargs)
        if sig == self._bound_sig:
            return
        state = self._last_state
        bound = self._model.bind(*model_args, **model_kwargs)
        if bound.dim != self._bound.dim or bound.trace_sites != self._bound.trace_sites:
            raise ValueError("model arguments changed the latent sites; call `init` again")
        self._create(bound, self._num_chains, self._num_warmup)
        self._bound_sig = sig
        self._sample_fn = self._sample_one
        self._last_state = None
        if state is not None:
            self._push_state(state)

    def _sample_one(self, state, model_args, model_kwargs):
        self._check_args(model_args, model_kwargs)
        self._push_state(state)
        _lib.check(_lib.lib().b200mc_sample(self._handle, 1, 1, 1, None, None))
        return self._to_state(*self._snapshot())

    def sample(self, state, model_args, model_kwargs):
        """``hmc.py:806-816``: one transition."""
        if self._handle is None:
            raise RuntimeError("call `init` before `sample`")
        return self._sample_one(state, model_args, model_kwargs)

    # ------------------------------------------------------------------ fused loop (fori_collect)
    def run_loop(self, state, num_iters, collect_start, thinning, n_keep, fields,
                 model_args=(), model_kwargs=None):
        """Run ``num_iters`` transitions in ONE launch, collecting ``fields`` (a subset of
        ``("z",) + EXTRA_FIELDS``) exactly as ``fori_collect`` would
        (``numpyro/util.py:277-407``).  Returns ``(dict of [n_keep, C, ...] arrays, last state)``."""
        self._check_args(model_args, model_kwargs or {})
        self._push_state(state)
        bufs, col = self._collect_buffers(n_keep, fields)
        _lib.check(_lib.lib().b200mc_sample(self._handle, int(num_iters), int(collect_start),
                                            int(thinning), C.byref(col), None))
        new_state = self._to_state(*self._snapshot())
        out = {f: b.numpy() for f, b in bufs.items()}
        if "diverging" in out:
            out["diverging"] = out["diverging"].astype(bool)
        return out, new_state

    def _collect_buffers(self, n_keep, fields):
        """Device buffers ``[n_keep, C, ...]`` for the collected ``fields`` and the C struct that
        points at them."""
        Cn, D = self._num_chains, self._bound.dim
        bufs, col = {}, _lib.Collect(n_keep=n_keep)
        spec = {"z": ("z_dev", (n_keep, Cn, D), np.float32),
                "potential_energy": ("potential_energy_dev", (n_keep, Cn), np.float32),
                "energy": ("energy_dev", (n_keep, Cn), np.float32),
                "num_steps": ("num_steps_dev", (n_keep, Cn), np.int32),
                "accept_prob": ("accept_prob_dev", (n_keep, Cn), np.float32),
                "mean_accept_prob": ("mean_accept_prob_dev", (n_keep, Cn), np.float32),
                "diverging": ("diverging_dev", (n_keep, Cn), np.int32),
                "adapt_state.step_size": ("step_size_dev", (n_keep, Cn), np.float32)}
        if n_keep > 0:
            for f in fields:
                attr, shape, dt = spec[f]
                bufs[f] = _lib.DeviceArray(shape, dt).zero_()
                setattr(col, attr, bufs[f].ptr)
        return bufs, col

    def grad_evals(self):
        """Cumulative potential+gradient evaluations per chain (init search included)."""
        return None if getattr(self, "_last_grad_evals", None) is None else self._last_grad_evals.copy()

    def potential_and_grad(self, z):
        """Tier-2 hook: ``(U, grad)`` at flat unconstrained points ``z[n, D]``."""
        z = np.ascontiguousarray(z, np.float32).reshape(-1, self._bound.dim)
        zd = _lib.DeviceArray.from_numpy(z)
        U = _lib.DeviceArray((z.shape[0],), np.float32)
        g = _lib.DeviceArray(z.shape, np.float32)
        _lib.check(_lib.lib().b200mc_potential_and_grad(self._handle, zd.ptr, z.shape[0], U.ptr,
                                                        g.ptr, None))
        return U.numpy(), g.numpy()


    def potential_and_grad_passes(self, z, passes):
        """Tier-2 hook of the streaming kernel: the same point on ``passes`` consecutive passes
        of one launch (forward / reversed sweeps, ring wrap).  Returns ``(U[passes],
        grad[passes, D])``."""
        z = np.ascontiguousarray(z, np.float32).reshape(self._bound.dim)
        zd = _lib.DeviceArray.from_numpy(z)
        U = _lib.DeviceArray((passes,), np.float32)
        g = _lib.DeviceArray((passes, z.shape[0]), np.float32)
        _lib.check(_lib.lib().b200mc_potential_and_grad_passes(self._handle, zd.ptr, int(passes),
                                                               U.ptr, g.ptr, None))
        return U.numpy(), g.numpy()


class NUTS(HMC):
    """No-U-Turn Sampler (``hmc.py:825-948``)."""

    def __init__(self, model=None, potential_fn=None, kinetic_fn=None, step_size=1.0,
                 inverse_mass_matrix=None, adapt_step_size=True, adapt_mass_matrix=True,
                 dense_mass=False, target_accept_prob=0.8, trajectory_length=None,
                 max_tree_depth=10, init_strategy=init_to_uniform, find_heuristic_step_size=False,
                 forward_mode_differentiation=False, regularize_mass_matrix=True):
        super().__init__(potential_fn=potential_fn, model=model, kinetic_fn=kinetic_fn,
                         step_size=step_size, inverse_mass_matrix=inverse_mass_matrix,
                         adapt_step_size=adapt_step_size, adapt_mass_matrix=adapt_mass_matrix,
                         dense_mass=dense_mass, target_accept_prob=target_accept_prob,
                         trajectory_length=trajectory_length, init_strategy=init_strategy,
                         find_heuristic_step_size=find_heuristic_step_size,
                         forward_mode_differentiation=forward_mode_differentiation,
                         regularize_mass_matrix=regularize_mass_matrix)
        self._max_tree_depth = max_tree_depth
        self._algo = "NUTS"
