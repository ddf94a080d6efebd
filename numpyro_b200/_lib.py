"""ctypes binding of ``libb200mc.so`` (C ABI declared in ``include/b200mc.h``).

There is no CPU fallback: importing this module without the built CUDA library, or
calling into it without a CUDA device, raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200MC_LIB") or os.path.join(_HERE, "libb200mc.so")


ABI_VERSION = 4  # include/b200mc.h B200MC_ABI_VERSION


class B200Error(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("family", C.c_int), ("num_chains", C.c_int), ("N", C.c_int), ("ncols", C.c_int),
        ("data0_dev", C.c_void_p), ("data1_dev", C.c_void_p),
        ("algo", C.c_int), ("step_size", C.c_float),
        ("adapt_step_size", C.c_int), ("adapt_mass_matrix", C.c_int), ("dense_mass", C.c_int),
        ("regularize_mass_matrix", C.c_int), ("target_accept_prob", C.c_float),
        ("max_tree_depth_warmup", C.c_int), ("max_tree_depth_sample", C.c_int),
        ("num_warmup", C.c_int), ("has_trajectory_length", C.c_int),
        ("trajectory_length", C.c_float), ("num_steps", C.c_int),
        ("threefry_partitionable", C.c_int), ("init_radius", C.c_float),
        ("inverse_mass_matrix_dev", C.c_void_p), ("find_heuristic_step_size", C.c_int),
        ("likelihood", C.c_int), ("nranks", C.c_int), ("rank", C.c_int),
        ("num_blocks", C.c_int), ("block_perm_host", C.c_void_p),
        ("block_sizes_host", C.c_void_p), ("block_dense_host", C.c_void_p),
        ("subsample_size", C.c_int), ("ecs_degree", C.c_int), ("ecs_ref_dev", C.c_void_p),
        ("centered", C.c_float), ("cond_count", C.c_int), ("cond_index_host", C.c_void_p),
    ]


class ChainScalars(C.Structure):
    _fields_ = [
        ("i", C.c_int32), ("potential_energy", C.c_float), ("energy", C.c_float),
        ("num_steps", C.c_int32), ("accept_prob", C.c_float), ("mean_accept_prob", C.c_float),
        ("diverging", C.c_int32), ("rng_key", C.c_uint32 * 2), ("step_size", C.c_float),
        ("ss_x_t", C.c_float), ("ss_x_avg", C.c_float), ("ss_g_avg", C.c_float),
        ("ss_t", C.c_int32), ("ss_prox_center", C.c_float), ("mm_n", C.c_int32),
        ("window_idx", C.c_int32), ("adapt_key", C.c_uint32 * 2), ("status", C.c_int32),
        ("init_tries", C.c_int32), ("grad_evals", C.c_int64),
    ]


SCALARS_DTYPE = np.dtype([
    ("i", "<i4"), ("potential_energy", "<f4"), ("energy", "<f4"), ("num_steps", "<i4"),
    ("accept_prob", "<f4"), ("mean_accept_prob", "<f4"), ("diverging", "<i4"),
    ("rng_key", "<u4", (2,)), ("step_size", "<f4"), ("ss_x_t", "<f4"), ("ss_x_avg", "<f4"),
    ("ss_g_avg", "<f4"), ("ss_t", "<i4"), ("ss_prox_center", "<f4"), ("mm_n", "<i4"),
    ("window_idx", "<i4"), ("adapt_key", "<u4", (2,)), ("status", "<i4"), ("init_tries", "<i4"),
    ("grad_evals", "<i8"),
], align=True)
assert SCALARS_DTYPE.itemsize == C.sizeof(ChainScalars)


class Collect(C.Structure):
    _fields_ = [
        ("n_keep", C.c_int), ("z_dev", C.c_void_p), ("potential_energy_dev", C.c_void_p),
        ("energy_dev", C.c_void_p), ("num_steps_dev", C.c_void_p),
        ("accept_prob_dev", C.c_void_p), ("mean_accept_prob_dev", C.c_void_p),
        ("diverging_dev", C.c_void_p), ("step_size_dev", C.c_void_p),
    ]


# every symbol include/b200mc.h declares: (restype, argtypes)
_vp, _i, _f, _u64 = C.c_void_p, C.c_int, C.c_float, C.c_uint64
SYMBOLS = {
    "b200mc_last_error": (C.c_char_p, []),
    "b200mc_version": (_i, []),
    "b200mc_create": (_i, [C.POINTER(Config), C.POINTER(_vp)]),
    "b200mc_destroy": (None, [_vp]),
    "b200mc_dim": (_i, [_vp]),
    "b200mc_init": (_i, [_vp, _vp, _vp, _vp]),
    "b200mc_sample": (_i, [_vp, _i, _i, _i, C.POINTER(Collect), _vp]),
    "b200mc_get_scalars": (_i, [_vp, _vp]),
    "b200mc_set_scalars": (_i, [_vp, _vp]),
    "b200mc_get_vector": (_i, [_vp, _i, _vp, _vp]),
    "b200mc_set_vector": (_i, [_vp, _i, _vp, _vp]),
    "b200mc_get_profile": (_i, [_vp, _vp, _i]),
    "b200mc_comm_export": (_i, [_vp, _vp]),
    "b200mc_comm_import": (_i, [_vp, _vp]),
    "b200mc_obs_forward": (_i, [_vp, _i, _i, _vp, _i, _vp, _i, _vp, _vp, _vp, _i, _i, _vp, _vp]),
    "b200mc_diag_ess_rhat": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "b200mc_glm_layout": (_i, [_vp, C.POINTER(_i)]),
    "b200mc_membench": (_i, [_u64, _i, _i, _i, _i, C.POINTER(_f), _vp]),
    "b200mc_potential_and_grad": (_i, [_vp, _vp, _i, _vp, _vp, _vp]),
    "b200mc_potential_and_grad_passes": (_i, [_vp, _vp, _i, _vp, _vp, _vp]),
    "b200mc_horseshoe_potential_noalloc": (_i, [_vp, _vp, _i, _i, _i, _vp, _i, _vp, _vp, _vp]),
    "b200mc_threefry2x32": (_i, [_vp, _vp, _vp, _i, _vp]),
    "b200mc_random_randint": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
    "b200mc_ecs_init_subsample": (_i, [_vp, _vp, _vp]),
    "b200mc_ecs_get_subsample": (_i, [_vp, _vp, _vp]),
    "b200mc_ecs_set_subsample": (_i, [_vp, _vp, _vp]),
    "b200mc_ecs_update": (_i, [_vp, _vp, _i, _vp, _vp]),
    "b200mc_ecs_sample": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "b200mc_set_conditioned": (_i, [_vp, _vp, _vp, _vp]),
    "b200mc_random_split": (_i, [_vp, _i, _i, _vp, _i, _vp]),
    "b200mc_random_bits": (_i, [_vp, _i, _vp, _i, _vp]),
    "b200mc_random_uniform": (_i, [_vp, _i, _f, _f, _vp, _i, _vp]),
    "b200mc_random_normal": (_i, [_vp, _i, _vp, _i, _vp]),
    "b200mc_gemm3xtf32": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "b200mc_adaptation_schedule": (_i, [_i, C.POINTER(_i), C.POINTER(_i), _i]),
    "b200mc_device_count": (_i, []),
    "b200mc_set_device": (_i, [_i]),
    "b200mc_malloc": (_vp, [_u64]),
    "b200mc_free": (_i, [_vp]),
    "b200mc_pool_trim": (_i, []),
    "b200mc_memcpy_h2d": (_i, [_vp, _vp, _u64, _vp]),
    "b200mc_memcpy_d2h": (_i, [_vp, _vp, _u64, _vp]),
    "b200mc_memset": (_i, [_vp, _i, _u64, _vp]),
    "b200mc_stream_synchronize": (_i, [_vp]),
    "b200mc_host_alloc": (_vp, [_u64]),
    "b200mc_host_free": (_i, [_vp]),
    "b200mc_event_create": (_vp, []),
    "b200mc_event_record": (_i, [_vp, _vp]),
    "b200mc_event_elapsed_ms": (_i, [_vp, _vp, C.POINTER(_f)]),
    "b200mc_event_destroy": (_i, [_vp]),
    "b200mc_kernel_launches": (C.c_int64, []),
}

_lib = None


def lib():
    """The loaded library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200Error(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (nvcc, sm_100a). numpyro_b200 has no CPU fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            if "B200MC_LIB" in os.environ and not hasattr(handle, name):
                continue  # development A/B runs against an older build
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        got = handle.b200mc_version()
        if got != ABI_VERSION:
            raise B200Error(f"{LIB_PATH} speaks ABI version {got}, this binding was written for "
                            f"{ABI_VERSION} (include/b200mc.h B200MC_ABI_VERSION): rebuild the library")
        _lib = handle
    return _lib


def check(rc):
    if rc != 0:
        raise B200Error(lib().b200mc_last_error().decode())


def require_device():
    if lib().b200mc_device_count() < 1:
        raise B200Error("no CUDA device visible: numpyro_b200 runs on the GPU only (no CPU fallback)")


class DeviceArray:
    """A device buffer with a numpy-side shape/dtype; owns its memory."""

    def __init__(self, shape, dtype):
        self.shape = tuple(int(s) for s in np.atleast_1d(shape)) if np.ndim(shape) else (int(shape),)
        self.dtype = np.dtype(dtype)
        self.nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        self.ptr = lib().b200mc_malloc(max(self.nbytes, 16))
        if not self.ptr:
            raise B200Error(lib().b200mc_last_error().decode())

    @classmethod
    def from_numpy(cls, a, dtype=None, stream=None):
        a = np.ascontiguousarray(a, dtype=dtype)
        out = cls(a.shape if a.ndim else (1,), a.dtype)
        if a.nbytes:
            check(lib().b200mc_memcpy_h2d(out.ptr, a.ctypes.data, a.nbytes, stream))
            check(lib().b200mc_stream_synchronize(stream))
        return out

    def numpy(self, stream=None):
        out = np.empty(self.shape, self.dtype)
        if self.nbytes:
            check(lib().b200mc_memcpy_d2h(out.ctypes.data, self.ptr, self.nbytes, stream))
            check(lib().b200mc_stream_synchronize(stream))
        return out

    def zero_(self, stream=None):
        check(lib().b200mc_memset(self.ptr, 0, self.nbytes, stream))
        return self

    def free(self):
        if getattr(self, "ptr", None) and _lib is not None:
            _lib.b200mc_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
