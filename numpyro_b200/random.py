"""Host-side handles on the device Threefry generator (``jax.random`` surface used by
the reference: ``PRNGKey``, ``split``, ``uniform``, ``normal``, ``bernoulli``).  Keys are
``uint32[2]`` numpy arrays; every draw runs in the CUDA library (no CPU generator)."""
import numpy as np

from . import _lib

#: counter layout of ``jax.random``: True = ``jax_threefry_partitionable`` (JAX >= 0.5 default)
DEFAULT_PARTITIONABLE = True


def PRNGKey(seed):
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    return np.array([(seed >> 32) & 0xFFFFFFFF, seed & 0xFFFFFFFF], dtype=np.uint32)


def is_prng_key(key):
    """numpyro/util.py:154-160."""
    try:
        key = np.asarray(key)
        return key.shape == (2,) and key.dtype == np.uint32
    except Exception:
        return False


def _part(partitionable):
    return int(DEFAULT_PARTITIONABLE if partitionable is None else partitionable)


def split(key, num=2, partitionable=None):
    _lib.require_device()
    key = np.ascontiguousarray(key, np.uint32)
    batch = key.shape[:-1]
    flat = key.reshape(-1, 2)
    kd = _lib.DeviceArray.from_numpy(flat)
    out = _lib.DeviceArray((flat.shape[0], num, 2), np.uint32)
    _lib.check(_lib.lib().b200mc_random_split(kd.ptr, flat.shape[0], num, out.ptr,
                                              _part(partitionable), None))
    return out.numpy().reshape(batch + (num, 2))


def _draw(fn, key, shape, dtype, *args, partitionable=None):
    _lib.require_device()
    shape = (shape,) if np.isscalar(shape) else tuple(shape)
    n = int(np.prod(shape)) if shape else 1
    kd = _lib.DeviceArray.from_numpy(np.ascontiguousarray(key, np.uint32))
    out = _lib.DeviceArray((n,), dtype)
    _lib.check(fn(kd.ptr, n, *args, out.ptr, _part(partitionable), None))
    return out.numpy().reshape(shape)


def bits(key, shape=(), partitionable=None):
    return _draw(_lib.lib().b200mc_random_bits, key, shape, np.uint32, partitionable=partitionable)


def uniform(key, shape=(), minval=0.0, maxval=1.0, partitionable=None):
    return _draw(_lib.lib().b200mc_random_uniform, key, shape, np.float32, float(minval),
                 float(maxval), partitionable=partitionable)


def normal(key, shape=(), partitionable=None):
    return _draw(_lib.lib().b200mc_random_normal, key, shape, np.float32,
                 partitionable=partitionable)


def bernoulli(key, p=0.5, shape=(), partitionable=None):
    return uniform(key, shape, partitionable=partitionable) < np.float32(p)


def threefry2x32(key, counter):
    _lib.require_device()
    key = np.ascontiguousarray(key, np.uint32).reshape(-1, 2)
    ctr = np.ascontiguousarray(counter, np.uint32).reshape(-1, 2)
    kd, cd = _lib.DeviceArray.from_numpy(key), _lib.DeviceArray.from_numpy(ctr)
    out = _lib.DeviceArray(key.shape, np.uint32)
    _lib.check(_lib.lib().b200mc_threefry2x32(kd.ptr, cd.ptr, out.ptr, key.shape[0], None))
    return out.numpy()
