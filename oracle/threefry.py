"""Oracle restatement of ``jax.random`` on the Threefry-2x32 generator.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Third-party algorithm: the
reference (NumPyro) calls ``jax.random.{split,uniform,normal,bernoulli}``
(call sites: ``numpyro/infer/hmc.py:102,335,472``;
``numpyro/infer/hmc_util.py:355,565,656,804,920,1005,1161-1162``;
``numpyro/infer/util.py:392,429-432``; ``numpyro/infer/mcmc.py:662``).  JAX is
not vendored under /root/reference (``setup.py:12-13`` says ``jax>=0.4.25``), so
this restates the published algorithm:

* Threefry-2x32, 20 rounds (Salmon et al., "Parallel random numbers: as easy as
  1, 2, 3", SC'11; Random123 ``threefry2x32_20``), rotation constants
  ``[13, 15, 26, 6]`` / ``[17, 29, 16, 24]``, key schedule constant 0x1BD11BDA.
* ``jax/_src/prng.py`` counter layouts: the *legacy* layout
  (``jax_threefry_partitionable=False``, default before JAX 0.5.0) and the
  *partitionable* layout (default from 0.5.0).
* ``jax/_src/random.py``: ``uniform`` (mantissa trick), ``normal``
  (``sqrt(2) * erf_inv(uniform(nextafter(-1, 0), 1))``), ``bernoulli``.
* XLA's float32 ``erf_inv`` (Giles, "Approximating the erfinv function",
  single-precision polynomial -- the one in ``xla/client/lib/math.cc``).

Parity: pinned by Random123 KATs + values printed in JAX's documentation
(``tests/test_oracle_golden.py``); otherwise "parity unpinned" vs a live JAX.
"""
import numpy as np

_U32 = np.uint32
_ROT = ((13, 15, 26, 6), (17, 29, 16, 24))


def _rotl(x, r):
    return ((x << _U32(r)) | (x >> _U32(32 - r))).astype(_U32)


def threefry2x32(k0, k1, x0, x1):
    """Threefry-2x32-20 block function on uint32 arrays (broadcasting)."""
    with np.errstate(over="ignore"):
        k0 = np.asarray(k0, _U32)
        k1 = np.asarray(k1, _U32)
        x0 = np.asarray(x0, _U32).copy()
        x1 = np.asarray(x1, _U32).copy()
        ks = (k0, k1, (k0 ^ k1 ^ _U32(0x1BD11BDA)).astype(_U32))
        x0 = (x0 + ks[0]).astype(_U32)
        x1 = (x1 + ks[1]).astype(_U32)
        for i in range(5):
            for r in _ROT[i % 2]:
                x0 = (x0 + x1).astype(_U32)
                x1 = _rotl(x1, r)
                x1 = (x1 ^ x0).astype(_U32)
            x0 = (x0 + ks[(i + 1) % 3]).astype(_U32)
            x1 = (x1 + ks[(i + 2) % 3] + _U32(i + 1)).astype(_U32)
    return x0, x1


def PRNGKey(seed):
    """``jax.random.PRNGKey`` for the default (32-bit mode) threefry impl."""
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    # with x64 disabled the seed is an int32: the high word is the sign extension
    # shifted out, i.e. 0 for non-negative python ints below 2**32.
    return np.array([(seed >> 32) & 0xFFFFFFFF, seed & 0xFFFFFFFF], dtype=_U32)


def _threefry_2x32_counts(key, counts):
    """``jax._src.prng.threefry_2x32``: halves trick with odd-size padding."""
    counts = np.asarray(counts, _U32).ravel()
    n = counts.size
    odd = n % 2
    if odd:
        counts = np.concatenate([counts, np.zeros(1, _U32)])
    m = counts.size // 2
    o0, o1 = threefry2x32(key[0], key[1], counts[:m], counts[m:])
    out = np.concatenate([o0, o1])
    return out[:n] if odd else out


def split(key, num=2, partitionable=False):
    """``jax.random.split(key, num)`` -> uint32[num, 2]."""
    key = np.asarray(key, _U32)
    if partitionable:
        # _threefry_split_foldlike: key_i = threefry(key, (hi=0, lo=i))
        lo = np.arange(num, dtype=_U32)
        b0, b1 = threefry2x32(key[0], key[1], np.zeros(num, _U32), lo)
        return np.stack([b0, b1], axis=-1)
    out = _threefry_2x32_counts(key, np.arange(2 * num, dtype=_U32))
    return out.reshape(num, 2)


def random_bits(key, shape=(), partitionable=False):
    """32-bit ``jax.random.bits`` for a 1-key threefry generator."""
    key = np.asarray(key, _U32)
    shape = tuple(shape) if not np.isscalar(shape) else (int(shape),)
    n = int(np.prod(shape)) if len(shape) else 1
    if partitionable:
        lo = np.arange(n, dtype=_U32)
        b0, b1 = threefry2x32(key[0], key[1], np.zeros(n, _U32), lo)
        bits = (b0 ^ b1).astype(_U32)
    else:
        bits = _threefry_2x32_counts(key, np.arange(n, dtype=_U32))
    return bits.reshape(shape)


def _bits_to_unit_float(bits):
    """float32 in [0, 1): ``bitcast((bits >> 9) | 0x3F800000) - 1``."""
    f = ((bits >> _U32(9)) | _U32(0x3F800000)).astype(_U32).view(np.float32)
    return f - np.float32(1.0)


def uniform(key, shape=(), minval=0.0, maxval=1.0, partitionable=False):
    """``jax.random.uniform`` (float32)."""
    minval = np.float32(minval)
    maxval = np.float32(maxval)
    u = _bits_to_unit_float(random_bits(key, shape, partitionable))
    scale = np.float32(maxval - minval)
    return np.maximum(minval, (u * scale + minval).astype(np.float32)).astype(np.float32)


def bernoulli(key, p=0.5, shape=(), partitionable=False):
    """``jax.random.bernoulli``: ``uniform(key) < p``."""
    p = np.asarray(p, np.float32)
    return uniform(key, shape, partitionable=partitionable) < p


def erf_inv_f32(x):
    """XLA's float32 ``erf_inv`` (Giles' single-precision polynomial)."""
    x = np.asarray(x, np.float32)
    f = np.float32
    with np.errstate(divide="ignore", invalid="ignore"):
        w = -np.log1p((-x * x).astype(f)).astype(f)
        lt = w < f(5.0)
        w1 = (w - f(2.5)).astype(f)
        w2 = (np.sqrt(np.where(lt, f(9.0), w)).astype(f) - f(3.0)).astype(f)
        ww = np.where(lt, w1, w2).astype(f)
        c_lt = (2.81022636e-08, 3.43273939e-07, -3.5233877e-06, -4.39150654e-06,
                0.00021858087, -0.00125372503, -0.00417768164, 0.246640727, 1.50140941)
        c_ge = (-0.000200214257, 0.000100950558, 0.00134934322, -0.00367342844,
                0.00573950773, -0.0076224613, 0.00943887047, 1.00167406, 2.83297682)
        p = np.where(lt, f(c_lt[0]), f(c_ge[0])).astype(f)
        for a, b in zip(c_lt[1:], c_ge[1:]):
            p = (np.where(lt, f(a), f(b)) + (p * ww).astype(f)).astype(f)
        out = (p * x).astype(f)
        out = np.where(np.abs(x) == f(1.0), np.copysign(f(np.inf), x), out)
    return out.astype(f)


def normal(key, shape=(), partitionable=False):
    """``jax.random.normal`` (float32): ``sqrt(2) * erf_inv(u)``, u in (-1, 1)."""
    lo = np.nextafter(np.float32(-1.0), np.float32(0.0))
    u = uniform(key, shape, minval=lo, maxval=1.0, partitionable=partitionable)
    return (np.float32(np.sqrt(2)) * erf_inv_f32(u)).astype(np.float32)
