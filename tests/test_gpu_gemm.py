"""GPU test of the split-precision (3xTF32) tcgen05 contraction behind the many-chain
potentials: out[c][m] = sum_k A[m][k] B[c][k] against float64 NumPy.

Tolerance: every product carries ~2^-22 relative error (hi*hi + hi*lo + lo*hi, the lo*lo
term dropped) and the tensor core accumulates in f32 (with truncation), so we require
|err| <= 2e-6 * (|A| |B|^T) (measured: <= 7e-7); a plain TF32 product (2^-11 = 5e-4) would
miss this bound by more than two orders of magnitude.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g

    g.build()
    from numpyro_b200 import _lib

    _lib.require_device()
    return _lib


def run_gemm(_lib, A, B, splits):
    lib = _lib.lib()
    M, K = A.shape
    Nc = B.shape[0]
    a = _lib.DeviceArray.from_numpy(np.ascontiguousarray(A, np.float32))
    b = _lib.DeviceArray.from_numpy(np.ascontiguousarray(B, np.float32))
    out = _lib.DeviceArray((Nc, M), np.float32)
    _lib.check(lib.b200mc_gemm3xtf32(a.ptr, b.ptr, M, Nc, K, splits, out.ptr, None))
    return out.numpy()


@pytest.mark.parametrize("M,K,Nc,splits", [
    (128, 32, 256, 1),        # one tile, one k-block
    (300, 100, 256, 1),       # ragged rows and k
    (1000, 10000, 200, 18),   # X^T R shape of the horseshoe config, split-K
    (10000, 1000, 256, 1),    # X beta shape of the horseshoe config
    (55, 4096, 64, 7),        # logistic X^T R shape (D = 55)
])
def test_gemm3xtf32_matches_float64(lib, M, K, Nc, splits):
    rng = np.random.default_rng(M + K)
    A = rng.normal(size=(M, K)).astype(np.float32)
    B = (rng.normal(size=(Nc, K)) * np.exp(rng.normal(size=(Nc, 1)))).astype(np.float32)
    got = run_gemm(lib, A, B, splits)
    want = B.astype(np.float64) @ A.astype(np.float64).T
    scale = np.abs(B).astype(np.float64) @ np.abs(A).astype(np.float64).T
    err = np.abs(got - want)
    assert np.all(err <= 2e-6 * scale + 1e-30), float((err / scale).max())


def test_gemm3xtf32_exact_on_small_integers(lib):
    # integers < 2^10 are exact in TF32: the result must be exact
    rng = np.random.default_rng(3)
    A = rng.integers(-8, 9, size=(256, 96)).astype(np.float32)
    B = rng.integers(-8, 9, size=(256, 96)).astype(np.float32)
    got = run_gemm(lib, A, B, 3)
    assert np.array_equal(got, (B.astype(np.int64) @ A.astype(np.int64).T).astype(np.float32))
