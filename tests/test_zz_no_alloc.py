"""The allocation-free entry point ``b200mc_horseshoe_potential_noalloc`` (scratch in a
file-scope ``__device__`` array, no handle, no cudaMalloc) against the existing handle-based
``b200mc_potential_and_grad``, through the C ABI, on two calls in a row.  Collected last on
purpose."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


@pytest.fixture(scope="module")
def nb():
    import __graft_entry__ as g

    return g.build()


def _data(N, Dh, likelihood, seed):
    rng = np.random.RandomState(seed)
    X = rng.randn(N, Dh).astype(np.float32)
    Y = X[:, :3] @ np.array([2.0, -1.0, 0.5])
    Y = Y - Y.mean()
    if likelihood == 1:
        Y = Y + 0.05 * rng.randn(N)
    else:
        Y = rng.binomial(1, 1 / (1 + np.exp(-Y))).astype(np.float64)
    return X, Y.astype(np.float32)


@pytest.mark.parametrize("N,Dh,likelihood", [(700, 12, 1), (300, 300, 0), (50, 5, 0)])
def test_noalloc_matches_handle_based_entry_point_twice(nb, N, Dh, likelihood):
    from numpyro_b200 import _lib
    from numpyro_b200.random import PRNGKey

    lib = _lib.lib()
    X, Y = _data(N, Dh, likelihood, seed=N)
    model = nb.models.horseshoe_normal if likelihood == 1 else nb.models.horseshoe_bernoulli
    kern = nb.infer.NUTS(model)
    D = 2 * Dh + 1 + likelihood
    # existing entry point: a handle (its scratch comes from the caching allocator)
    z0 = {"lambdas": np.zeros(Dh, np.float32), "tau": np.zeros(1, np.float32),
          "unscaled_betas": np.zeros(Dh, np.float32)}
    if likelihood == 1:
        z0["prec_obs"] = np.float32(0.0)
    kern.init(PRNGKey(0), 0, z0, (X, Y), {})
    assert kern._bound.dim == D
    Xd, Yd = _lib.DeviceArray.from_numpy(X), _lib.DeviceArray.from_numpy(Y)
    n = 7
    Ud, gd = _lib.DeviceArray((n,), np.float32), _lib.DeviceArray((n, D), np.float32)
    rng = np.random.default_rng(5)
    for call in range(2):  # two calls in a row: the static scratch keeps the first call's contents
        zs = (rng.normal(size=(n, D)) * (0.3 + 0.4 * call)).astype(np.float32)
        U_ref, g_ref = kern.potential_and_grad(zs)
        zd = _lib.DeviceArray.from_numpy(zs)
        Ud.zero_()
        gd.zero_()
        _lib.check(lib.b200mc_horseshoe_potential_noalloc(Xd.ptr, Yd.ptr, N, Dh, likelihood, zd.ptr,
                                                          n, Ud.ptr, gd.ptr, None))
        _lib.check(lib.b200mc_stream_synchronize(None))
        U, g = Ud.numpy(), gd.numpy()
        assert np.isfinite(U).all() and np.isfinite(g).all()
        # same kernel, same grouping: tier-2 tolerance (rtol 1e-5) is the bar, equality expected
        np.testing.assert_allclose(U, U_ref, rtol=1e-5)
        np.testing.assert_allclose(g, g_ref, rtol=1e-5, atol=1e-5 * np.abs(g_ref).max())


def test_noalloc_rejects_sizes_beyond_the_static_scratch(nb):
    from numpyro_b200 import _lib

    lib = _lib.lib()
    d = _lib.DeviceArray((16,), np.float32)
    rc = lib.b200mc_horseshoe_potential_noalloc(d.ptr, d.ptr, 1 << 19, 3, 0, d.ptr, 3, d.ptr, d.ptr, None)
    assert rc != 0
    msg = lib.b200mc_last_error().decode()
    assert "exceed the static scratch" in msg and str(1 << 20) in msg
