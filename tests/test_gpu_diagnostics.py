"""On-device diagnostics (csrc/diag_kernel.cuh, ``b200mc_diag_ess_rhat``) against the reference's
own outputs (``tests/golden/refpin.npz`` ``diag/*``: numpyro/diagnostics.py executed from
/root/reference) and against the host restatement on the same float32 data."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
G = np.load(os.path.join(HERE, "golden", "refpin.npz"))


@pytest.fixture(scope="module")
def nb():
    import __graft_entry__ as g

    return g.build()


@pytest.mark.parametrize("name", ["ar", "anti"])
def test_device_diagnostics_match_reference_outputs(nb, name):
    x = G[f"diag/{name}/x"]  # [chain, draw, D] float64; the device works on its float32 rounding
    got = nb.diagnostics.device_ess_rhat(x)
    np.testing.assert_allclose(got["n_eff"], G[f"diag/{name}/ess"], rtol=2e-5)
    np.testing.assert_allclose(got["r_hat"], G[f"diag/{name}/gelman_rubin"], rtol=1e-6)
    np.testing.assert_allclose(got["split_r_hat"], G[f"diag/{name}/split_gelman_rubin"], rtol=1e-6)


@pytest.mark.parametrize("C,N,D", [(1, 50, 3), (4, 1000, 55), (3, 257, 70), (64, 301, 10), (2, 5000, 5)])
def test_device_diagnostics_match_host_restatement(nb, C, N, D):
    from oracle import diagnostics as od

    rng = np.random.default_rng(C * 1000 + N)
    phi = rng.uniform(-0.5, 0.95, size=D)
    x = np.zeros((C, N, D))
    e = rng.normal(size=(C, N, D))
    for i in range(1, N):
        x[:, i] = phi * x[:, i - 1] + e[:, i]
    x = (x * rng.uniform(0.1, 100.0, size=D) + rng.normal(size=D) * 10).astype(np.float32)
    got = nb.diagnostics.device_ess_rhat(x)
    x64 = x.astype(np.float64)
    np.testing.assert_allclose(got["n_eff"], od.effective_sample_size(x64), rtol=1e-7)
    np.testing.assert_allclose(got["split_r_hat"], od.split_gelman_rubin(x64), rtol=1e-9)
    if C > 1:
        np.testing.assert_allclose(got["r_hat"], od.gelman_rubin(x64), rtol=1e-9)
    # a window of the draws, straight from a device buffer in the collection layout
    from numpyro_b200 import _lib

    xd = _lib.DeviceArray.from_numpy(np.ascontiguousarray(np.swapaxes(x, 0, 1)))
    first, n = N // 5, N // 2
    win = nb.diagnostics.device_ess_rhat(xd, first=first, n=n, layout="draw_chain")
    np.testing.assert_allclose(win["n_eff"], od.effective_sample_size(x64[:, first:first + n]), rtol=1e-7)


def test_device_diagnostics_many_chains(nb):
    """16,384 chains x 10 parameters x 200 draws (the eight-schools bench shape): finite, and equal
    to the host restatement on a subset of the parameters."""
    from oracle import diagnostics as od

    rng = np.random.default_rng(7)
    x = rng.normal(size=(16384, 200, 10)).astype(np.float32).cumsum(axis=1) * 0.05 + rng.normal(size=(16384, 200, 10)).astype(np.float32)
    got = nb.diagnostics.device_ess_rhat(x)
    assert np.isfinite(got["n_eff"]).all() and np.isfinite(got["split_r_hat"]).all()
    ref = od.effective_sample_size(x[:, :, :2].astype(np.float64))
    np.testing.assert_allclose(got["n_eff"][:2], ref, rtol=1e-7)
