"""Row-sharded logistic NUTS over 2 GPUs (BASELINE configs[2]'s partitioning; the reference's
analogue is GSPMD row sharding, numpyro/infer/mcmc.py:239-265), as a collected GPU test: it
launches tests/multigpu/rowshard_check.py under torchrun with one rank per GPU and skips on a box
with fewer than two devices.  The script checks that every rank ends in the bit-identical chain
state, that the sharded chain reproduces the single-GPU chain's tree sizes and energies, and
that U and grad match the float64 oracle on the full data."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _device_count():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g

    g.build()
    from numpyro_b200 import _lib

    return _lib.lib().b200mc_device_count()


@pytest.mark.parametrize("world", [2])
def test_row_sharded_chain_is_replicated_and_matches_single_gpu(world):
    if _device_count() < world:
        pytest.skip(f"needs {world} CUDA devices")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "multigpu", "rowshard_check.py")]
    out = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert "rowshard_check ok" in out.stdout


def test_parallel_chain_method_over_local_gpus_equals_vectorized():
    """chain_method="parallel" (mcmc.py:707-708): chains spread over the local GPUs from one
    process, one handle per device; chain c equals the vectorized one-GPU run bit for bit."""
    import numpy as np

    if _device_count() < 2:
        pytest.skip("needs 2 CUDA devices")
    import numpyro_b200 as nb

    y = np.array([28.0, 8.0, -3.0, 7.0, -1.0, 1.0, 18.0, 12.0])
    sg = np.array([15.0, 10.0, 16.0, 11.0, 9.0, 11.0, 10.0, 18.0])
    out = {}
    for method in ("vectorized", "parallel"):
        mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.eight_schools_noncentered), num_warmup=100,
                           num_samples=50, num_chains=6, chain_method=method)
        mc.warmup(nb.random.PRNGKey(3), 8, sg, y=y)
        mc.run(mc.post_warmup_state.rng_key, 8, sg, y=y, extra_fields=("num_steps",))
        out[method] = (mc.get_samples(group_by_chain=True), mc.get_extra_fields(group_by_chain=True))
        assert np.asarray(mc.last_state.i).shape == (6,)
    for site in ("mu", "tau", "theta"):
        assert np.array_equal(out["vectorized"][0][site], out["parallel"][0][site]), site
    assert np.array_equal(out["vectorized"][1]["num_steps"], out["parallel"][1]["num_steps"])
