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
