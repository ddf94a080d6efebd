"""World-size-2 ``gloo`` test of the chain-sharding host logic (numpyro_b200.distributed),
run on CPU: a tiny pure-Python ``MCMCKernel`` stands in for the CUDA kernel, so what is
checked is the rank slicing of per-chain keys, the driver and the final gather -- chain c
must be identical to the single-process run whatever the rank count."""
from collections import namedtuple
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

State = namedtuple("State", ["z", "rng_key"])


def _make_mcmc(n_local):
    from numpyro_b200.infer import MCMC, MCMCKernel

    class RandomWalk(MCMCKernel):
        """z' = z + N(0,1) driven only by the chain's key (deterministic per chain)."""
        sample_field = "z"

        def init(self, rng_key, num_warmup, init_params, model_args, model_kwargs):
            keys = np.asarray(rng_key, np.uint32).reshape(-1, 2)
            self._sample_fn = True
            return State(np.zeros((keys.shape[0], 3)), keys.copy())

        def sample(self, state, model_args, model_kwargs):
            z, keys = state.z.copy(), state.rng_key.copy()
            for c in range(keys.shape[0]):
                rng = np.random.default_rng(keys[c].astype(np.uint64))
                z[c] += rng.normal(size=3)
                keys[c] = rng.integers(0, 2**32, size=2, dtype=np.uint32)
            return State(z, keys)

    return MCMC(RandomWalk(), num_warmup=3, num_samples=7, num_chains=n_local,
                chain_method="vectorized")


def _chain_keys(n):
    from oracle import threefry as T

    return T.split(T.PRNGKey(5), n, True)


def _worker(rank, world, port, n_chains, out_path):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from numpyro_b200 import distributed as D
    from tools import torch_transport as TT

    samples, extras = D.run_sharded(_make_mcmc, _chain_keys(n_chains), rank=rank, world_size=world,
                                    gather=TT.gather())
    if rank == 0:
        np.save(out_path, samples)
    else:
        assert samples is None and extras is None
    dist.barrier()
    dist.destroy_process_group()


def test_shard_bounds_cover_all_chains():
    from numpyro_b200 import distributed as D

    for n in (1, 2, 7, 8, 16384):
        for w in (1, 2, 3, 8):
            b = [D.shard_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        D.shard_bounds(4, 2, 2)
    with pytest.raises(ValueError):
        D.shard_keys(np.zeros(2, np.uint32), 0, 1)


def test_package_imports_no_communication_library():
    # north_star: PyTorch is plumbing for bench/tests, never part of the product package
    import subprocess

    code = ("import sys; import numpyro_b200, numpyro_b200.distributed, numpyro_b200.infer; "
            "bad = [m for m in sys.modules if m == 'torch' or m.startswith('torch.')]; "
            "assert not bad, bad")
    subprocess.check_call([sys.executable, "-c", code], cwd=ROOT)
    src = open(os.path.join(ROOT, "numpyro_b200", "distributed.py")).read()
    assert "import torch" not in src
    from numpyro_b200 import distributed as D

    with pytest.raises(ValueError):
        D.run_sharded(_make_mcmc, _chain_keys(4), rank=0, world_size=2)


@pytest.mark.parametrize("n_chains", [5, 8])
def test_two_rank_gloo_matches_single_process(tmp_path, n_chains):
    import torch.multiprocessing as mp

    from numpyro_b200 import distributed as D

    ref, _ = D.run_sharded(_make_mcmc, _chain_keys(n_chains), rank=0, world_size=1)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "gathered.npy")
    mp.spawn(_worker, args=(2, port, n_chains, out), nprocs=2, join=True)
    got = np.load(out)
    assert got.shape == (n_chains, 7, 3)
    assert np.array_equal(got, ref)


def test_sequential_chain_method_equals_vectorized():
    # mcmc.py:704-705 vs :709-710: per-chain results do not depend on the method
    from numpyro_b200.infer import MCMC

    keys = _chain_keys(3)
    out = {}
    for method in ("vectorized", "sequential"):
        mc = _make_mcmc(3)
        mc.chain_method = method
        mc.run(keys)
        out[method] = mc.get_samples(group_by_chain=True)
        assert out[method].shape == (3, 7, 3)
        assert np.asarray(mc.last_state.z).shape == (3, 3)
    assert np.array_equal(out["vectorized"], out["sequential"])


def test_parallel_warns_like_the_reference_when_devices_are_short():
    import warnings

    from numpyro_b200.infer import MCMC

    mc = _make_mcmc(1)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        MCMC(mc.sampler, num_warmup=1, num_samples=1, num_chains=64, chain_method="parallel")
    assert any("not enough devices to run parallel chains" in str(x.message) for x in w)
