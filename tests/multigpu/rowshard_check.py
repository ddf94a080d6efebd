"""Row-sharded logistic NUTS over N GPUs (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \\
        --master-port 29511 tests/multigpu/rowshard_check.py

Checks: (1) every rank ends in the bit-identical chain state (replicated sampler, fixed-order
exchange); (2) the sharded run reproduces the single-GPU run's tree sizes for the first
transitions and its potential energies to float32 round-off (the partition changes the
summation order, nothing else); (3) U and grad at fixed points match the float64 oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import __graft_entry__ as g

    nb = g.build()
    from numpyro_b200 import _lib, distributed as D
    from oracle import potentials as opot

    _lib.check(_lib.lib().b200mc_set_device(local))
    rng = np.random.default_rng(0)
    N, Dm = 200003, 55
    X = rng.normal(size=(N, Dm)).astype(np.float32)
    X[:, -1] = 1.0
    y = (rng.random(N) < 1 / (1 + np.exp(-X @ (rng.normal(size=Dm) * 0.3)))).astype(np.uint8)
    key = nb.random.PRNGKey(1)
    kw = dict(step_size=0.003, adapt_step_size=False, adapt_mass_matrix=False)
    fields = ("num_steps", "potential_energy")

    kern = nb.infer.NUTS(nb.models.logistic_regression, **kw)
    from tools import torch_transport as TT

    Xl, yl = D.shard_rows(kern, X, y, rank, world, TT.allgather_bytes())
    mc = nb.infer.MCMC(kern, num_warmup=0, num_samples=40)
    mc.run(key, Xl, yl, extra_fields=fields)
    ex = mc.get_extra_fields()
    z_last = np.asarray(mc.last_state.z["coefs"])
    zs = (rng.normal(size=(2, Dm)) * 0.2).astype(np.float32)
    U, gr = kern.potential_and_grad(zs)

    # (1) all ranks identical
    allz = [None] * world
    dist.all_gather_object(allz, (z_last.tobytes(), ex["num_steps"].tobytes()))
    assert all(a == allz[0] for a in allz), "ranks diverged"
    # (3) potential / gradient against the float64 oracle (full data)
    fam = opot.Logistic(X, y, np.float64)
    for i in range(2):
        U0, g0 = fam.potential_and_grad(zs[i].astype(np.float64))
        np.testing.assert_allclose(U[i], U0, rtol=1e-5)
        np.testing.assert_allclose(gr[i], g0, rtol=1e-5, atol=1e-5 * np.abs(g0).max())
    # (2) against the single-GPU run (rank 0 only)
    if rank == 0:
        k1 = nb.infer.NUTS(nb.models.logistic_regression, **kw)
        m1 = nb.infer.MCMC(k1, num_warmup=0, num_samples=40)
        m1.run(key, X, y, extra_fields=fields)
        e1 = m1.get_extra_fields()
        assert np.array_equal(ex["num_steps"][:10], e1["num_steps"][:10]), (ex["num_steps"],
                                                                          e1["num_steps"])
        np.testing.assert_allclose(ex["potential_energy"][:10], e1["potential_energy"][:10],
                                   rtol=2e-5)
        print("rowshard_check ok: world", world, "num_steps", ex["num_steps"][:12].tolist())
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
