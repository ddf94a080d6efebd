"""Chain and row sharding over ranks: the multi-GPU forms of ``chain_method="parallel"``
(``mcmc.py:700-712``) and of GSPMD row sharding (``mcmc.py:239-265``), one process per GPU.

This module holds the host LOGIC only (which rank owns which chains / rows, how results are
put back together).  It imports no communication library: the caller injects the transport as
plain callables --

``gather(obj) -> list[obj] | None``
    gather one picklable object per rank on rank 0 (rank order); ``None`` on the other ranks.
``allgather(payload: bytes) -> list[bytes]``
    every rank gets every rank's payload (rank order); used once, for the CUDA IPC handles of
    row sharding.

``tools/torch_transport.py`` builds both from ``torch.distributed`` for ``bench.py`` and the
tests; any other launcher (MPI, a socket) can supply its own.

Chains: the reference draws chains in parallel with ``pmap`` over devices after
``rng_key = split(rng_key, num_chains)`` (``mcmc.py:661-662``).  Here every rank takes a
contiguous slice of those per-chain keys and runs it with the vectorized kernel; there is no
data-path communication, only one gather of the collected draws at the end, and chain c is
bit-identical whatever the number of ranks.
"""
import numpy as np


def shard_bounds(num_chains, rank, world_size):
    """Contiguous slice [lo, hi) of the chains (or rows) owned by ``rank``: the first
    ``num_chains % world_size`` ranks take one extra."""
    if not (0 <= rank < world_size):
        raise ValueError("rank must be in [0, world_size)")
    base, extra = divmod(int(num_chains), int(world_size))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_keys(chain_keys, rank, world_size):
    """Rows of the ``[num_chains, 2]`` per-chain key array owned by ``rank``."""
    chain_keys = np.asarray(chain_keys, np.uint32)
    if chain_keys.ndim != 2 or chain_keys.shape[1] != 2:
        raise ValueError("chain_keys must be uint32[num_chains, 2] (already split)")
    lo, hi = shard_bounds(chain_keys.shape[0], rank, world_size)
    return chain_keys[lo:hi]


def _tree_map2(fn, trees):
    t0 = trees[0]
    if isinstance(t0, dict):
        return {k: _tree_map2(fn, [t[k] for t in trees]) for k in t0}
    return fn(trees)


def concat_chains(per_rank):
    """Concatenate per-rank results (pytrees of ``[local_chains, draws, ...]`` arrays, rank
    order == chain order) along the chain axis; ranks that owned no chain contribute ``{}``."""
    per_rank = [o for o in per_rank if o is not None and not (isinstance(o, dict) and not o)]
    if not per_rank:
        return {}
    return _tree_map2(lambda xs: np.concatenate([x for x in xs if x.shape[0]], axis=0), per_rank)


def run_sharded(make_mcmc, chain_keys, *args, rank=0, world_size=1, gather=None,
                extra_fields=(), **kwargs):
    """Run ``len(chain_keys)`` chains sharded over ``world_size`` ranks.

    ``make_mcmc(num_local_chains)`` builds this rank's ``MCMC`` (``chain_method="vectorized"``);
    ``gather`` is the transport described in the module docstring (needed when
    ``world_size > 1``).  Returns ``(samples, extra_fields)`` grouped by chain on rank 0,
    ``(None, None)`` elsewhere.
    """
    if world_size > 1 and gather is None:
        raise ValueError("run_sharded over more than one rank needs a `gather` transport")
    keys = shard_keys(chain_keys, rank, world_size)
    samples, extras = {}, {}
    if keys.shape[0]:
        mcmc = make_mcmc(keys.shape[0])
        mcmc.run(keys, *args, extra_fields=tuple(extra_fields), **kwargs)
        samples = mcmc.get_samples(group_by_chain=True)
        extras = mcmc.get_extra_fields(group_by_chain=True)
    if world_size == 1:
        return samples, extras
    got = gather({"samples": samples, "extras": extras})
    if got is None:
        return None, None
    return (concat_chains([g["samples"] for g in got]), concat_chains([g["extras"] for g in got]))


def shard_rows(kernel, data, labels, rank, world_size, allgather=None):
    """Configure ``kernel`` for row sharding and return this rank's ``(data, labels)`` slice
    (contiguous rows, same split as ``shard_bounds``)."""
    if world_size > 1 and allgather is None:
        raise ValueError("row sharding over more than one rank needs an `allgather` transport")
    lo, hi = shard_bounds(len(data), rank, world_size)
    kernel.shard_rows(rank, world_size, allgather if world_size > 1 else None)
    return data[lo:hi], labels[lo:hi]
