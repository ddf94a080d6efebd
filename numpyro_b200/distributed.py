"""Chain sharding over ranks: the multi-GPU form of ``chain_method="parallel"``.

The reference draws chains in parallel with ``pmap`` over devices (``mcmc.py:700-712``) after
``rng_key = split(rng_key, num_chains)`` (``mcmc.py:661-662``).  Here every rank (one process
per GPU, ``torch.distributed``) takes a contiguous slice of those per-chain keys and runs its
slice with the vectorized kernel; there is no data-path communication, only one gather of the
collected draws at the end.  Chain c is bit-identical whatever the number of ranks.

``torch.distributed`` is plumbing only and is imported lazily, when world_size > 1.
"""
import numpy as np


def shard_bounds(num_chains, rank, world_size):
    """Contiguous slice [lo, hi) of the chains owned by ``rank``: the first
    ``num_chains % world_size`` ranks take one extra chain."""
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


def gather_chains(local, group=None, dst=0):
    """Concatenate per-rank results (pytrees of ``[local_chains, draws, ...]`` arrays) along
    the chain axis on rank ``dst`` (None elsewhere).  Rank order == chain order."""
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    out = [None] * world if dist.get_rank(group) == dst else None
    dist.gather_object(local, out, dst=dst, group=group)
    if out is None:
        return None
    out = [o for o in out if o is not None]
    return _tree_map2(lambda xs: np.concatenate([x for x in xs if x.shape[0]], axis=0), out)


def run_sharded(make_mcmc, chain_keys, *args, rank=None, world_size=None, group=None,
                extra_fields=(), **kwargs):
    """Run ``len(chain_keys)`` chains sharded over the ranks of ``group``.

    ``make_mcmc(num_local_chains)`` builds this rank's ``MCMC`` (``chain_method="vectorized"``).
    Returns ``(samples, extra_fields)`` grouped by chain on rank 0, ``(None, None)`` elsewhere.
    """
    if rank is None or world_size is None:
        import torch.distributed as dist

        rank, world_size = dist.get_rank(group), dist.get_world_size(group)
    keys = shard_keys(chain_keys, rank, world_size)
    samples, extras = {}, {}
    if keys.shape[0]:
        mcmc = make_mcmc(keys.shape[0])
        mcmc.run(keys, *args, extra_fields=tuple(extra_fields), **kwargs)
        samples = mcmc.get_samples(group_by_chain=True)
        extras = mcmc.get_extra_fields(group_by_chain=True)
    if world_size == 1:
        return samples, extras
    got = gather_chains({"samples": samples, "extras": extras}, group=group)
    if got is None:
        return None, None
    return got["samples"], got["extras"]


def allgather_bytes(group=None):
    """Host transport for ``HMC.shard_rows``: ``torch.distributed.all_gather_object``."""
    import torch.distributed as dist

    def allgather(payload):
        out = [None] * dist.get_world_size(group)
        dist.all_gather_object(out, bytes(payload), group=group)
        return out

    return allgather


def shard_rows(kernel, data, labels, rank=None, world_size=None, group=None):
    """Configure ``kernel`` for row sharding and return this rank's ``(data, labels)`` slice
    (contiguous rows, same split as ``shard_bounds``)."""
    if rank is None or world_size is None:
        import torch.distributed as dist

        rank, world_size = dist.get_rank(group), dist.get_world_size(group)
    lo, hi = shard_bounds(len(data), rank, world_size)
    kernel.shard_rows(rank, world_size, allgather_bytes(group) if world_size > 1 else None)
    return data[lo:hi], labels[lo:hi]
