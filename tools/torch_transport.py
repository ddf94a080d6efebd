"""``torch.distributed`` transports for ``numpyro_b200.distributed`` (bench / test plumbing;
the product package itself imports no communication library)."""
import os


def env():
    """(rank, world_size, local_rank) as torchrun exports them."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def gather(group=None, dst=0):
    import torch.distributed as dist

    def fn(obj):
        world = dist.get_world_size(group)
        out = [None] * world if dist.get_rank(group) == dst else None
        dist.gather_object(obj, out, dst=dst, group=group)
        return out

    return fn


def allgather_bytes(group=None):
    import torch.distributed as dist

    def fn(payload):
        out = [None] * dist.get_world_size(group)
        dist.all_gather_object(out, bytes(payload), group=group)
        return out

    return fn
