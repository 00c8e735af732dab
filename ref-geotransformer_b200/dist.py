"""Pair sharding across ranks.  Registration pairs are independent, so the hot path needs no
collective: pair p runs on rank p mod W with a full weight replica, and the per-pair outputs are gathered once
at the end (NCCL all_gather over NVLink on GPUs; gloo in the CPU tests).  The reference itself runs batch size 1
per GPU under DDP (/root/reference/geotransformer/engine/base_trainer.py:66-72,190-198)."""
import torch
import torch.distributed as dist


def shard_pairs(num_pairs, rank, world_size):
    """Indices of the pairs rank `rank` owns (round-robin: p mod W == rank)."""
    return list(range(rank, num_pairs, world_size))


def pairs_per_rank(num_pairs, world_size):
    return (num_pairs + world_size - 1) // world_size


def gather_pair_outputs(local_outputs, num_pairs, group=None):
    """local_outputs: list (one per owned pair, in shard_pairs order) of equally-shaped tensors.
    Returns a tensor (num_pairs, *shape) on every rank with pair p at index p.  One all_gather."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    per = pairs_per_rank(num_pairs, world)
    if not local_outputs:
        raise ValueError('every rank must own at least one pair (num_pairs >= world_size)')
    shape = local_outputs[0].shape
    buf = local_outputs[0].new_zeros((per,) + tuple(shape))
    for slot, t in enumerate(local_outputs):
        buf[slot].copy_(t)
    if world == 1:
        return buf[:num_pairs]
    gathered = buf.new_empty((world * per,) + tuple(shape))
    dist.all_gather_into_tensor(gathered, buf, group=group)
    # rank r, slot s  ->  pair s*world + r
    out = gathered.view(world, per, *shape).transpose(0, 1).reshape(world * per, *shape)
    return out[:num_pairs]
