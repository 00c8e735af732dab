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


class OverlappedGather:
    """The same gather in chunks, overlapped with the computation that produces the pairs.

    `local` (per, *shape) is filled pair by pair by the caller (slot s = the rank's s-th pair, shard_pairs order), in
    slot order; after writing slot s on some stream the caller calls `pair_done(s)` under that stream.  Every `chunk`
    pairs one all_gather of that chunk is enqueued on a communication stream behind the events of its pairs, so all
    chunks but the last travel while later pairs are still being computed.  `finish()` makes the current stream wait
    for the last chunk and returns the (num_pairs, *shape) tensor with pair p at index p (a view of the gather buffer,
    which is laid out (slot, rank): no final reordering copy).
    Works without torch.distributed (world 1: no communication) and with gloo (CPU tensors: synchronous)."""

    def __init__(self, num_pairs, shape, device, dtype=torch.float32, chunk=4, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.per = pairs_per_rank(num_pairs, self.world)
        self.num_pairs, self.chunk = num_pairs, max(1, int(chunk))
        self.local = torch.zeros((self.per,) + tuple(shape), dtype=dtype, device=device)
        # (per, world, *shape): slot s of rank r = pair s * world + r, i.e. flat index s * world + r
        self.gathered = self.local.new_empty((self.per, self.world) + tuple(shape)) if self.world > 1 else None
        self.cuda = self.local.is_cuda
        self.comm = torch.cuda.Stream(device=device) if self.cuda and self.world > 1 else None
        self._events, self._sent = {}, 0

    def pair_done(self, slot):
        """Slot `slot` of `local` has been written on the current stream (slots complete in order)."""
        if self.world == 1:
            return
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record()
            self._events[slot] = ev
        if slot + 1 - self._sent >= self.chunk:
            self._send(self._sent, slot + 1)

    def _send(self, lo, hi):
        tail = tuple(self.local.shape[1:])
        stage = self.local.new_empty((self.world * (hi - lo),) + tail)  # all_gather concatenates rank-major along dim 0
        by_rank = stage.view((self.world, hi - lo) + tail)
        if self.cuda:
            with torch.cuda.stream(self.comm):
                for s in range(lo, hi):
                    self.comm.wait_event(self._events.pop(s))
                dist.all_gather_into_tensor(stage, self.local[lo:hi], group=self.group)
                self.gathered[lo:hi].copy_(by_rank.transpose(0, 1), non_blocking=True)
            stage.record_stream(self.comm)
        else:
            dist.all_gather_into_tensor(stage, self.local[lo:hi].contiguous(), group=self.group)
            self.gathered[lo:hi].copy_(by_rank.transpose(0, 1))
        self._sent = hi

    def finish(self):
        if self.world == 1:
            return self.local[:self.num_pairs]
        if self._sent < self.per:
            self._send(self._sent, self.per)
        if self.cuda:
            torch.cuda.current_stream(self.local.device).wait_stream(self.comm)
        self._events, self._sent = {}, 0
        return self.gathered.view((self.per * self.world,) + tuple(self.local.shape[1:]))[:self.num_pairs]
