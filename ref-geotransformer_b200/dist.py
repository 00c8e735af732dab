"""Pair sharding across ranks.  Registration pairs are independent, so the hot path needs no
collective: pair p runs on rank p mod W with a full weight replica, and the per-pair outputs are gathered once
at the end (NCCL all_gather over NVLink on GPUs; gloo in the CPU tests).  The reference itself runs batch size 1
per GPU under DDP (/root/reference/geotransformer/engine/base_trainer.py:66-72,190-198)."""
import os

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
                    ev = self._events.pop(s, None)  # a rank that owns one pair fewer has no event for the padding slot
                    if ev is not None:
                        self.comm.wait_event(ev)
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


class PeerGather(OverlappedGather):
    """The chunked gather without a collective kernel: every rank maps the other ranks' gather buffers through CUDA IPC
    (one node, NVLink / NVSwitch peer access) and WRITES its chunk into each of them with copy-engine copies
    (geotx_copy_async) on a communication stream, behind the events of the pairs of the chunk.  No NCCL kernel sits on
    SMs while pairs are being computed.  OPT-IN (GEOTX_GATHER=peer): correct (tools/peer_gather_check.py) but slower than
    the NCCL chunks as issued from Python today, see make_gather.  One barrier at the end (a 4-byte all_reduce on
    the communication stream of every rank) tells each rank that all peers' copies into its buffer have completed.
    CUDA tensors, world > 1, all ranks on one node; `make_gather` falls back to OverlappedGather otherwise."""

    def __init__(self, num_pairs, shape, device, dtype=torch.float32, chunk=4, group=None):
        super().__init__(num_pairs, shape, device, dtype=dtype, chunk=chunk, group=group)
        if not (self.cuda and self.world > 1):
            raise RuntimeError('PeerGather needs CUDA tensors and more than one rank')
        from . import _abi
        self._abi = _abi
        self.rank = dist.get_rank(group)
        st = self.gathered.untyped_storage()
        mine = (st._share_cuda_(), self.gathered.storage_offset())
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=group)
        self.peers, self._keep = [], []
        for r, (handle, offset) in enumerate(everyone):
            if r == self.rank:
                self.peers.append(self.gathered)
                continue
            storage = torch.UntypedStorage._new_shared_cuda(*handle)
            t = torch.empty(0, dtype=dtype, device=storage.device).set_(storage, offset, self.gathered.shape,
                                                                          self.gathered.stride())
            self._keep.append(storage)
            self.peers.append(t)
        self.slot_bytes = self.local[0].numel() * self.local.element_size()
        self._flag = torch.zeros(1, dtype=torch.int32, device=device)

    def _send(self, lo, hi):
        import ctypes
        with torch.cuda.stream(self.comm):
            for s in range(lo, hi):
                ev = self._events.pop(s, None)
                if ev is not None:
                    self.comm.wait_event(ev)
            stream = ctypes.c_void_p(self.comm.cuda_stream)
            for r in range(self.world):  # slot s of this rank -> gathered[s, rank] of rank r: one contiguous block each
                dst = self.peers[(self.rank + r) % self.world]  # start with the own buffer, then spread over the peers
                for s in range(lo, hi):
                    self._abi.call('geotx_copy_async', ctypes.c_void_p(dst[s, self.rank].data_ptr()),
                                   ctypes.c_void_p(self.local[s].data_ptr()), self.slot_bytes, stream)
        self._sent = hi

    def finish(self):
        if self._sent < self.per:
            self._send(self._sent, self.per)
        with torch.cuda.stream(self.comm):  # all ranks' copy streams meet here: every peer's copies into MY buffer are done
            dist.all_reduce(self._flag, group=self.group)
        torch.cuda.current_stream(self.local.device).wait_stream(self.comm)
        self._events, self._sent = {}, 0
        return self.gathered.view((self.per * self.world,) + tuple(self.local.shape[1:]))[:self.num_pairs]


def make_gather(num_pairs, shape, device, dtype=torch.float32, chunk=4, group=None):
    """OverlappedGather (NCCL all_gather per chunk: the default), or with GEOTX_GATHER=peer PeerGather where it applies
    (CUDA, several ranks, peer mapping succeeds on EVERY rank).  Measured, 3DMatch 1024, 20 steps: 2 GPUs 2447 (NCCL) /
    2413 (peer) pairs/s, 8 GPUs 9403 / 5011: the copy-engine version issues world x chunk copies per chunk from the
    Python thread that also launches the steps, and loses more there than the collective kernel costs on the SMs."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world > 1 and torch.device(device).type == 'cuda' and os.environ.get('GEOTX_GATHER', 'nccl') == 'peer':
        g, ok = None, 1
        try:
            g = PeerGather(num_pairs, shape, device, dtype=dtype, chunk=chunk, group=group)
        except Exception as exc:  # no IPC (containers without /dev/shm, several nodes ...): say so once, use NCCL
            ok = 0
            if dist.get_rank(group) == 0:
                import warnings
                warnings.warn(f'ref-geotransformer_b200: peer-memory gather unavailable ({type(exc).__name__}: {exc}); '
                              'using NCCL all_gather chunks')
        flag = torch.tensor([ok], dtype=torch.int32, device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 1:
            return g
    return OverlappedGather(num_pairs, shape, device, dtype=dtype, chunk=chunk, group=group)
