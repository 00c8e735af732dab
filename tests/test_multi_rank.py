"""world_size-2 gloo test (CPU) of the pair sharding + single gather used for multi-GPU runs."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, num_pairs, results):
    sys.path.insert(0, ROOT)
    import pkgpath
    pkg = pkgpath.load_package()
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    mine = pkg.dist.shard_pairs(num_pairs, rank, world)
    # stand-in for the per-pair forward: a deterministic function of the pair index
    outs = [torch.full((3, 4), float(p)) + torch.arange(4.0) for p in mine]
    full = pkg.dist.gather_pair_outputs(outs, num_pairs)
    expect = torch.stack([torch.full((3, 4), float(p)) + torch.arange(4.0) for p in range(num_pairs)])
    # the chunked gather that overlaps with the computation (what bench.py times): same result, any chunk size
    ok = True
    for chunk in (1, 2, 4):
        og = pkg.dist.OverlappedGather(num_pairs, (3, 4), torch.device('cpu'), chunk=chunk)
        for slot, p in enumerate(mine):
            og.local[slot] = torch.full((3, 4), float(p)) + torch.arange(4.0)
            og.pair_done(slot)
        ok = ok and bool(torch.equal(og.finish(), expect))
    results[rank] = bool(torch.equal(full, expect)) and mine == list(range(rank, num_pairs, world)) and ok
    dist.barrier()
    dist.destroy_process_group()


def _run(num_pairs, port):
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(_worker, args=(2, port, num_pairs, results), nprocs=2, join=True)
    assert results[0] and results[1]


def test_shard_and_gather_even():
    _run(6, 29611)


def test_shard_and_gather_ragged():
    _run(5, 29612)  # rank 1 owns one pair fewer; padding slot must not leak into the result


def test_single_process_gather():
    sys.path.insert(0, ROOT)
    import pkgpath
    pkg = pkgpath.load_package()
    outs = [torch.ones(2) * p for p in range(3)]
    assert torch.equal(pkg.dist.gather_pair_outputs(outs, 3), torch.stack(outs))
    assert pkg.dist.shard_pairs(7, 1, 4) == [1, 5]
    og = pkg.dist.OverlappedGather(3, (2,), torch.device('cpu'), chunk=2)
    for slot in range(3):
        og.local[slot] = outs[slot]
        og.pair_done(slot)
    assert torch.equal(og.finish(), torch.stack(outs))
