"""torchrun --nproc-per-node N tools/peer_gather_check.py: dist.PeerGather (CUDA IPC + copy engines) against the expected
result and against dist.OverlappedGather (NCCL), ragged pair counts included."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import pkgpath
pkg = pkgpath.load_package()
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
ok = True
for num_pairs in (world * 5, world * 5 - 1, 256):
    mine = pkg.dist.shard_pairs(num_pairs, rank, world)
    expect = torch.stack([torch.full((64, 32), float(p), device=dev) + torch.arange(32.0, device=dev) for p in range(num_pairs)])
    for cls in (pkg.dist.make_gather, pkg.dist.OverlappedGather):
        for rep in range(2):
            g = cls(num_pairs, (64, 32), dev, chunk=4)
            for slot, p in enumerate(mine):
                g.local[slot] = torch.full((64, 32), float(p), device=dev) + torch.arange(32.0, device=dev)
                g.pair_done(slot)
            out = g.finish()
            torch.cuda.synchronize()
            good = bool(torch.equal(out, expect))
            ok = ok and good
            if rank == 0:
                print(f'{type(g).__name__:16s} pairs {num_pairs:4d} rep {rep}: {"ok" if good else "MISMATCH"}', flush=True)
            del g
t = torch.tensor([int(ok)], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print('ALL OK' if int(t.item()) else 'FAILED', flush=True)
dist.barrier()
dist.destroy_process_group()
