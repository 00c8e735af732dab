"""Two pyramid builds for one 3DMatch-sized pair (for ncu launch lists)."""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
import pkgpath

pkg = pkgpath.load_package()
sizes = (20000, 18000)
p = torch.cat([pkg.synth.make_dense_cloud(n, 3.0, 800 + b) for b, n in enumerate(sizes)]).cuda()
for _ in range(2):
    out = pkg.precompute_data_stack_mode(p, list(sizes), 4, 0.025, 0.0625, [38, 36, 36, 38])
torch.cuda.synchronize()
print([tuple(x.shape) for x in out['neighbors']], [l.tolist() for l in out['lengths']])
