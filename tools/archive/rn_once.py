"""One stage-0-sized radius search with both methods (for ncu launch lists)."""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
import pkgpath

pkg = pkgpath.load_package()
sizes = (20000, 18000)
s = torch.cat([pkg.synth.make_dense_cloud(n, 3.0, 800 + b) for b, n in enumerate(sizes)]).cuda()
for method in ('grid', 'grid', 'brute'):
    out = pkg.ops.radius_neighbors(s, s, list(sizes), list(sizes), 0.0625, 38, method=method)
torch.cuda.synchronize()
print(int(out[2]), float(out[1].float().mean()))
