"""Three launches of the tabulated embedding kernel at n = 1024, C = 256 (for ncu captures)."""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
import pkgpath

pkg = pkgpath.load_package()
torch.manual_seed(1)
m = pkg.GeometricStructureEmbedding(256, 0.2, 15, 3).cuda().eval()
pts = pkg.synth.make_pair('3dmatch', 1024)[0].cuda()
with torch.no_grad():
    geom = m.geometry(pts)
    out = torch.empty(1, 1024, 1024, 256, dtype=torch.float16, device='cuda')
    for _ in range(3):
        m.embed(geom, torch.float16, out=out)
torch.cuda.synchronize()
print('ok', float(out.float().abs().max()))
