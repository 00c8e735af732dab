import os, sys, torch
sys.path.insert(0, '/root/repo')
import pkgpath
pkg = pkgpath.load_package()
for n in (1024, 4096):
    torch.manual_seed(n)
    ref, src = torch.randn(n, 256, device='cuda'), torch.randn(n, 256, device='cuda')
    for _ in range(3):
        pkg.ops.superpoint_matching(ref, src, 256, True, normalize=True)
    torch.cuda.synchronize()
