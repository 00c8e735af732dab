"""Two launches of the attention kernel at n = m = 1024, H = 4, C = 256: with the additive scores (self block) and
without (cross block); for ncu captures."""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
import pkgpath

pkg = pkgpath.load_package()
n, C, H = 1024, 256, 4
torch.manual_seed(1)
y = torch.randn(1, n, 3 * C, device='cuda')
q, k, v = y[:, :, :C], y[:, :, C:2 * C], y[:, :, 2 * C:]
sp = torch.randn(1, H, n, n, device='cuda')
out = torch.empty(1, n, C, device='cuda')
for _ in range(2):
    pkg.ops.attention(q, k, v, H, scores_p=sp, out=out)
    pkg.ops.attention(q, k, v, H, scores_p=None, out=out)
torch.cuda.synchronize()
print('ok')
