"""A few stand-alone launches of the half-operand attention kernel (K3h) for ncu: python tools/att_once.py [n] [split 0|1]
launch 0, 2, ..: with additive scores (a 'self' block), launch 1, 3, ..: without (a 'cross' block)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
split = bool(int(sys.argv[2])) if len(sys.argv) > 2 else False
H, c = 4, 64
dev = torch.device('cuda', 0)
g = torch.Generator().manual_seed(1)
y = torch.randn(1, n, 3 * H * c, generator=g).to(dev).to(torch.float16)
sp = (torch.randn(1, H, n, n, generator=g) * 3).to(dev)
q, k, v = y[:, :, :H * c], y[:, :, H * c:2 * H * c], y[:, :, 2 * H * c:]
for it in range(3):
    a = pkg.ops.attention(q, k, v, H, sp, split_keys=split)
    b = pkg.ops.attention(q, k, v, H, None, split_keys=split)
torch.cuda.synchronize()
t0, t1, t2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
t0.record()
for it in range(20):
    pkg.ops.attention(q, k, v, H, sp, split_keys=split)
t1.record()
for it in range(20):
    pkg.ops.attention(q, k, v, H, None, split_keys=split)
t2.record()
torch.cuda.synchronize()
print(f'n={n} split={split}: with scores {t0.elapsed_time(t1) / 20 * 1e3:.1f} us, without {t1.elapsed_time(t2) / 20 * 1e3:.1f} us '
      f'(back-to-back launches, eager)')
