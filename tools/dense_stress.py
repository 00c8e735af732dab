"""Race hunt for K13 / K4f: the same call many times, every result compared bit for bit with the first."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pkgpath
pkg = pkgpath.load_package()
dev = torch.device('cuda')
torch.manual_seed(0)
for (T, K, N) in ((2048, 256, 1792), (2048, 256, 2112), (4096, 256, 2112), (1024, 1024, 256), (2048, 256, 768), (4096, 128, 896), (8192, 256, 256)):
    x = torch.randn(T, K, device=dev)
    w = torch.randn(N, K, device=dev) / K ** 0.5
    b = torch.randn(N, device=dev)
    wh = pkg.ops.pack_linear(w)
    ref = torch.nn.functional.linear(x, w, b)
    first = pkg.ops.linear_f16(x, wh, b).clone()
    print(T, K, N, 'rel err vs torch', float((first - ref).norm() / ref.norm()))
    bad = 0
    for it in range(300):
        y = pkg.ops.linear_f16(x, wh, b)
        if not torch.equal(y, first):
            bad += 1
            d = (y != first)
            rows = d.any(1).nonzero().flatten()
            cols = d.any(0).nonzero().flatten()
            if bad <= 5:
                print('  iter', it, 'mismatch: rows', rows[:4].tolist(), '..', rows[-1].item(), 'n', rows.numel(), '| cols', cols[:4].tolist(), '..',
                      cols[-1].item(), 'n', cols.numel(), '| max diff', float((y - first).abs().max()))
    print('  mismatching iterations:', bad, 'of 300')
C = 256
g = torch.Generator().manual_seed(1)
hidden = torch.randn(1, 16384, C, generator=g).to(dev); x_in = torch.randn(1, 16384, C, generator=g).to(dev)
wl = pkg.ops.pack_linear((torch.randn(C, C, generator=g) / 16).to(dev)); we = pkg.ops.pack_linear((torch.randn(2 * C, C, generator=g) / 16).to(dev))
ws = pkg.ops.pack_linear((torch.randn(C, 2 * C, generator=g) / 22).to(dev))
vec = lambda n: torch.randn(n, generator=g).to(dev)
args = (wl, vec(C), vec(C), vec(C), 1e-5, we, vec(2 * C), ws, vec(C), vec(C), vec(C), 1e-5)
first = pkg.ops.post_attention(hidden, x_in, *args).clone()
bad = 0
for it in range(300):
    y = pkg.ops.post_attention(hidden, x_in, *args)
    if not torch.equal(y, first):
        bad += 1
        d = (y != first)[0]
        rows = d.any(1).nonzero().flatten()
        if bad <= 5:
            print('  K4f iter', it, 'rows', rows[:4].tolist(), 'n', rows.numel(), 'max diff', float((y - first).abs().max()))
print('K4f mismatching iterations:', bad, 'of 300')
