"""Stand-alone timing of the attention kernels (CUDA events around graph replays of 20 back-to-back launches: the host
cannot issue eager launches fast enough).  python tools/att_time.py [n ...]   (GEOTX_ATT_GROUPS / _MAX_SPLITS / _BIG apply)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
dev = torch.device('cuda', 0)
H, c = 4, 64


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20):
            fn()
    g.replay()
    torch.cuda.synchronize()
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        g.replay()
    e.record()
    torch.cuda.synchronize()
    return a.elapsed_time(e) / 100 * 1e3


for n in [int(a) for a in sys.argv[1:]] or [1024]:
    gen = torch.Generator().manual_seed(1)
    y = torch.randn(1, n, 3 * H * c, generator=gen).to(dev)
    yh = y.to(torch.float16)
    sp = (torch.randn(1, H, n, n, generator=gen) * 3).to(dev)
    out = torch.empty(1, n, H * c, device=dev)
    for name, t in (('half (K3h)', yh), ('fp32 (K3b)', y)):
        q, k, v = t[:, :, :H * c], t[:, :, H * c:2 * H * c], t[:, :, 2 * H * c:]
        row = []
        for split in (False, True):
            for s in (sp, None):
                row.append(timed(lambda: pkg.ops.attention(q, k, v, H, s, split_keys=split, out=out)))
        print(f'n={n} {name}: no split: {row[0]:.1f} us with scores, {row[1]:.1f} without;  cluster split: {row[2]:.1f} / {row[3]:.1f}',
              flush=True)
