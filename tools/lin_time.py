"""Stand-alone timing of K13 shapes of the forward (CUDA events around graph replays of 20 back-to-back launches)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pkgpath
pkg = pkgpath.load_package()
dev = torch.device('cuda')
for (T, K, N) in ((1024, 1024, 256), (2048, 256, 2112), (2048, 256, 768), (1024, 256, 512), (2048, 256, 256), (2048, 2048, 128)):
    x = torch.randn(T, K, device=dev); w = torch.randn(N, K, device=dev) / K ** 0.5; b = torch.randn(N, device=dev)
    wh = pkg.ops.pack_linear(w); y = torch.empty(T, N, device=dev)
    for _ in range(3): pkg.ops.linear_f16(x, wh, b, out=y)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):  # the host cannot issue eager launches fast enough: 20 launches per graph replay
        for _ in range(20): pkg.ops.linear_f16(x, wh, b, out=y)
    g.replay(); torch.cuda.synchronize()
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5): g.replay()
    e.record(); torch.cuda.synchronize()
    print(f'T={T} K={K} N={N}: {a.elapsed_time(e) / 100 * 1e3:.1f} us')
