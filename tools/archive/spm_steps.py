"""Cumulative in-graph time of the matching call's kernels (debug library built with -DSPM_DBG_STOP)."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgpath
pkg = pkgpath.load_package()
pkg._abi.LIB_PATH = os.path.abspath(sys.argv[1])
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
torch.manual_seed(n)
ref, src = torch.randn(n, 256, device='cuda'), torch.randn(n, 256, device='cuda')
names = ['prepare', 'scores', 'sums', 'pass0', 'pass1', 'pass2', 'pass3', 'pass4', 'pass5', 'collect']
prev = 0.0
for k in range(1, 11):
    os.environ['GEOTX_SPM_STOP_AFTER'] = str(k)
    fn = lambda: pkg.ops.superpoint_matching(ref, src, 256, True, normalize=True)
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(10):
            keep = fn()
    g.replay(); torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(5):
        g.replay()
    ev[1].record(); torch.cuda.synchronize()
    us = ev[0].elapsed_time(ev[1]) * 1e3 / 50
    print(f'{names[k-1]:8s} cumulative {us:6.1f} us  (+{us - prev:5.1f})')
    prev = us
