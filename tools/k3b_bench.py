"""GPU timing of the attention kernel alone (geotx_attention through the C ABI), host overhead removed: 20 calls are
captured into one CUDA graph and the graph is replayed.  usage: k3b_bench.py [lib.so]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
if len(sys.argv) > 1:
    pkg._abi.LIB_PATH = os.path.abspath(sys.argv[1])
dev = 'cuda:0'
CALLS = 20
for C, H in ((256, 4), (128, 4)):
    for n in (512, 1024, 2048, 4096):
        torch.manual_seed(n)
        y = torch.randn(1, n, 3 * C, device=dev)
        q, k, v = y[:, :, :C], y[:, :, C:2 * C], y[:, :, 2 * C:]
        sp = torch.randn(1, H, n, n, device=dev)
        out = torch.empty(1, n, C, device=dev)
        line = f'K3b C={C} n={n}:'
        for with_p in (True, False):
            for split in (True, False):
                args = dict(scores_p=sp if with_p else None, split_keys=split, out=out)
                with torch.no_grad():
                    side = torch.cuda.Stream()
                    side.wait_stream(torch.cuda.current_stream())
                    with torch.cuda.stream(side):
                        pkg.ops.attention(q, k, v, H, **args)
                    torch.cuda.current_stream().wait_stream(side)
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        for _ in range(CALLS):
                            pkg.ops.attention(q, k, v, H, **args)
                    g.replay()
                    torch.cuda.synchronize()
                    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
                    ev[0].record()
                    for _ in range(5):
                        g.replay()
                    ev[1].record()
                    torch.cuda.synchronize()
                us = ev[0].elapsed_time(ev[1]) * 1e3 / (5 * CALLS)
                line += f"  {'sp' if with_p else 'no-sp'}/{'split' if split else 'whole'} {us:6.1f} us"
        print(line, flush=True)
