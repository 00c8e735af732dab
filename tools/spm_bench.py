"""GPU timing of the coarse matching call (geotx_superpoint_matching through the C ABI: F.normalize + scores + dual
normalisation + global top-K), captured into a CUDA graph and replayed, beside the reference's formulation written
with stock torch ops on the same GPU (fp32, allow_tf32 off; information only, not the baseline of record).
usage: spm_bench.py [lib.so]"""
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
if len(sys.argv) > 1:
    pkg._abi.LIB_PATH = os.path.abspath(sys.argv[1])
dev = 'cuda:0'
CALLS = 10
torch.backends.cuda.matmul.allow_tf32 = False


def stock(ref, src, K):
    rn, sn = F.normalize(ref, p=2, dim=1), F.normalize(src, p=2, dim=1)
    s = torch.exp(-(2.0 - 2.0 * torch.matmul(rn, sn.transpose(-1, -2))).clamp(min=0.0))
    s = (s / s.sum(dim=1, keepdim=True)) * (s / s.sum(dim=0, keepdim=True))
    v, i = s.view(-1).topk(k=K, largest=True)
    return torch.div(i, s.shape[1], rounding_mode='floor'), i % s.shape[1], v


def graph_time(fn):
    with torch.no_grad():
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(CALLS):
                keep = fn()
        g.replay()
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        for _ in range(5):
            g.replay()
        ev[1].record()
        torch.cuda.synchronize()
    del keep
    return ev[0].elapsed_time(ev[1]) * 1e3 / (5 * CALLS)


rows = []
for n, d in ((512, 256), (717, 256), (1024, 256), (2048, 256), (4096, 256)):
    torch.manual_seed(n)
    ref, src = torch.randn(n, d, device=dev), torch.randn(n, d, device=dev)
    ours = graph_time(lambda: pkg.ops.superpoint_matching(ref, src, 256, True, normalize=True))
    theirs = graph_time(lambda: stock(ref, src, 256))
    row = dict(n=n, d=d, K=256, geotx_us=round(ours, 1), torch_eager_graph_us=round(theirs, 1))
    rows.append(row)
    print(json.dumps(row), flush=True)
