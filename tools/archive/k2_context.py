"""Why is the embedding kernel ~8 % slower inside a forward than alone?  Times geotx_gse_embed (n = 1024, C = 256)
alone, with rotating point sets, and with other kernels of the forward interleaved between the timed launches."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
dev = 'cuda:0'
C, n, H = 256, 1024, 4
torch.manual_seed(1)
m = pkg.GeometricStructureEmbedding(C, 0.2, 15, 3).to(dev).eval()
with torch.no_grad():
    geoms = [m.geometry((torch.rand(1, n, 3) * 3.0).to(dev)) for _ in range(4)]
    syn = pkg.synth.make_pair('3dmatch', n, n, seed=7351)
    geoms_syn = [m.geometry(syn[0].to(dev)), m.geometry(syn[1].to(dev))]
    outs = [torch.empty(1, n, n, C, dtype=torch.float16, device=dev) for _ in range(2)]
    qt = torch.randn(1, H, n, C, device=dev)
    sp = torch.empty(1, H, n, n, device=dev)
    big = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def run(label, geom_list, between):
    with torch.no_grad():
        for i in range(3):
            m.embed(geom_list[i % len(geom_list)], torch.float16, out=outs[i % 2])
        torch.cuda.synchronize()
        evs = []
        torch.cuda._sleep(int(4e6))
        for i in range(10):
            between(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            m.embed(geom_list[i % len(geom_list)], torch.float16, out=outs[i % 2])
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in evs)
    print(f'{label:46s} median {ms[len(ms) // 2]:.4f} ms  min {ms[0]:.4f}  max {ms[-1]:.4f}', flush=True)


run('alone, one point set', geoms[:1], lambda i: None)
run('alone, 4 random point sets', geoms, lambda i: None)
run('alone, synthetic 3DMatch pair (bench inputs)', geoms_syn, lambda i: None)
run('after rpe_scores (reads the other E)', geoms_syn, lambda i: pkg.ops.rpe_scores(qt, outs[(i + 1) % 2], out=sp))
run('after a 256 MB memset', geoms_syn, lambda i: big.zero_())
run('after a 1 ms device sleep', geoms_syn, lambda i: torch.cuda._sleep(int(2e6)))
