"""GPU timing of the fused embedding kernel alone (geotx_gse_embed through the C ABI).
usage: k2_bench.py [lib.so] -- optional alternative library build (timing experiments, tools/k2_variants.sh)."""
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
DT = {'f16': torch.float16, 'bf16': torch.bfloat16, 'f32': torch.float32}[os.environ.get('K2_DTYPE', 'f16')]
CASES = ((256, 1024, 0.2), (256, 2048, 0.2), (128, 2048, 4.8))
if os.environ.get('K2_CASE'):
    CASES = (CASES[int(os.environ['K2_CASE'])],)
for C, n, sd in CASES:
    torch.manual_seed(1)
    m = pkg.GeometricStructureEmbedding(C, sd, 15, 3).to(dev).eval()
    pts = (torch.rand(1, n, 3) * (3.0 if C == 256 else 100.0)).to(dev)
    with torch.no_grad():
        geom = m.geometry(pts)
        nbuf = int(os.environ.get('K2_BUFFERS', '1'))  # >1: rotate output buffers (as the two clouds of a pair do)
        outs = [torch.empty(1, n, n, C, dtype=DT, device=dev) for _ in range(nbuf)]
        out = outs[0]
        for i in range(3):
            m.embed(geom, DT, out=outs[i % nbuf])
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        reps = 10
        ev[0].record()
        for i in range(reps):
            m.embed(geom, DT, out=outs[i % nbuf])
        ev[1].record()
        torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / reps
    fl = 2.0 * n * n * C * C * 4
    print(f'K2 C={C} n={n}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s  cycles/tile/SM@1.965GHz '
          f'{ms * 1e-3 * 1.965e9 / (n * n / 128 / 148):.0f}  watchdog {pkg._abi.watchdog()}', flush=True)
