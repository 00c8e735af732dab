"""GPU debug aid: run the tcgen05 embedding kernel on a small cloud, report the protocol watchdog and the error
against the fp32 CUDA-core kernel (both through the C ABI)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
C = int(sys.argv[1]) if len(sys.argv) > 1 else 256
n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
red = sys.argv[3] if len(sys.argv) > 3 else 'max'
dev = 'cuda:0'
torch.manual_seed(1)
m = pkg.GeometricStructureEmbedding(C, 0.2, 15, 3, reduction_a=red).to(dev).eval()
pts = (torch.rand(1, n, 3) * 2).to(dev)
with torch.no_grad():
    geom = m.geometry(pts)
    ref = pkg.ops.gse_embed_fp32(geom, m.embedding.div_term, m.proj_d.weight, m.proj_d.bias, m.proj_a.weight,
                                 m.proj_a.bias, m.sigma_d, m.factor_a, red == 'mean')
    torch.cuda.synchronize()
    print('fp32 kernel ok', float(ref.abs().mean()))
    try:
        e = m.embed(geom, torch.float32)
        torch.cuda.synchronize()
    except Exception as ex:  # noqa: BLE001
        print('tc kernel raised:', repr(ex)[:300])
        sys.exit(3)
    print('watchdog:', pkg._abi.watchdog())
    err = ((e - ref).norm() / ref.norm()).item()
    print(f'C={C} n={n} {red}: tc vs fp32 rel err {err:.3e}  max|diff| {float((e - ref).abs().max()):.3e}')
    if err > 1e-2:
        d = (e - ref)[0].abs()
        print('per-row-of-tile err (first 16 pairs):', d.view(-1, C)[:16].mean(1).tolist())
        print('per-column err (first 16 cols):', d.view(-1, C).mean(0)[:16].tolist())
        print('tc sample:', e.view(-1, C)[1, :8].tolist())
        print('ref sample:', ref.view(-1, C)[1, :8].tolist())
    sys.exit(0 if err <= 1e-2 else 1)
