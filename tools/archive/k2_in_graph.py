"""Duration of the embedding kernel INSIDE the CUDA-graph replay of a forward, by difference: replay time of the full
forward minus replay time of the same forward with the embedding kernel's launches removed (E left as it was)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
dev = 'cuda:0'
shape, n = os.environ.get('SHAPE', '3dmatch'), int(os.environ.get('N', '1024'))
cfg = dict(pkg.synth.CONFIGS[shape])
torch.manual_seed(pkg.synth.SEED)
model = pkg.GeometricTransformer(**cfg).to(dev).eval()
pair = [t.to(dev) for t in pkg.synth.make_pair(shape, n, n, seed=pkg.synth.SEED)]


def replay_ms(fast, reps=30):
    for _ in range(5):
        fast(*pair)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fast(*pair)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


full = replay_ms(pkg.GraphedGeometricTransformer(model, *pair))
orig = model.embedding.embed
cache = {}


def no_kernel(geom, out_dtype, row_begin=0, row_count=None, out=None, qt=None):
    key = (geom.n, out_dtype)
    if key not in cache:
        cache[key] = torch.zeros(geom.B, geom.n, geom.n, model.embedding.hidden_dim, dtype=out_dtype, device=dev)
    return cache[key]


with torch.no_grad():
    no_kernel(model.embedding.geometry(pair[0]), model.embedding.storage_dtype())
model.embedding.embed = no_kernel
without = replay_ms(pkg.GraphedGeometricTransformer(model, *pair))
model.embedding.embed = orig
fl = 2.0 * n * n * cfg['hidden_dim'] ** 2 * 4
k2 = (full - without) / 2
print(f'{shape} n={n}: forward {full:.4f} ms, without the embedding kernel {without:.4f} ms -> {k2:.4f} ms per launch '
      f'in the graph = {fl / k2 / 1e9:.0f} TFLOP/s')
