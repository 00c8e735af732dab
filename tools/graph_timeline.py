"""Per-launch timeline of one forward INSIDE its CUDA graph: every kernel-launching call of the forward is bracketed by a
pair of external CUDA events captured into the graph; after a few replays the per-call durations and the gaps between
consecutive calls are printed.  (The event-record nodes add ~1 us of serialisation each: totals are slightly high.)
usage: SHAPE=3dmatch N=1024 python tools/graph_timeline.py"""
import collections
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
modules = sys.modules[pkg.__name__ + '.modules']
dev = 'cuda:0'
shape, n = os.environ.get('SHAPE', '3dmatch'), int(os.environ.get('N', '1024'))
cfg = dict(pkg.synth.CONFIGS[shape])
torch.manual_seed(pkg.synth.SEED)
model = pkg.GeometricTransformer(**cfg).to(dev).eval()
model.transformer.overlap_attention = False  # one stream: a linear timeline
pair = [t.to(dev) for t in pkg.synth.make_pair(shape, n, n, seed=pkg.synth.SEED)]
events = []


def wrap(label, fn):
    def inner(*a, **kw):
        e0 = torch.cuda.Event(enable_timing=True, external=True)
        e1 = torch.cuda.Event(enable_timing=True, external=True)
        e0.record()
        out = fn(*a, **kw)
        e1.record()
        events.append((label, e0, e1))
        return out
    return inner


for name in ('gse_knn', 'gse_embed', 'rpe_scores', 'attention', 'add_layernorm'):
    setattr(pkg.ops, name, wrap(name, getattr(pkg.ops, name)))
real_linear, real_addmm, real_addmm_act = F.linear, torch.addmm, torch._addmm_activation
modules.F.linear = wrap('linear (cuBLAS)', real_linear)
torch.addmm = wrap('addmm (cuBLAS)', real_addmm)
torch._addmm_activation = wrap('addmm+relu (cuBLAS)', real_addmm_act)

fast = pkg.GraphedGeometricTransformer(model, *pair)
F.linear, torch.addmm, torch._addmm_activation = real_linear, real_addmm, real_addmm_act
per_forward = [e for e in events]  # includes the warm-up forwards; keep the captured one = the last third
k = len(per_forward) // 3
captured = per_forward[-k:]
tot = collections.defaultdict(lambda: [0, 0.0])
gaps, wall = 0.0, 0.0
reps = 10
for r in range(3 + reps):
    fast(*pair)
    torch.cuda.synchronize()
    if r < 3:
        continue
    for i, (label, e0, e1) in enumerate(captured):
        d = e0.elapsed_time(e1)
        tot[label][0] += 1
        tot[label][1] += d
    wall += captured[0][1].elapsed_time(captured[-1][2])
    for i in range(len(captured) - 1):
        gaps += max(captured[i][2].elapsed_time(captured[i + 1][1]), 0.0)
print(f'{shape} n={n}: {len(captured)} timed calls per forward, first start -> last end {wall / reps * 1e3:.1f} us, '
      f'sum of gaps between calls {gaps / reps * 1e3:.1f} us')
for label, (c, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f'  {label:24s} {c / reps:5.1f} calls  {t / reps * 1e3:8.1f} us/forward  {t / c * 1e3:7.2f} us each')
