"""Do two in-flight forwards (two CUDA graphs on two streams) overlap on one GPU?  Pairs/s with 1, 2, 3 graphs."""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
import pkgpath

pkg = pkgpath.load_package()
cfg = pkg.synth.CONFIGS['3dmatch']
torch.manual_seed(0)
model = pkg.GeometricTransformer(**cfg).cuda().eval()
pairs = [tuple(t.cuda() for t in pkg.synth.make_pair('3dmatch', 1024, seed=7351 + p)[:4]) for p in range(4)]
for depth in (1, 2, 3):
    graphs = [pkg.GraphedGeometricTransformer(model, *pairs[0]) for _ in range(depth)]
    streams = [torch.cuda.Stream() for _ in range(depth)]
    def run(steps):
        for s in range(steps):
            with torch.cuda.stream(streams[s % depth]):
                graphs[s % depth].graph.replay()
    run(6)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for st in streams:
        st.wait_stream(torch.cuda.current_stream())
    run(60)
    for st in streams:
        torch.cuda.current_stream().wait_stream(st)
    b.record()
    torch.cuda.synchronize()
    print(f'depth {depth}: {60 / (a.elapsed_time(b) * 1e-3):.1f} pairs/s', flush=True)
    del graphs
