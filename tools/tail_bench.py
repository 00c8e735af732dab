"""Times the stages after the transformer chained on the GPU (tests/test_tail_chain.py's glue) at the 3DMatch size:
1024 nodes, 16k fine points, 256 node correspondences, 64 points per patch.  Eager, CUDA events, us per pair."""
import json
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'tests'))
import torch
import pkgpath
import test_tail_chain as tc

pkg = pkgpath.load_package()
tc.NUM_CORR, tc.POINT_LIMIT = 256, 64
d = {k: v.cuda() for k, v in pkg.synth.make_tail_inputs(num_nodes=1024, points_per_node=16, seed=21).items()}
stages = (pkg.point_to_node_partition, pkg.SuperPointMatching(256, True), pkg.LearnableLogOptimalTransport(100).cuda(),
          pkg.LocalGlobalRegistration(3, 0.1), pkg.patch_matching_scores)
out = tc.run_tail(d, *stages)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 20
a.record()
for _ in range(n):
    out = tc.run_tail(d, *stages)
b.record()
torch.cuda.synchronize()
print(json.dumps({'us_per_pair': a.elapsed_time(b) * 1e3 / n, 'correspondences': int(out[0].shape[0]),
                  'transform_error': float((out[3] - d['transform']).abs().max()),
                  'nodes': 1024, 'fine_points': [int(d['ref_points_f'].shape[0]), int(d['src_points_f'].shape[0])],
                  'timing': f'eager, {n} pairs, CUDA events; two host reads per pair (coarse-match count under masks, '
                            'correspondence count)'}))
