"""Golden vectors for SURVEY.md 8(f)-2 (point_to_node_partition) and 8(f)-3 (LearnableLogOptimalTransport), written by
running the UNMODIFIED reference functions on CPU.  Runs only in the build container:

    python tests/golden/make_golden_next.py

Both reference functions create their helper tensors with `.cuda()` (pointcloud_partition.py:85,89,95;
learnable_sinkhorn.py:35-62), which cannot run without a GPU.  For the duration of the calls `torch.Tensor.cuda` is
replaced by the identity, so the reference's own code runs unchanged on CPU tensors: same arithmetic, same torch ops.

Partition cases (inputs stored; every case is checked to be tie-free, oracle.partition_is_tie_free):
  p2n_small   400 points, 37 nodes, limit 16   -- nodes with more and with fewer points than the limit
  p2n_empty   300 points, 60 nodes (20 of them far away: empty nodes), limit 8
  p2n_all     50 points, 3 nodes, limit 50     -- limit == number of points
  p2n_big     3000 points, 4 nodes, limit 100  -- nodes far larger than the sort buffer (chunked merge in the kernel)
  p2n_3dmatch 1024 nodes x 16 points, limit 64 -- BASELINE-size 3DMatch shape (model.py:98-103, config.py:93)
  p2n_kitti   512 nodes x 24 points, limit 128 -- KITTI shape
Transport cases:
  lot_small   (6, 9, 7) scores, masks, 20 iterations; lot_nomask (4, 12, 12), 100 iterations
  lot_3dmatch (32, 64, 64), masks, 100 iterations (config.py:94); lot_kitti (8, 128, 128), masks, 100 iterations
  lot_alpha   (5, 10, 14) with alpha = -0.37
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
from load_reference import load_reference  # noqa: E402
import geotx_oracle as oracle  # noqa: E402
import pkgpath  # noqa: E402


class cuda_is_identity:
    def __enter__(self):
        self.saved = torch.Tensor.cuda
        torch.Tensor.cuda = lambda self, *a, **k: self
        return self

    def __exit__(self, *exc):
        torch.Tensor.cuda = self.saved
        return False


def main():
    load_reference()
    from geotransformer.modules.ops.pointcloud_partition import point_to_node_partition
    from geotransformer.modules.sinkhorn.learnable_sinkhorn import LearnableLogOptimalTransport
    synth = pkgpath.load_package().synth
    torch.set_grad_enabled(False)
    out = dict(meta=repr(dict(torch=torch.__version__)))

    def partition_case(tag, points, nodes, K):
        assert oracle.partition_is_tie_free(points, nodes), tag
        with cuda_is_identity():
            p2n, sizes, masks, knn_idx, knn_masks = point_to_node_partition(points, nodes, K, return_count=True)
        out.update({f'{tag}.points': points.numpy(), f'{tag}.nodes': nodes.numpy(), f'{tag}.K': np.int64(K),
                    f'{tag}.point_to_node': p2n.numpy().astype(np.int32), f'{tag}.node_sizes': sizes.numpy().astype(np.int32),
                    f'{tag}.node_masks': masks.numpy(), f'{tag}.node_knn_indices': knn_idx.numpy().astype(np.int32),
                    f'{tag}.node_knn_masks': knn_masks.numpy()})
        print(tag, tuple(points.shape), tuple(nodes.shape), K, 'sizes', int(sizes.min()), int(sizes.max()),
              'empty nodes', int((~masks).sum()))

    g = torch.Generator().manual_seed(11)
    nodes = torch.rand(37, 3, generator=g) * 2
    partition_case('p2n_small', synth.make_fine_cloud('3dmatch', nodes, 11, g)[:400].contiguous(), nodes, 16)
    nodes = torch.cat([torch.rand(40, 3, generator=g), torch.rand(20, 3, generator=g) + 50.0])
    partition_case('p2n_empty', torch.rand(300, 3, generator=g), nodes, 8)
    partition_case('p2n_all', torch.rand(50, 3, generator=g), torch.rand(3, 3, generator=g), 50)
    for t in range(64):
        pts, nds = torch.rand(3000, 3, generator=g), torch.rand(4, 3, generator=g)
        if oracle.partition_is_tie_free(pts, nds):
            break
    partition_case('p2n_big', pts, nds, 100)
    for t in range(16):
        pts, nds = synth.make_partition_inputs('3dmatch', 1024, 16, synth.SEED + 1000 * t)
        if oracle.partition_is_tie_free(pts, nds):
            break
    partition_case('p2n_3dmatch', pts, nds, 64)
    for t in range(16):
        pts, nds = synth.make_partition_inputs('kitti', 512, 24, synth.SEED + 1000 * t)
        if oracle.partition_is_tie_free(pts, nds):
            break
    partition_case('p2n_kitti', pts, nds, 128)

    def transport_case(tag, scores, rm, cm, iters, alpha=1.0):
        mod = LearnableLogOptimalTransport(iters)
        mod.alpha.fill_(alpha)
        with cuda_is_identity():
            res = mod(scores, rm, cm)
        out.update({f'{tag}.scores': scores.numpy(), f'{tag}.iters': np.int64(iters), f'{tag}.alpha': np.float32(alpha),
                    f'{tag}.out': res.numpy()})
        if rm is not None:
            out.update({f'{tag}.row_masks': rm.numpy(), f'{tag}.col_masks': cm.numpy()})
        print(tag, tuple(res.shape), float(res[res > -1e11].min()), float(res.max()))

    transport_case('lot_small', *synth.make_patch_scores(6, 9, 7, 32, seed=21), 20)
    transport_case('lot_nomask', synth.make_patch_scores(4, 12, 12, 32, seed=22)[0], None, None, 100)
    transport_case('lot_3dmatch', *synth.make_patch_scores(32, 64, 64, 256, seed=23), 100)
    transport_case('lot_kitti', *synth.make_patch_scores(8, 128, 128, 256, seed=24), 100)
    transport_case('lot_alpha', *synth.make_patch_scores(5, 10, 14, 16, seed=25), 50, alpha=-0.37)
    np.savez_compressed(os.path.join(HERE, 'next_cases.npz'), **out)
    print('bytes', os.path.getsize(os.path.join(HERE, 'next_cases.npz')))


if __name__ == '__main__':
    main()
