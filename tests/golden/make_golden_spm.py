"""Golden vectors for the coarse matching step, written by running the UNMODIFIED reference
`SuperPointMatching` (geotransformer/modules/geotransformer/superpoint_matching.py) and `F.normalize` on CPU.
Runs only in the build container:

    python tests/golden/make_golden_spm.py

Cases
  * small_*: seeded random features (stored), with and without masks / dual normalisation, K larger than the matrix;
  * 3dmatch_1024: inputs are the committed reference outputs of GeometricTransformer at BASELINE configs[1]
    (tests/golden/3dmatch_n1024_out.npz), i.e. exactly what experiments/*/model.py:141-157 feeds the matcher.
The reference creates its default masks with `.cuda()` (superpoint_matching.py:27-30), which cannot run here, so
all-True masks are passed explicitly: same arithmetic.
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
from load_reference import load_reference  # noqa: E402


def correlated_features(n, m, d, seed, noise=0.35):
    """src rows = a permutation of the first m ref rows plus noise: a matching problem with structure."""
    g = torch.Generator().manual_seed(seed)
    ref = torch.randn(n, d, generator=g)
    perm = torch.randperm(n, generator=g)[:m]
    src = ref[perm] + noise * torch.randn(m, d, generator=g)
    return ref, src


def main():
    load_reference()
    from geotransformer.modules.geotransformer.superpoint_matching import SuperPointMatching
    torch.set_grad_enabled(False)
    out = dict(meta=repr(dict(torch=torch.__version__)))
    cases = [
        # tag, n, m, d, K, dual, masked, seed
        ('small_dual', 96, 80, 32, 64, True, False, 1),
        ('small_plain', 96, 80, 32, 64, False, False, 2),
        ('small_masked', 120, 100, 48, 96, True, True, 3),
        ('small_allpairs', 9, 7, 16, 256, True, False, 4),
        ('mid_dual', 717, 650, 256, 256, True, True, 5),
    ]
    for tag, n, m, d, K, dual, masked, seed in cases:
        ref, src = correlated_features(n, m, d, seed)
        g = torch.Generator().manual_seed(100 + seed)
        rm = torch.rand(n, generator=g) > 0.15 if masked else torch.ones(n, dtype=torch.bool)
        sm = torch.rand(m, generator=g) > 0.15 if masked else torch.ones(m, dtype=torch.bool)
        rn, sn = F.normalize(ref, p=2, dim=1), F.normalize(src, p=2, dim=1)  # model.py:141-142
        ri, si, sc = SuperPointMatching(K, dual)(rn, sn, rm, sm)
        out.update({f'{tag}.ref_feats': ref.numpy(), f'{tag}.src_feats': src.numpy(), f'{tag}.ref_masks': rm.numpy(),
                    f'{tag}.src_masks': sm.numpy(), f'{tag}.ref_idx': ri.numpy(), f'{tag}.src_idx': si.numpy(),
                    f'{tag}.scores': sc.numpy(), f'{tag}.cfg': repr(dict(K=K, dual=dual))})
        print(tag, len(sc))
    z = np.load(os.path.join(HERE, '3dmatch_n1024_out.npz'))
    ref, src = torch.from_numpy(z['out_ref'][0]), torch.from_numpy(z['out_src'][0])
    rn, sn = F.normalize(ref, p=2, dim=1), F.normalize(src, p=2, dim=1)
    ones = torch.ones(ref.shape[0], dtype=torch.bool)
    ri, si, sc = SuperPointMatching(256, True)(rn, sn, ones, torch.ones(src.shape[0], dtype=torch.bool))
    out.update({'3dmatch_1024.ref_idx': ri.numpy(), '3dmatch_1024.src_idx': si.numpy(), '3dmatch_1024.scores': sc.numpy(),
                '3dmatch_1024.cfg': repr(dict(K=256, dual=True, inputs='3dmatch_n1024_out.npz'))})
    print('3dmatch_1024', len(sc))
    np.savez_compressed(os.path.join(HERE, 'spm_cases.npz'), **out)


if __name__ == '__main__':
    main()
