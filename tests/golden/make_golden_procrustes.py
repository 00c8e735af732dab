"""Golden vectors for SURVEY.md 8(f)-4 (weighted_procrustes), written by running the UNMODIFIED reference function
(geotransformer/modules/registration/procrustes.py:6-73) on CPU.  Runs only in the build container:

    python tests/golden/make_golden_procrustes.py

The reference hops H.cpu() -> torch.svd -> .cuda() (procrustes.py:53-56); `torch.Tensor.cuda` is the identity for the
duration of the call, so its own code runs unchanged on CPU tensors.  Inputs are regenerated from the seeds by
synth.make_procrustes_inputs (stored too for the small cases).

Cases: local (64 problems x 64 points, weights, zero-padded tails: what local_global_registration.py:169 feeds),
thresh (weight_thresh = 0.3), noweights (weights=None), planar (rank-2 covariances), reflect (mirrored targets: the
determinant correction picks the proper rotation), global (1 problem x 5000 points: the refinement steps :181-191).
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
import pkgpath  # noqa: E402

CASES = {
    # tag: (make_procrustes_inputs kwargs, weight_thresh, use weights)
    'local': (dict(batch=64, n=64, seed=31), 0.0, True),
    'thresh': (dict(batch=32, n=48, seed=32), 0.3, True),
    'noweights': (dict(batch=16, n=40, seed=33, outliers=0.0), 0.0, False),
    'planar': (dict(batch=16, n=64, seed=34, planar=True, outliers=0.0), 0.0, True),
    'reflect': (dict(batch=16, n=64, seed=35, reflect=True, outliers=0.0), 0.0, True),
    'global': (dict(batch=1, n=5000, seed=36), 0.0, True),
}


def main():
    load_reference()
    from geotransformer.modules.registration.procrustes import weighted_procrustes
    synth = pkgpath.load_package().synth
    torch.set_grad_enabled(False)
    out = dict(meta=repr(dict(torch=torch.__version__)))
    saved = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        for tag, (kw, thresh, use_w) in CASES.items():
            src, ref, w, _ = synth.make_procrustes_inputs(**kw)
            T = weighted_procrustes(src, ref, w if use_w else None, weight_thresh=thresh, return_transform=True)
            R, t = weighted_procrustes(src, ref, w if use_w else None, weight_thresh=thresh)
            assert torch.equal(T[:, :3, :3], R) and torch.equal(T[:, :3, 3], t)
            out[f'{tag}.transform'] = T.numpy()
            print(tag, tuple(T.shape), float(torch.det(R).min()), float(torch.det(R).max()))
    finally:
        torch.Tensor.cuda = saved
    np.savez_compressed(os.path.join(HERE, 'procrustes_cases.npz'), **out)


if __name__ == '__main__':
    main()
