"""Golden vectors for LocalGlobalRegistration (the consumer of the transport and Procrustes rows), written by running
the UNMODIFIED reference module (geotransformer/modules/geotransformer/local_global_registration.py) on CPU with
`torch.Tensor.cuda` made the identity for the duration of the call (the module creates its index helpers with
`.cuda()` and hops to the host for the SVD).  Runs only in the build container:

    python tests/golden/make_golden_lgr.py

Inputs are regenerated from the seeds by synth.make_lgr_inputs.  Cases: 3dmatch (128 patch pairs x 64 points, top-3,
radius 0.1: config.py:116-125), kitti (48 x 128 points, top-2), small (6 x 12), nonmutual, global_score, degenerate
(no patch pair reaches the correspondence threshold: local_global_registration.py:185-190)."""
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
    # tag: (make_lgr_inputs kwargs, LocalGlobalRegistration kwargs)
    '3dmatch': (dict(num_patches=128, K=64, seed=51), dict(k=3, acceptance_radius=0.1)),
    'kitti': (dict(num_patches=48, K=128, seed=52), dict(k=2, acceptance_radius=0.1)),
    'small': (dict(num_patches=6, K=12, seed=53), dict(k=3, acceptance_radius=0.1)),
    'nonmutual': (dict(num_patches=16, K=32, seed=54), dict(k=2, acceptance_radius=0.1, mutual=False)),
    'global_score': (dict(num_patches=16, K=32, seed=55), dict(k=3, acceptance_radius=0.1, use_global_score=True)),
    'degenerate': (dict(num_patches=5, K=8, seed=56, valid_frac=0.3), dict(k=1, acceptance_radius=0.1,
                                                                         correspondence_threshold=9)),
}


def main():
    load_reference()
    from geotransformer.modules.geotransformer.local_global_registration import LocalGlobalRegistration
    synth = pkgpath.load_package().synth
    torch.set_grad_enabled(False)
    out = dict(meta=repr(dict(torch=torch.__version__)))
    saved = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        for tag, (kw, mkw) in CASES.items():
            inp = synth.make_lgr_inputs(**kw)
            mod = LocalGlobalRegistration(**mkw)
            rc, sc, s, T = mod(*inp[:6])
            out.update({f'{tag}.ref_corr_points': rc.numpy(), f'{tag}.src_corr_points': sc.numpy(),
                        f'{tag}.corr_scores': s.numpy(), f'{tag}.transform': T.numpy()})
            print(tag, tuple(rc.shape), float((T - inp[6]).abs().max()))
    finally:
        torch.Tensor.cuda = saved
    np.savez_compressed(os.path.join(HERE, 'lgr_cases.npz'), **out)
    print(os.path.getsize(os.path.join(HERE, 'lgr_cases.npz')))


if __name__ == '__main__':
    main()
