"""Runs K11 a few times on a 3DMatch- and a KITTI-sized input (for ncu launch lists)."""
import sys
import os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
import pkgpath

pkg = pkgpath.load_package()
for K, k in ((64, 3), (128, 2)):
    inp = [t.cuda() for t in pkg.synth.make_lgr_inputs(num_patches=256, K=K, seed=pkg.synth.SEED)[:6]]
    for _ in range(3):
        out = pkg.ops.local_global_registration(*inp[:5], None, k, 0.1)
    torch.cuda.synchronize()
    print(K, k, int(out[3].item()))
