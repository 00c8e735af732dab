"""One call of K12 and K11 at the 3DMatch and KITTI sizes (for ncu captures)."""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
import pkgpath

pkg = pkgpath.load_package()
for K, k, n, M in ((64, 3, 16384, 1024), (128, 2, 49152, 2048)):
    g = torch.Generator().manual_seed(pkg.synth.SEED)
    feats = [torch.randn(n, 256, generator=g).cuda() for _ in range(2)]
    knn = [torch.randint(0, n + 1, (M, K), generator=g).cuda() for _ in range(2)]
    corr = [torch.randint(0, M, (256,), generator=g).cuda() for _ in range(2)]
    inp = [t.cuda() for t in pkg.synth.make_lgr_inputs(num_patches=256, K=K, seed=pkg.synth.SEED)[:6]]
    for _ in range(2):
        s = pkg.ops.patch_scores(feats[0], feats[1], knn[0], knn[1], corr[0], corr[1])
        out = pkg.ops.local_global_registration(*inp[:5], None, k, 0.1)
    torch.cuda.synchronize()
    print(K, float(s.abs().max()), int(out[3].item()))
