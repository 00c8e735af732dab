"""Generates the committed golden vectors by running the UNMODIFIED reference (imported from
/root/reference with the four sys.modules stubs of oracle/load_reference.py) on seeded synthetic
inputs.  Runs only in the build container; the .npz files it writes travel with the repo.

    python tests/golden/make_golden.py

Recorded with every file: torch version, seeds, config.  Weights = default nn.Linear/LayerNorm init
under torch.manual_seed(7351) (the seed of experiments/*/config.py:13).
"""
import hashlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import pkgpath  # noqa: E402
from load_reference import load_reference  # noqa: E402



def _synth():
    import importlib.util
    spec = importlib.util.spec_from_file_location('geotx_synth', os.path.join(pkgpath.PKG_DIR, 'synth.py'))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def digest(*tensors):
    h = hashlib.sha256()
    for t in tensors:
        h.update(np.ascontiguousarray(t.detach().numpy()).tobytes())
    return h.hexdigest()


def build(GT, cfg, seed=7351):
    torch.manual_seed(seed)
    return GT(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['num_heads'], cfg['blocks'],
              cfg['sigma_d'], cfg['sigma_a'], cfg['angle_k'], reduction_a=cfg.get('reduction_a', 'max')).eval()


def main():
    S = _synth()
    GT, GSE = load_reference()
    torch.set_grad_enabled(False)
    meta = dict(torch=torch.__version__, weight_seed=7351)

    # ---- case 1: tiny, everything stored (inputs, weights, indices, E, per-block trace, outputs) ----
    cfg = dict(input_dim=16, output_dim=24, hidden_dim=32, num_heads=2, blocks=['self', 'cross', 'self', 'cross'],
               sigma_d=0.2, sigma_a=15, angle_k=3, reduction_a='max')
    for tag, red in (('tiny_max', 'max'), ('tiny_mean', 'mean')):
        cfg['reduction_a'] = red
        m = build(GT, cfg)
        rp, sp, rf, sf = S.make_pair('modelnet', 40, 33, seed=11, input_dim=16)
        d0, a0 = m.embedding.get_embedding_indices(rp)
        e0 = m.embedding(rp)
        e1 = m.embedding(sp)
        o0, o1 = m(rp, sp, rf, sf)
        out = {f'sd.{k}': v.numpy() for k, v in m.state_dict().items()}
        out.update(ref_points=rp.numpy(), src_points=sp.numpy(), ref_feats=rf.numpy(), src_feats=sf.numpy(),
                   d_idx=d0.numpy(), a_idx=a0.numpy(), E_ref=e0.numpy(), E_src=e1.numpy(),
                   out_ref=o0.numpy(), out_src=o1.numpy(), cfg=repr(cfg), meta=repr(meta))
        np.savez_compressed(os.path.join(HERE, f'{tag}.npz'), **out)
        print(tag, 'done')

    # ---- case 2: the three shipped configs at reduced n: knn, sampled E rows, outputs; inputs by seed ----
    for shape, n, mm in (('3dmatch', 192, 160), ('kitti', 256, 256), ('modelnet', 128, 128)):
        cfg = dict(S.CONFIGS[shape])
        m = build(GT, cfg)
        rp, sp, rf, sf, used = S.make_tie_free_pair(shape, n, mm, seed=S.SEED)
        assert used == S.SEED
        d0, a0 = m.embedding.get_embedding_indices(rp)
        knn0 = torch.sqrt(__import__('geotransformer.modules.ops', fromlist=['pairwise_distance'])
                          .pairwise_distance(rp, rp)).topk(k=4, dim=2, largest=False)[1][:, :, 1:]
        rows = torch.arange(0, n, max(1, n // 8))[:8]
        e0 = m.embedding(rp)
        o0, o1 = m(rp, sp, rf, sf)
        np.savez_compressed(
            os.path.join(HERE, f'{shape}_n{n}.npz'),
            knn_ref=knn0[0].numpy().astype(np.int32), rows=rows.numpy(), E_rows=e0[0, rows].numpy(),
            d_idx_rows=d0[0, rows].numpy(), a_idx_rows=a0[0, rows].numpy(),
            out_ref=o0.numpy(), out_src=o1.numpy(), n=n, m=mm, seed=S.SEED, shape=shape,
            input_digest=digest(rp, sp, rf, sf), weight_digest=digest(*m.state_dict().values()),
            cfg=repr(cfg), meta=repr(meta))
        print(shape, n, 'done')

    # ---- case 3: k-NN indices only at the BASELINE.json sizes (bit-exact integer work) ----
    pd = __import__('geotransformer.modules.ops', fromlist=['pairwise_distance']).pairwise_distance
    knn_cases = {}
    for shape, n in (('modelnet', 717), ('3dmatch', 1024), ('kitti', 2048), ('3dmatch', 4096)):
        for seed in (S.SEED, S.SEED + 1):
            rp, sp, _, _, used = S.make_tie_free_pair(shape, n, n, seed=seed, input_dim=4)
            knn_cases[f'{shape}_{n}_{seed}_seed_used'] = used
            for nm, pts in (('ref', rp), ('src', sp)):
                knn = torch.sqrt(pd(pts, pts)).topk(k=4, dim=2, largest=False)[1][:, :, 1:]
                knn_cases[f'{shape}_{n}_{seed}_{nm}'] = knn[0].numpy().astype(np.int32)
                knn_cases[f'{shape}_{n}_{seed}_{nm}_digest'] = digest(pts)
        print('knn', shape, n)
    np.savez_compressed(os.path.join(HERE, 'knn_cases.npz'), meta=repr(meta), **knn_cases)

    # ---- case 4: full-size outputs of BASELINE configs[0] (ModelNet 717) and configs[1] (3DMatch 1024) ----
    for shape, n in (('modelnet', 717), ('3dmatch', 1024)):
        cfg = dict(S.CONFIGS[shape])
        m = build(GT, cfg)
        rp, sp, rf, sf, used = S.make_tie_free_pair(shape, n, n, seed=S.SEED)
        assert used == S.SEED
        o0, o1 = m(rp, sp, rf, sf)
        np.savez_compressed(os.path.join(HERE, f'{shape}_n{n}_out.npz'), out_ref=o0.numpy().astype(np.float32),
                            out_src=o1.numpy(), input_digest=digest(rp, sp, rf, sf),
                            weight_digest=digest(*m.state_dict().values()), cfg=repr(cfg), meta=repr(meta),
                            n=n, seed=S.SEED, shape=shape)
        print('full', shape, n)


if __name__ == '__main__':
    main()
