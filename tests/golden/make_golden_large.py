"""Golden outputs of the UNMODIFIED reference at the BASELINE.json sizes it cannot run in one piece:
KITTI 2048 (configs[2]), 3DMatch 2048 / 4096 (configs[3], north_star's target size) and 3DMatch 1024 / 921 (M != N).

The reference's GeometricTransformer.forward materialises (n, n, k, C) fp32 tensors (51.5 GB at n = 4096), so it is
driven here in QUERY-ROW CHUNKS through its own, unmodified submodules (imported from /root/reference with the stubs of
oracle/load_reference.py):
    get_embedding_indices(points)                                  geotransformer.py:25-55   (O(n^2 k) memory, unchunked)
    proj_d(embedding(d_idx[:, I])) + max_k proj_a(embedding(a_idx[:, I]))     :60-70, rows I of E
    RPETransformerLayer(x[:, I], x, E[:, I])                       rpe_transformer.py:113-131, rows I of a 'self' block
    TransformerLayer(x0, x1)                                       vanilla_transformer.py:111-129, whole ('cross')
Rows of E and of an attention layer are independent, so this is the same arithmetic; `--check` proves it against the
unchunked reference forward at n = 512.  E is recomputed for each of the three 'self' blocks instead of stored (17 GB
per cloud at n = 4096).

Stored (to keep the repository small): 128 evenly spaced rows of both output feature matrices, and for 4 query rows
every 8th column of E.  Inputs and weights are regenerated from seeds by the tests (synth.make_tie_free_pair, seed 7351;
default init under torch.manual_seed(7351)), their digests are stored.

    python tests/golden/make_golden_large.py [--check] [case ...]        (build container only; ~25 min on 8 cores)
"""
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
sys.path.insert(0, HERE)
from load_reference import load_reference  # noqa: E402
from make_golden import _synth, build, digest  # noqa: E402

CASES = {  # name: (shape, n, m, rows per chunk)
    'kitti_n2048': ('kitti', 2048, 2048, 128),
    '3dmatch_n2048': ('3dmatch', 2048, 2048, 64),
    '3dmatch_n4096': ('3dmatch', 4096, 4096, 32),
    '3dmatch_n1024_m921': ('3dmatch', 1024, 921, 128),
}


def e_rows(gse, d_idx, a_idx, I):
    """geotransformer.py:60-70 for the query rows I (the reference's own modules and operation order)."""
    d_embeddings = gse.proj_d(gse.embedding(d_idx[:, I]))
    a_embeddings = gse.proj_a(gse.embedding(a_idx[:, I]))
    if gse.reduction_a == 'max':
        a_embeddings = a_embeddings.max(dim=3)[0]
    else:
        a_embeddings = a_embeddings.mean(dim=3)
    return d_embeddings + a_embeddings


def chunked_forward(model, rp, sp, rf, sf, rows):
    """GeometricTransformer.forward (geotransformer.py:114-155) in query-row chunks; returns the outputs."""
    gse = model.embedding
    idx = [gse.get_embedding_indices(p) for p in (rp, sp)]
    feats = [model.in_proj(rf), model.in_proj(sf)]
    for i, block in enumerate(model.transformer.blocks):
        layer = model.transformer.layers[i]
        if block == 'self':  # conditional_transformer.py:101-102
            for c in (0, 1):
                x, (d_idx, a_idx) = feats[c], idx[c]
                out = []
                for lo in range(0, x.shape[1], rows):
                    I = slice(lo, min(lo + rows, x.shape[1]))
                    out.append(layer(x[:, I], x, e_rows(gse, d_idx, a_idx, I))[0])
                feats[c] = torch.cat(out, dim=1)
        else:  # conditional_transformer.py:110-111 (parallel=False): the second step sees the UPDATED first cloud
            feats[0] = layer(feats[0], feats[1])[0]
            feats[1] = layer(feats[1], feats[0])[0]
    return model.out_proj(feats[0]), model.out_proj(feats[1]), idx


def main():
    S = _synth()
    GT, _ = load_reference()
    torch.set_grad_enabled(False)
    meta = dict(torch=torch.__version__, weight_seed=7351, how='reference submodules in query-row chunks')
    args = [a for a in sys.argv[1:] if not a.startswith('--')]
    if '--check' in sys.argv:
        for shape, n, m in (('3dmatch', 512, 460), ('kitti', 384, 384)):
            cfg = dict(S.CONFIGS[shape])
            model = build(GT, cfg)
            rp, sp, rf, sf, _ = S.make_tie_free_pair(shape, n, m, seed=S.SEED)
            w0, w1 = model(rp, sp, rf, sf)
            g0, g1, _ = chunked_forward(model, rp, sp, rf, sf, 96)
            e = max(float((g0 - w0).norm() / w0.norm()), float((g1 - w1).norm() / w1.norm()))
            print(f'check {shape} {n}/{m}: chunked vs unchunked reference forward rel {e:.2e}')
            assert e < 2e-6
    for name in (args or list(CASES)):
        shape, n, m, rows = CASES[name]
        cfg = dict(S.CONFIGS[shape])
        model = build(GT, cfg)
        rp, sp, rf, sf, used = S.make_tie_free_pair(shape, n, m, seed=S.SEED)
        t0 = time.time()  # `used` = the seed make_tie_free_pair settled on (it bumps the seed until no k-NN row has a tie)
        o0, o1, idx = chunked_forward(model, rp, sp, rf, sf, rows)
        out_rows_ref = torch.arange(0, n, max(1, n // 128))[:128]
        out_rows_src = torch.arange(0, m, max(1, m // 128))[:128]
        erows = torch.tensor([0, n // 3, n // 2 + 1, n - 1])
        ecols = torch.arange(0, n, 8)
        E = torch.cat([e_rows(model.embedding, idx[0][0], idx[0][1], slice(int(r), int(r) + 1)) for r in erows], dim=1)
        np.savez_compressed(
            os.path.join(HERE, f'{name}_rows.npz'),
            out_rows_ref=out_rows_ref.numpy(), out_rows_src=out_rows_src.numpy(),
            out_ref=o0[0, out_rows_ref].numpy(), out_src=o1[0, out_rows_src].numpy(),
            out_ref_norm=float(o0.norm()), out_src_norm=float(o1.norm()),
            E_rows=erows.numpy(), E_cols=ecols.numpy(), E=E[0][:, ecols].numpy(),
            n=n, m=m, seed=S.SEED, seed_used=used, shape=shape, input_digest=digest(rp, sp, rf, sf),
            weight_digest=digest(*model.state_dict().values()), cfg=repr(cfg), meta=repr(meta))
        print(f'{name}: done in {time.time() - t0:.0f} s', flush=True)


if __name__ == '__main__':
    main()
