"""-m gpu: K1 (fused pairwise distance + top-(k+1)) and the d_idx / a_idx kernel, through the C ABI.
Bit-exact against (a) golden indices written by the reference itself, (b) the plain-C oracle, (c) the torch oracle."""
import ctypes
import hashlib

import numpy as np
import pytest
import torch

from gpu_common import DEV

pytestmark = pytest.mark.gpu


def _digest(t):
    return hashlib.sha256(np.ascontiguousarray(t.numpy()).tobytes()).hexdigest()


@pytest.mark.parametrize('shape,n', [('modelnet', 717), ('3dmatch', 1024), ('kitti', 2048), ('3dmatch', 4096)])
def test_knn_bit_exact_vs_reference_golden(pkg, golden, shape, n):
    """BASELINE.json sizes (717 / 1024 / 2048 / 4096), two seeds, both clouds: indices equal the reference's."""
    z = golden('knn_cases.npz')
    for seed in (pkg.synth.SEED, pkg.synth.SEED + 1):
        rp, sp, _, _, used = pkg.synth.make_tie_free_pair(shape, n, n, seed=seed, input_dim=4)
        assert used == int(z[f'{shape}_{n}_{seed}_seed_used'])
        for nm, pts in (('ref', rp), ('src', sp)):
            assert _digest(pts) == str(z[f'{shape}_{n}_{seed}_{nm}_digest']), 'synthetic input drifted'
            geom = pkg.ops.gse_knn(pts.to(DEV), 3)
            got = geom.knn[0].cpu().numpy().astype(np.int32)
            assert np.array_equal(got, z[f'{shape}_{n}_{seed}_{nm}']), f'{shape} n={n} seed={seed} {nm}'


@pytest.mark.parametrize('k', [1, 3, 7])
@pytest.mark.parametrize('n', [8, 33, 500])
def test_knn_vs_c_oracle_ragged(pkg, knn_c, n, k):
    """Ragged sizes / every supported k / batch > 1 against the plain-C restatement."""
    g = torch.Generator().manual_seed(n * 10 + k)
    pts = torch.rand(3, n, 3, generator=g) * 2.0
    geom = pkg.ops.gse_knn(pts.to(DEV), k)
    got = geom.knn.cpu()
    for b in range(3):
        exp = np.empty((n, k), dtype=np.int64)
        p = np.ascontiguousarray(pts[b].numpy())
        knn_c.geotx_oracle_knn(p.ctypes.data_as(ctypes.c_void_p), n, k, ctypes.c_float(0.2),
                               exp.ctypes.data_as(ctypes.c_void_p), None)
        assert np.array_equal(got[b].numpy(), exp)


def test_knn_minimum_size_and_errors(pkg):
    pts = torch.rand(1, 4, 3, generator=torch.Generator().manual_seed(1))
    geom = pkg.ops.gse_knn(pts.to(DEV), 3)  # n == k+1: every other point is a neighbour
    assert sorted(geom.knn[0, 0].tolist()) == [1, 2, 3]
    with pytest.raises(RuntimeError, match='out of range'):
        pkg.ops.gse_knn(pts[:, :3].contiguous().to(DEV), 3)  # torch.topk raises here too (geotransformer.py:42)
    with pytest.raises(RuntimeError, match='CUDA tensor'):
        pkg.ops.gse_knn(pts, 3)  # no CPU path
    with pytest.raises(RuntimeError, match='angle_k'):
        pkg.ops.gse_knn(pts.to(DEV), 9)


@pytest.mark.parametrize('tag', ['tiny_max', '3dmatch_n192', 'kitti_n256'])
def test_embedding_indices_vs_golden(pkg, golden, tag):
    """d_idx within 1 ulp (torch-CPU's vectorised sqrt is itself 1 ulp off IEEE on 0.6 % of entries), a_idx <= 2e-5."""
    z = golden(f'{tag}.npz')
    cfg = eval(str(z['cfg']))
    if tag == 'tiny_max':
        pts = torch.from_numpy(z['ref_points'])
        d_exp, a_exp = torch.from_numpy(z['d_idx'])[0], torch.from_numpy(z['a_idx'])[0]
        rows = torch.arange(pts.shape[1])
    else:
        pts = pkg.synth.make_tie_free_pair(str(z['shape']), int(z['n']), int(z['m']), seed=int(z['seed']))[0]
        d_exp, a_exp = torch.from_numpy(z['d_idx_rows']), torch.from_numpy(z['a_idx_rows'])
        rows = torch.from_numpy(z['rows'])
    m = pkg.GeometricStructureEmbedding(cfg['hidden_dim'], cfg['sigma_d'], cfg['sigma_a'], cfg['angle_k']).to(DEV)
    d_idx, a_idx = m.get_embedding_indices(pts.to(DEV))
    d_got, a_got = d_idx[0].cpu()[rows], a_idx[0].cpu()[rows]
    ulp = torch.abs(d_exp) * 2.0 ** -23 + 1e-30
    # off the diagonal; on the diagonal d2 is a cancellation residue whose sqrt/divide is still the same op chain
    assert (torch.abs(d_got - d_exp) <= 2.0 * ulp).all()
    assert torch.abs(a_got - a_exp).max().item() <= 2e-5 * (180.0 / (cfg['sigma_a'] * np.pi)) * np.pi
