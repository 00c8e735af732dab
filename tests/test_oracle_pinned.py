"""CPU tests: the oracle (oracle/geotx_oracle.py + oracle/knn_oracle.c) is pinned against golden vectors produced by
the UNMODIFIED reference (tests/golden/make_golden.py) and, when /root/reference is present, against the live
reference.  Bit-exact when the torch build matches the one that wrote the fixtures."""
import ctypes
import os

import numpy as np
import pytest
import torch

from load_reference import load_reference, reference_available


def _same_torch(meta):
    return torch.__version__ in str(meta)


def _close(a, b, exact, tol=1e-5):
    a = torch.as_tensor(np.asarray(a))
    b = torch.as_tensor(np.asarray(b))
    if exact:
        assert torch.equal(a, b)
    else:
        assert ((a - b).norm() / b.norm().clamp_min(1e-30)).item() <= tol


def _sd_from_npz(z):
    return {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith('sd.')}


@pytest.mark.parametrize('tag', ['tiny_max', 'tiny_mean'])
def test_tiny_golden_bit_exact(oracle, golden, tag):
    z = golden(f'{tag}.npz')
    cfg = eval(str(z['cfg']))
    exact = _same_torch(z['meta'])
    sd = _sd_from_npz(z)
    rp, sp = torch.from_numpy(z['ref_points']), torch.from_numpy(z['src_points'])
    rf, sf = torch.from_numpy(z['ref_feats']), torch.from_numpy(z['src_feats'])
    d_idx, a_idx, _ = oracle.embedding_indices(rp, cfg['sigma_d'], cfg['sigma_a'], cfg['angle_k'])
    _close(d_idx, z['d_idx'], exact)
    _close(a_idx, z['a_idx'], exact)
    e0 = oracle.gse(rp, sd, 'embedding.', cfg['sigma_d'], cfg['sigma_a'], cfg['angle_k'], cfg['reduction_a'])
    _close(e0, z['E_ref'], exact)
    o0, o1 = oracle.geometric_transformer(sd, cfg, rp, sp, rf, sf)
    _close(o0, z['out_ref'], exact)
    _close(o1, z['out_src'], exact)
    # init_state_dict reproduces the reference's default init under seed 7351 without importing it
    sd2 = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    assert list(sd2.keys()) == list(sd.keys())
    if exact:
        assert all(torch.equal(sd[k], sd2[k]) for k in sd)


@pytest.mark.parametrize('name', ['3dmatch_n192', 'kitti_n256', 'modelnet_n128'])
def test_config_golden(oracle, golden, pkg, name):
    z = golden(f'{name}.npz')
    cfg = eval(str(z['cfg']))
    exact = _same_torch(z['meta'])
    rp, sp, rf, sf, used = pkg.synth.make_tie_free_pair(str(z['shape']), int(z['n']), int(z['m']), seed=int(z['seed']))
    assert used == int(z['seed'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    knn = oracle.knn_indices(rp, cfg['angle_k'])
    assert np.array_equal(knn[0].numpy().astype(np.int32), z['knn_ref'])
    rows = torch.from_numpy(z['rows'])
    e = oracle.gse(rp, sd, 'embedding.', cfg['sigma_d'], cfg['sigma_a'], cfg['angle_k'])
    _close(e[0, rows], z['E_rows'], exact)
    o0, o1 = oracle.geometric_transformer(sd, cfg, rp, sp, rf, sf)
    _close(o0, z['out_ref'], exact)
    _close(o1, z['out_src'], exact)
    # the row-chunked oracle used for large n is the same math
    c0, c1 = oracle.geometric_transformer_chunked(sd, cfg, rp, sp, rf, sf, rows=50)
    _close(c0, o0, False, 1e-5)
    _close(c1, o1, False, 1e-5)


def test_modelnet717_full_golden(oracle, golden, pkg):
    """BASELINE.json configs[0]: ModelNet config on one 717-point pair, fp32 on CPU."""
    z = golden('modelnet_n717_out.npz')
    cfg = eval(str(z['cfg']))
    rp, sp, rf, sf, _ = pkg.synth.make_tie_free_pair('modelnet', 717, 717, seed=int(z['seed']))
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    o0, o1 = oracle.geometric_transformer_chunked(sd, cfg, rp, sp, rf, sf, rows=128)
    _close(o0, z['out_ref'], False, 1e-5)
    _close(o1, z['out_src'], False, 1e-5)


def _c_knn(knn_c, pts, k, sigma_d=1.0, want_d=False):
    n = pts.shape[0]
    p = np.ascontiguousarray(pts.numpy(), dtype=np.float32)
    out = np.zeros((n, k), np.int64)
    dd = np.zeros((n, n), np.float32) if want_d else None
    knn_c.geotx_oracle_knn(p.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(n), ctypes.c_int(k), ctypes.c_float(sigma_d),
                           out.ctypes.data_as(ctypes.c_void_p),
                           dd.ctypes.data_as(ctypes.c_void_p) if want_d else ctypes.c_void_p(0))
    return out, dd


def test_knn_cases_torch_and_c_oracle(oracle, golden, pkg, knn_c):
    """Integer work: both oracles reproduce the reference's k-NN indices bit for bit at every BASELINE size."""
    z = golden('knn_cases.npz')
    for shape, n in (('modelnet', 717), ('3dmatch', 1024), ('kitti', 2048), ('3dmatch', 4096)):
        for seed in (pkg.synth.SEED, pkg.synth.SEED + 1):
            rp, sp, _, _, used = pkg.synth.make_tie_free_pair(shape, n, n, seed=seed, input_dim=4)
            assert used == int(z[f'{shape}_{n}_{seed}_seed_used'])
            for nm, pts in (('ref', rp), ('src', sp)):
                want = z[f'{shape}_{n}_{seed}_{nm}']
                got_c, _ = _c_knn(knn_c, pts[0], 3)
                assert np.array_equal(got_c.astype(np.int32), want), (shape, n, seed, nm, 'C oracle')
                if n <= 2048:
                    got_t = oracle.knn_indices(pts, 3)[0].numpy().astype(np.int32)
                    assert np.array_equal(got_t, want), (shape, n, seed, nm, 'torch oracle')


def test_c_oracle_distance_recipe(oracle, pkg, knn_c):
    """d_idx of the C restatement vs torch-CPU: identical d2 recipe; torch's vectorised sqrt is 1 ulp off IEEE
    in <1 % of entries (documented in knn_oracle.c), never more."""
    rp, _, _, _ = pkg.synth.make_pair('kitti', 512, 512, seed=3, input_dim=4)
    d_t, _, _ = oracle.embedding_indices(rp, 4.8, 15, 3)
    _, d_c = _c_knn(knn_c, rp[0], 3, sigma_d=4.8, want_d=True)
    a = d_t[0].numpy()
    assert np.abs(a - d_c).max() <= 4e-7 * np.abs(a).max()
    d2_t = oracle.pairwise_sqdist(rp, rp)[0].numpy()
    _, d_c1 = _c_knn(knn_c, rp[0], 3, sigma_d=1.0, want_d=True)
    assert np.array_equal(np.sqrt(d2_t), d_c1)  # IEEE sqrt of the identical d2


def test_edge_cases_oracle(oracle, knn_c):
    pts = torch.rand(1, 4, 3, generator=torch.Generator().manual_seed(0))
    knn = oracle.knn_indices(pts, 3)  # n == k+1
    got, _ = _c_knn(knn_c, pts[0], 3)
    assert np.array_equal(got, knn[0].numpy())
    with pytest.raises(RuntimeError):
        oracle.knn_indices(pts[:, :3], 3)  # n < k+1: torch.topk raises, as the reference does


@pytest.mark.skipif(not reference_available(), reason='/root/reference not present (GPU box)')
def test_oracle_equals_live_reference(oracle, pkg):
    GT, _ = load_reference()
    cfg = dict(pkg.synth.CONFIGS['kitti'])
    torch.manual_seed(7351)
    ref = GT(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['num_heads'], cfg['blocks'], cfg['sigma_d'],
             cfg['sigma_a'], cfg['angle_k'], reduction_a=cfg['reduction_a']).eval()
    sd = {k: v.detach().clone() for k, v in ref.state_dict().items()}
    rp, sp, rf, sf = pkg.synth.make_pair('kitti', 150, 120, seed=5)
    masks0 = torch.zeros(1, 150, dtype=torch.bool)
    masks0[0, 140:] = True
    with torch.no_grad():
        a0, a1 = ref(rp, sp, rf, sf, ref_masks=masks0)
    b0, b1 = oracle.geometric_transformer(sd, cfg, rp, sp, rf, sf, ref_masks=masks0)
    assert torch.equal(a0, b0) and torch.equal(a1, b1)


SPM_CASES = ['small_dual', 'small_plain', 'small_masked', 'small_allpairs', 'mid_dual']


@pytest.mark.parametrize('tag', SPM_CASES)
def test_superpoint_matching_oracle_vs_golden(oracle, golden, tag):
    """the oracle's F.normalize + SuperPointMatching restatement against vectors written by the
    reference class itself (tests/golden/make_golden_spm.py)."""
    z = golden('spm_cases.npz')
    cfg = eval(str(z[f'{tag}.cfg']))
    exact = _same_torch(z['meta'])
    ri, si, sc = oracle.superpoint_matching(
        torch.from_numpy(z[f'{tag}.ref_feats']), torch.from_numpy(z[f'{tag}.src_feats']), cfg['K'], cfg['dual'],
        torch.from_numpy(z[f'{tag}.ref_masks']), torch.from_numpy(z[f'{tag}.src_masks']), normalize=True)
    _close(sc, z[f'{tag}.scores'], exact)
    if exact:
        assert np.array_equal(ri.numpy(), z[f'{tag}.ref_idx']) and np.array_equal(si.numpy(), z[f'{tag}.src_idx'])


def test_superpoint_matching_oracle_on_transformer_output(oracle, golden):
    z, zin = golden('spm_cases.npz'), golden('3dmatch_n1024_out.npz')
    exact = _same_torch(z['meta'])
    ri, si, sc = oracle.superpoint_matching(torch.from_numpy(zin['out_ref'][0]), torch.from_numpy(zin['out_src'][0]), 256,
                                            True, normalize=True)
    _close(sc, z['3dmatch_1024.scores'], exact)
    if exact:
        assert np.array_equal(ri.numpy(), z['3dmatch_1024.ref_idx'])
        assert np.array_equal(si.numpy(), z['3dmatch_1024.src_idx'])


@pytest.mark.skipif(not reference_available(), reason='/root/reference not present (GPU box)')
def test_superpoint_matching_oracle_vs_live_reference(oracle):
    load_reference()
    from geotransformer.modules.geotransformer.superpoint_matching import SuperPointMatching
    g = torch.Generator().manual_seed(99)
    ref, src = torch.randn(150, 64, generator=g), torch.randn(131, 64, generator=g)
    rm, sm = torch.rand(150, generator=g) > 0.1, torch.rand(131, generator=g) > 0.1
    rn, sn = torch.nn.functional.normalize(ref, p=2, dim=1), torch.nn.functional.normalize(src, p=2, dim=1)
    for dual in (True, False):
        want = SuperPointMatching(77, dual)(rn, sn, rm, sm)
        got = oracle.superpoint_matching(ref, src, 77, dual, rm, sm, normalize=True)
        for a, b in zip(got, want):
            assert torch.equal(a, b)
