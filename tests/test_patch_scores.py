"""K12 patch matching scores (experiments/*/model.py:169-170,176-179,187-188: the glue between coarse matching and
optimal transport).  The reference has no function for these lines, so the oracle is their line-by-line restatement
(oracle.patch_matching_scores; parity unpinned by golden vectors -- the lines are four torch one-liners); -m gpu tests
compare the CUDA kernel with it.  Tolerance: fp32 sums of C products in a different order, |diff| <= 1e-5 of the largest
score."""
import pytest
import torch


def _inputs(n_ref, n_src, C, M, K, P, seed, pad_frac=0.2):
    g = torch.Generator().manual_seed(seed)
    ref_feats, src_feats = torch.randn(n_ref, C, generator=g), torch.randn(n_src, C, generator=g)
    ref_knn = torch.randint(0, n_ref, (M, K), generator=g)
    src_knn = torch.randint(0, n_src, (M, K), generator=g)
    ref_knn[torch.rand(M, K, generator=g) < pad_frac] = n_ref  # padded slots (pointcloud_partition.py:97)
    src_knn[torch.rand(M, K, generator=g) < pad_frac] = n_src
    return (ref_feats, src_feats, ref_knn, src_knn, torch.randint(0, M, (P,), generator=g),
            torch.randint(0, M, (P,), generator=g))


def test_oracle_restatement(oracle):
    inp = _inputs(500, 400, 32, 40, 12, 9, seed=1)
    s = oracle.patch_matching_scores(*inp)
    assert s.shape == (9, 12, 12)
    p, i, j = 3, 5, 7
    a, b = inp[2][inp[4][p], i], inp[3][inp[5][p], j]
    want = 0.0 if a == 500 or b == 400 else float(inp[0][a] @ inp[1][b]) / 32 ** 0.5
    assert abs(float(s[p, i, j]) - want) < 1e-5
    assert bool((s[inp[2][inp[4]] == 500] == 0).all())  # a padded ref point has a zero score row


def test_host_api_without_gpu(pkg):
    with pytest.raises(RuntimeError):
        pkg.patch_matching_scores(*_inputs(100, 100, 16, 10, 8, 4, seed=2))


@pytest.mark.gpu
@pytest.mark.parametrize('n_ref,n_src,C,M,K,P', [
    (16384, 14000, 256, 1024, 64, 256),   # 3DMatch (config.py:88,113: 64 points per patch, 256 correspondences)
    (40000, 38000, 256, 2048, 128, 256),  # KITTI (128 points per patch)
    (900, 700, 36, 50, 20, 7),            # C not a multiple of the 64-channel step, K not a multiple of 4
    (3000, 3000, 128, 200, 100, 33),
    (50, 60, 4, 5, 1, 3),
])
def test_gpu_vs_oracle(pkg, oracle, n_ref, n_src, C, M, K, P):
    from gpu_common import DEV
    inp = _inputs(n_ref, n_src, C, M, K, P, seed=n_ref + K)
    want = oracle.patch_matching_scores(*inp)
    got = pkg.patch_matching_scores(*[t.to(DEV) for t in inp]).cpu()
    assert got.shape == want.shape
    assert float((got - want).abs().max()) <= 1e-5 * float(want.abs().max())
    assert bool((got[inp[2][inp[4]] == n_ref] == 0).all())


@pytest.mark.gpu
def test_gpu_errors(pkg):
    from gpu_common import DEV
    inp = [t.to(DEV) for t in _inputs(100, 100, 16, 10, 8, 4, seed=3)]
    with pytest.raises(RuntimeError):
        pkg.patch_matching_scores(inp[0][:, :15].contiguous(), inp[1][:, :15].contiguous(), *inp[2:])  # C % 4
    with pytest.raises(RuntimeError):
        pkg.patch_matching_scores(inp[0], inp[1], inp[2].int(), *inp[3:])  # int32 indices
    empty = pkg.patch_matching_scores(*inp[:4], inp[4][:0], inp[5][:0])
    assert empty.shape == (0, 8, 8)
