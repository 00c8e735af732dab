"""-m gpu: K6 point-to-node partition and K7 log-domain optimal transport (8(f)-3), through the
C ABI.  Checkers: golden vectors written by the reference's own functions (tests/golden/make_golden_next.py), the
plain-C oracle (bit-exact, torch-independent) and the torch oracle.
Tolerances: partition outputs are index / byte work: bit-exact.  Transport: fp32 on both sides with different
exp/log implementations and summation orders: |diff| <= 1e-4 + 1e-4 |ref| on finite entries (measured ~3e-6);
masked entries (about -1e12) to 1e-6 relative."""
import numpy as np
import pytest
import torch

from gpu_common import DEV
from test_oracle_next import LOT_CASES, P2N_CASES, _p2n_inputs, c_partition

pytestmark = pytest.mark.gpu


def _gpu_partition(pkg, points, nodes, K):
    out = pkg.ops.point_to_node_partition(points.to(DEV), nodes.to(DEV), K, return_count=True, return_points=True)
    torch.cuda.synchronize()
    p2n, sizes, masks, idx, kmask, kpts = out
    assert p2n.dtype == torch.int64 and sizes.dtype == torch.int64 and idx.dtype == torch.int64
    assert masks.dtype == torch.bool and kmask.dtype == torch.bool and kpts.dtype == torch.float32
    return [t.cpu() for t in out]


def _check_patch_points(points, idx, kpts):
    padded = torch.cat([points, torch.zeros_like(points[:1])], dim=0)  # model.py:105
    assert torch.equal(kpts, padded[idx])  # model.py:107 index_select


@pytest.mark.parametrize('tag', P2N_CASES)
def test_partition_vs_reference_golden(pkg, golden, tag):
    z = golden('next_cases.npz')
    points, nodes, K = _p2n_inputs(z, tag)
    p2n, sizes, masks, idx, kmask, kpts = _gpu_partition(pkg, points, nodes, K)
    assert np.array_equal(p2n.numpy(), z[f'{tag}.point_to_node'])
    assert np.array_equal(sizes.numpy(), z[f'{tag}.node_sizes'])
    assert np.array_equal(masks.numpy(), z[f'{tag}.node_masks'])
    assert np.array_equal(idx.numpy(), z[f'{tag}.node_knn_indices'])
    assert np.array_equal(kmask.numpy(), z[f'{tag}.node_knn_masks'])
    _check_patch_points(points, idx, kpts)
    # the drop-in function returns the reference's 4-tuple / 5-tuple
    four = pkg.point_to_node_partition(points.to(DEV), nodes.to(DEV), K)
    assert len(four) == 4 and torch.equal(four[2].cpu(), idx)


@pytest.mark.parametrize('shape,num_nodes,ppn,K', [('3dmatch', 1024, 24, 64), ('kitti', 2048, 24, 128),
                                                     ('3dmatch', 4096, 16, 64), ('3dmatch', 300, 200, 64)])
def test_partition_full_size_vs_c_oracle(pkg, oracle, knn_c, shape, num_nodes, ppn, K):
    """BASELINE-size clouds (and a 4096-node sweep point, and 200-point nodes that overflow the limit): the GPU result
    equals the plain-C restatement bit for bit; plus the size-independent properties."""
    points, nodes = pkg.synth.make_partition_inputs(shape, num_nodes, ppn, seed=pkg.synth.SEED + 17)
    p2n, sizes, masks, idx, kmask, kpts = _gpu_partition(pkg, points, nodes, K)
    c_p2n, c_sizes, c_idx, c_kmask = c_partition(knn_c, points, nodes, K)
    N = points.shape[0]
    assert np.array_equal(p2n.numpy(), c_p2n) and np.array_equal(sizes.numpy(), c_sizes)
    assert np.array_equal(masks.numpy(), c_sizes > 0)
    assert np.array_equal(idx.numpy(), c_idx) and np.array_equal(kmask.numpy(), c_kmask)
    _check_patch_points(points, idx, kpts)
    assert int(sizes.sum()) == N
    assert torch.equal(kmask.sum(1), sizes.clamp(max=K))
    assert bool((idx[~kmask] == N).all()) and bool((idx[kmask] < N).all())
    rows = torch.arange(num_nodes)[:, None].expand(-1, K)
    assert torch.equal(p2n[idx[kmask]], rows[kmask])
    # ascending own distances inside every patch: exact in fp64, up to the rounding of the reference's fp32
    # x2 - 2xy + y2 (a few ulp of |point|^2 + |node|^2, pairwise_distance.py:27-29)
    d = (points.double()[idx.clamp(max=N - 1)] - nodes.double()[:, None, :]).pow(2).sum(-1)
    both = kmask[:, 1:] & kmask[:, :-1]
    slack = 1e-6 * float(points.pow(2).sum(-1).max() + nodes.pow(2).sum(-1).max())
    assert bool((d[:, 1:][both] >= d[:, :-1][both] - slack).all())


def test_partition_edge_cases(pkg, knn_c):
    g = torch.Generator().manual_seed(3)
    for N, M, K in ((1, 1, 1), (5, 1, 5), (7, 9, 3), (2049, 2, 1), (33, 2500, 2)):
        points, nodes = torch.rand(N, 3, generator=g), torch.rand(M, 3, generator=g)
        p2n, sizes, masks, idx, kmask, kpts = _gpu_partition(pkg, points, nodes, K)
        c_p2n, c_sizes, c_idx, c_kmask = c_partition(knn_c, points, nodes, K)
        assert np.array_equal(p2n.numpy(), c_p2n) and np.array_equal(idx.numpy(), c_idx), (N, M, K)
        assert np.array_equal(kmask.numpy(), c_kmask) and np.array_equal(sizes.numpy(), c_sizes)
    # exact duplicates: equal distances are ordered by index, a point equidistant from two nodes goes to the first
    points = torch.tensor([[0.5, 0, 0], [0.25, 0, 0], [0.25, 0, 0], [0.5, 0, 0]])
    nodes = torch.tensor([[0., 0, 0], [1., 0, 0]])
    p2n, sizes, masks, idx, kmask, _ = _gpu_partition(pkg, points, nodes, 4)
    assert p2n.tolist() == [0, 0, 0, 0] and idx[0].tolist() == [1, 2, 0, 3] and idx[1].tolist() == [4] * 4
    assert masks.tolist() == [True, False]
    with pytest.raises(RuntimeError, match='out of range'):
        pkg.ops.point_to_node_partition(points.to(DEV), nodes.to(DEV), 5)
    with pytest.raises(RuntimeError):
        pkg.ops.point_to_node_partition(points, nodes, 2)  # CPU tensors: no fallback


def _close_transport(got, want):
    finite = want > -1e11
    assert torch.equal(finite, got > -1e11)
    diff = (got[finite] - want[finite]).abs()
    assert bool((diff <= 1e-4 + 1e-4 * want[finite].abs()).all()), diff.max()
    if (~finite).any():
        assert torch.allclose(got[~finite], want[~finite], rtol=1e-6, atol=0)
    return float(diff.max()) if diff.numel() else 0.0


@pytest.mark.parametrize('tag', LOT_CASES)
def test_transport_vs_reference_golden(pkg, golden, tag):
    z = golden('next_cases.npz')
    scores = torch.from_numpy(z[f'{tag}.scores'])
    rm = torch.from_numpy(z[f'{tag}.row_masks']).to(DEV) if f'{tag}.row_masks' in z.files else None
    cm = torch.from_numpy(z[f'{tag}.col_masks']).to(DEV) if f'{tag}.col_masks' in z.files else None
    mod = pkg.LearnableLogOptimalTransport(int(z[f'{tag}.iters'])).to(DEV)
    with torch.no_grad():
        mod.alpha.fill_(float(z[f'{tag}.alpha']))
        got = mod(scores.to(DEV), rm, cm)
    assert got.dtype == torch.float32
    _close_transport(got.cpu(), torch.from_numpy(z[f'{tag}.out']))


@pytest.mark.parametrize('P,rows,cols', [(256, 64, 64), (256, 128, 128), (64, 64, 50), (3, 200, 180)])
def test_transport_full_size_vs_oracle(pkg, oracle, P, rows, cols):
    """The BASELINE shapes (256 correspondences x 64 / 128 patch points, 100 iterations; config.py:93-94,100), a ragged
    one and one near the shared-memory limit, against the torch oracle; plus the marginals of the transport plan."""
    scores, rm, cm = pkg.synth.make_patch_scores(P, rows, cols, 64, seed=99)
    alpha = torch.tensor(0.8)
    got = pkg.ops.log_optimal_transport(scores.to(DEV), alpha.to(DEV), 100, rm.to(DEV), cm.to(DEV)).cpu()
    want = oracle.log_optimal_transport(scores, alpha, 100, rm, cm)
    _close_transport(got, want)
    norm = -torch.log(rm.float().sum(1) + cm.float().sum(1))
    plan = torch.exp(got.double() + norm[:, None, None].double())
    mass = torch.exp(norm.double())[:, None].expand(-1, cols).clone()
    mass[~cm] = 0.0
    assert torch.allclose(plan.sum(1)[:, :-1], mass, rtol=2e-4, atol=1e-9)  # columns are exact after the last half-step
    assert torch.allclose(plan.sum(1)[:, -1], rm.double().sum(1) * torch.exp(norm.double()), rtol=2e-4)


def test_transport_edge_cases(pkg, oracle):
    scores, rm, cm = pkg.synth.make_patch_scores(4, 6, 5, 8, seed=1)
    alpha = torch.tensor(1.0)
    for iters in (0, 1):
        got = pkg.ops.log_optimal_transport(scores.to(DEV), alpha.to(DEV), iters, rm.to(DEV), cm.to(DEV)).cpu()
        _close_transport(got, oracle.log_optimal_transport(scores, alpha, iters, rm, cm))
    one = torch.randn(2, 1, 1)
    got = pkg.ops.log_optimal_transport(one.to(DEV), alpha.to(DEV), 10).cpu()
    _close_transport(got, oracle.log_optimal_transport(one, alpha, 10))
    mod = pkg.LearnableLogOptimalTransport(3).to(DEV)
    with pytest.raises(RuntimeError, match='forward-only'):
        mod(scores.to(DEV))  # alpha requires grad and grad mode is on
    with pytest.raises(RuntimeError, match='shared memory'):
        pkg.ops.log_optimal_transport(torch.zeros(1, 400, 400, device=DEV), alpha.to(DEV), 1)
