"""CPU tests: the oracle restatements of SURVEY.md 8(f)-2 (point_to_node_partition) and 8(f)-3
(LearnableLogOptimalTransport) are pinned against golden vectors written by the UNMODIFIED reference functions
(tests/golden/make_golden_next.py) and, when /root/reference is present, against the live reference."""
import ctypes

import numpy as np
import pytest
import torch

from load_reference import load_reference, reference_available

P2N_CASES = ['p2n_small', 'p2n_empty', 'p2n_all', 'p2n_big', 'p2n_3dmatch', 'p2n_kitti']
LOT_CASES = ['lot_small', 'lot_nomask', 'lot_3dmatch', 'lot_kitti', 'lot_alpha']


def _p2n_inputs(z, tag):
    return torch.from_numpy(z[f'{tag}.points']), torch.from_numpy(z[f'{tag}.nodes']), int(z[f'{tag}.K'])


def c_partition(knn_c, points, nodes, K):
    N, M = points.shape[0], nodes.shape[0]
    p2n = np.empty(N, np.int64)
    sizes = np.empty(M, np.int64)
    idx = np.empty((M, K), np.int64)
    msk = np.empty((M, K), np.uint8)
    scratch = np.empty(N, np.float32)
    P, Q = np.ascontiguousarray(points.numpy()), np.ascontiguousarray(nodes.numpy())
    ptr = lambda a: a.ctypes.data_as(ctypes.c_void_p)  # noqa: E731
    knn_c.geotx_oracle_point_to_node.restype = None
    knn_c.geotx_oracle_point_to_node(ptr(P), ctypes.c_int(N), ptr(Q), ctypes.c_int(M), ctypes.c_int(K), ptr(p2n),
                                     ptr(sizes), ptr(idx), ptr(msk), ptr(scratch))
    return p2n, sizes, idx, msk.astype(bool)


@pytest.mark.parametrize('tag', P2N_CASES)
def test_partition_oracles_vs_reference_golden(oracle, golden, knn_c, tag):
    z = golden('next_cases.npz')
    points, nodes, K = _p2n_inputs(z, tag)
    assert oracle.partition_is_tie_free(points, nodes)
    p2n, sizes, masks, idx, kmask = oracle.point_to_node_partition(points, nodes, K, return_count=True)
    c_p2n, c_sizes, c_idx, c_kmask = c_partition(knn_c, points, nodes, K)
    for got in ((p2n.numpy(), sizes.numpy(), masks.numpy(), idx.numpy(), kmask.numpy()),
                (c_p2n, c_sizes, c_sizes > 0, c_idx, c_kmask)):
        assert np.array_equal(got[0], z[f'{tag}.point_to_node'])
        assert np.array_equal(got[1], z[f'{tag}.node_sizes'])
        assert np.array_equal(got[2], z[f'{tag}.node_masks'])
        assert np.array_equal(got[3], z[f'{tag}.node_knn_indices'])
        assert np.array_equal(got[4], z[f'{tag}.node_knn_masks'])


def test_partition_distance_recipe_is_bit_exact(oracle, golden, knn_c):
    """The C restatement's fp32 operation order (fma chain for xy, x = node, y = point) reproduces torch-CPU's
    pairwise_distance(nodes, points) bit for bit, also for a non-square (M,3)x(3,N) product."""
    z = golden('next_cases.npz')
    points, nodes, _ = _p2n_inputs(z, 'p2n_small')
    want = oracle.pairwise_sqdist(nodes, points).numpy()
    knn_c.geotx_oracle_sqdist.restype = ctypes.c_float
    P, Q = points.numpy(), nodes.numpy()
    got = np.empty_like(want)
    for m in range(Q.shape[0]):
        for p in range(P.shape[0]):
            got[m, p] = knn_c.geotx_oracle_sqdist(Q[m].ctypes.data_as(ctypes.c_void_p), P[p].ctypes.data_as(ctypes.c_void_p))
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_partition_properties_of_golden(golden):
    """Size-independent properties the GPU tests reuse at full size: sizes sum to N; a node's list is its own points by
    ascending distance; the sentinel fills exactly the slots beyond min(size, K)."""
    z = golden('next_cases.npz')
    for tag in P2N_CASES:
        N, K = z[f'{tag}.points'].shape[0], int(z[f'{tag}.K'])
        sizes, idx, msk = z[f'{tag}.node_sizes'], z[f'{tag}.node_knn_indices'], z[f'{tag}.node_knn_masks']
        assert sizes.sum() == N
        assert np.array_equal(msk.sum(1), np.minimum(sizes, K))
        assert np.all(idx[~msk] == N) and np.all(idx[msk] < N)
        owner = z[f'{tag}.point_to_node']
        rows = np.repeat(np.arange(idx.shape[0]), K).reshape(idx.shape)
        assert np.array_equal(owner[idx[msk]], rows[msk])


@pytest.mark.parametrize('tag', LOT_CASES)
def test_transport_oracle_vs_reference_golden(oracle, golden, tag):
    z = golden('next_cases.npz')
    exact = torch.__version__ in str(z['meta'])
    scores = torch.from_numpy(z[f'{tag}.scores'])
    rm = torch.from_numpy(z[f'{tag}.row_masks']) if f'{tag}.row_masks' in z.files else None
    cm = torch.from_numpy(z[f'{tag}.col_masks']) if f'{tag}.col_masks' in z.files else None
    got = oracle.log_optimal_transport(scores, float(z[f'{tag}.alpha']), int(z[f'{tag}.iters']), rm, cm)
    want = torch.from_numpy(z[f'{tag}.out'])
    assert got.shape == want.shape
    if exact:
        assert torch.allclose(got, want, rtol=1e-6, atol=1e-6)  # same ops; the CPU's vector width may differ
    else:
        assert torch.allclose(got, want, rtol=1e-4, atol=1e-4)
    # transport properties (size independent): rows and columns of exp(out + norm) carry the prescribed masses
    valid_r = rm.float().sum(1) if rm is not None else torch.full((scores.shape[0],), float(scores.shape[1]))
    valid_c = cm.float().sum(1) if cm is not None else torch.full((scores.shape[0],), float(scores.shape[2]))
    norm = -torch.log(valid_r + valid_c)
    plan = torch.exp(want.double() + norm[:, None, None].double())
    col_mass = plan.sum(1)  # the last half-iteration fixes the columns exactly
    want_mass = torch.exp(norm.double())[:, None].expand(-1, scores.shape[2]).clone()
    if cm is not None:
        want_mass[~cm] = 0.0
    assert torch.allclose(col_mass[:, :-1], want_mass, rtol=1e-4, atol=1e-9)


@pytest.mark.skipif(not reference_available(), reason='/root/reference not present (GPU box)')
def test_next_oracles_equal_live_reference(oracle, pkg):
    load_reference()
    from geotransformer.modules.ops.pointcloud_partition import point_to_node_partition
    from geotransformer.modules.sinkhorn.learnable_sinkhorn import LearnableLogOptimalTransport
    saved = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self  # the reference calls .cuda() on fresh helper tensors
    try:
        with torch.no_grad():
            pts, nds = pkg.synth.make_partition_inputs('3dmatch', 96, 12, seed=5)
            assert oracle.partition_is_tie_free(pts, nds)
            want = point_to_node_partition(pts, nds, 10, return_count=True)
            got = oracle.point_to_node_partition(pts, nds, 10, return_count=True)
            assert all(torch.equal(a, b) for a, b in zip(got, want))
            scores, rm, cm = pkg.synth.make_patch_scores(7, 20, 17, 32, seed=6)
            mod = LearnableLogOptimalTransport(30)
            assert torch.equal(oracle.log_optimal_transport(scores, mod.alpha.detach(), 30, rm, cm), mod(scores, rm, cm))
    finally:
        torch.Tensor.cuda = saved


def test_next_host_api_without_gpu(pkg):
    """State-dict key of the transport module, error behaviour, no CPU fallback."""
    mod = pkg.LearnableLogOptimalTransport(100)
    assert list(mod.state_dict().keys()) == ['alpha'] and mod.alpha.shape == () and float(mod.alpha.detach()) == 1.0
    assert repr(mod) == 'LearnableLogOptimalTransport(num_iterations=100)'
    with pytest.raises(RuntimeError):
        with torch.no_grad():
            mod(torch.zeros(1, 4, 4))  # CPU tensor: no fallback
    with pytest.raises(RuntimeError):
        pkg.point_to_node_partition(torch.zeros(10, 3), torch.zeros(2, 3), 4)
    lib = pkg._abi.load()
    assert lib.geotx_p2n_workspace_bytes(100, 10, 0) == 0
    assert lib.geotx_p2n_workspace_bytes(100, 10, 8193) == 0
    assert lib.geotx_p2n_workspace_bytes(100, 10, 8) > 0
    p = ctypes.c_void_p(256)
    # point_limit > number of points: torch.topk's error in the reference, GEOTX_E_ARG here (checked before any launch)
    assert lib.geotx_point_to_node_partition(p, 10, p, 2, 11, p, None, p, p, p, None, p, None) == -1
    assert b'out of range' in lib.geotx_last_error()
    assert lib.geotx_log_optimal_transport(p, None, None, p, 100, 1e12, 1, 400, 400, p, None) == -1
    assert b'shared memory' in lib.geotx_last_error()
