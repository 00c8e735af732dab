"""-m gpu: the allocation-free entry point geotx_point_to_node_partition_static (include/geotx_b200.h), called through
the C ABI and compared with geotx_point_to_node_partition on the same inputs.  Both run the same kernels over the same
arithmetic, and the partition is index / byte work, so the tolerance is the one of tests/test_gpu_next.py: bit-exact.
Two calls in a row with DIFFERENT inputs: the built-in scratch array keeps the previous call's contents, so the second
call proves that what must start at zero is cleared.  Collected last (file name) so that a failure here cannot hide
other tests under `pytest -x`; passes alone in a fresh process and after the other tests."""
import ctypes

import pytest
import torch

from gpu_common import DEV

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _outputs(N, M, K):
    return dict(p2n=torch.full((N,), -7, dtype=torch.int64, device=DEV),
                sizes=torch.full((M,), -7, dtype=torch.int64, device=DEV),
                masks=torch.full((M,), 3, dtype=torch.uint8, device=DEV),
                idx=torch.full((M, K), -7, dtype=torch.int64, device=DEV),
                kmask=torch.full((M, K), 3, dtype=torch.uint8, device=DEV),
                kpts=torch.full((M, K, 3), float('nan'), dtype=torch.float32, device=DEV))


def _call_both(pkg, points, nodes, K):
    abi = pkg._abi
    lib = abi.load()
    points, nodes = points.to(DEV).contiguous(), nodes.to(DEV).contiguous()
    N, M = points.shape[0], nodes.shape[0]
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    a, b = _outputs(N, M, K), _outputs(N, M, K)
    work = torch.empty(lib.geotx_p2n_workspace_bytes(N, M, K), dtype=torch.uint8, device=DEV)
    abi.call('geotx_point_to_node_partition', _ptr(points), N, _ptr(nodes), M, K, _ptr(a['p2n']), _ptr(a['sizes']),
             _ptr(a['masks']), _ptr(a['idx']), _ptr(a['kmask']), _ptr(a['kpts']), _ptr(work), stream)
    abi.call('geotx_point_to_node_partition_static', _ptr(points), N, _ptr(nodes), M, K, _ptr(b['p2n']),
             _ptr(b['sizes']), _ptr(b['masks']), _ptr(b['idx']), _ptr(b['kmask']), _ptr(b['kpts']), stream)
    torch.cuda.synchronize()
    return a, b


def test_static_scratch_matches_workspace_entry_point_twice(pkg):
    lib = pkg._abi.load()
    assert lib.geotx_p2n_static_scratch_bytes() == 2 * 1024 * 1024
    # two calls in a row, different sizes and data (the second is smaller: stale histogram entries would show)
    cases = [('3dmatch', 1024, 24, 64, 17), ('kitti', 300, 40, 128, 18)]
    for shape, num_nodes, ppn, K, seed in cases:
        points, nodes = pkg.synth.make_partition_inputs(shape, num_nodes, ppn, seed=pkg.synth.SEED + seed)
        a, b = _call_both(pkg, points, nodes, K)
        for name in a:
            assert torch.equal(a[name], b[name], ), f'{shape}: {name} differs between the two entry points'
        assert int(b['sizes'].sum()) == points.shape[0]
        assert not torch.isnan(b['kpts']).any()
    # and the first case again: same result as before on a scratch array that now holds the second case's data
    shape, num_nodes, ppn, K, seed = cases[0]
    points, nodes = pkg.synth.make_partition_inputs(shape, num_nodes, ppn, seed=pkg.synth.SEED + seed)
    a, b = _call_both(pkg, points, nodes, K)
    for name in a:
        assert torch.equal(a[name], b[name])


def test_static_scratch_rejects_sizes_beyond_the_array(pkg):
    lib = pkg._abi.load()
    N, M, K = 200_000, 1024, 64  # workspace 4 MB > 2 MiB
    assert lib.geotx_p2n_workspace_bytes(N, M, K) > lib.geotx_p2n_static_scratch_bytes()
    points = torch.zeros(N, 3, device=DEV)
    nodes = torch.zeros(M, 3, device=DEV)
    o = _outputs(N, M, K)
    rc = lib.geotx_point_to_node_partition_static(_ptr(points), N, _ptr(nodes), M, K, _ptr(o['p2n']), _ptr(o['sizes']),
                                                  _ptr(o['masks']), _ptr(o['idx']), _ptr(o['kmask']), _ptr(o['kpts']),
                                                  ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -1  # GEOTX_E_ARG
    msg = lib.geotx_last_error().decode()
    assert 'scratch' in msg and str(lib.geotx_p2n_static_scratch_bytes()) in msg
    torch.cuda.synchronize()
    assert int(o['p2n'][0]) == -7  # nothing was launched
