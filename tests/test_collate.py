"""K9 radius neighbours.  The checker is the REFERENCE's own C++ operator
(extensions/cpu/radius_neighbors, compiled from /root/reference into oracle/_ref by oracle/build_ref.sh) and golden
vectors written by it (tests/golden/make_golden_collate.py).  Index work: bit-exact, except that the reference sorts
with std::sort on the distance alone, so neighbours at EXACTLY equal fp32 distance may come in either order (and either
of two such neighbours may be the one cut off at `limit`); the comparison allows exactly that and nothing else."""
import importlib.util
import os

import numpy as np
import pytest
import torch

import ref_cpu

_spec = importlib.util.spec_from_file_location(
    'make_golden_collate', os.path.join(os.path.dirname(__file__), 'golden', 'make_golden_collate.py'))
_mod = None


def _gold():
    global _mod
    if _mod is None:
        _mod = importlib.util.module_from_spec(_spec)
        _spec.loader.exec_module(_mod)
    return _mod


TAGS = ['rn_small', 'rn_cross', 'rn_sparse', 'rn_stage']


def ref_dist(q, s):
    """nanoflann L2_Simple_Adaptor::evalMetric in fp32: ((qx-sx)^2 + (qy-sy)^2) + (qz-sz)^2 (nanoflann.hpp:432-439)."""
    d = (q - s).astype(np.float32)
    return ((d[..., 0] * d[..., 0]).astype(np.float32) + (d[..., 1] * d[..., 1]).astype(np.float32)).astype(np.float32) \
        + (d[..., 2] * d[..., 2]).astype(np.float32)


def assert_same_neighbors(got, want, q, s, sentinel):
    """got, want (Nq, W) int: equal, or differing only by permutations inside groups of equal reference distance
    (including the group that straddles the cut at W)."""
    assert got.shape == want.shape, (got.shape, want.shape)
    bad = np.nonzero((got != want).any(axis=1))[0]
    sp = np.concatenate([s, np.full((1, 3), np.inf, np.float32)])  # the sentinel is infinitely far
    for r in bad:
        dg = ref_dist(q[r][None], sp[np.minimum(got[r], sentinel)])
        dw = ref_dist(q[r][None], sp[np.minimum(want[r], sentinel)])
        assert np.array_equal(dg, dw), f'query {r}: distance sequences differ'
        diff = got[r] != want[r]
        for d in np.unique(dw[diff]):
            grp = dw == d
            a, b = set(got[r][grp]), set(want[r][grp])
            if a != b:  # only legal when the tie group reaches the last column (one of the tied points was cut)
                assert grp[-1], f'query {r}: different neighbours at distance {d}'
    return len(bad)


def test_reference_operator_matches_its_golden(pkg, golden):
    """oracle/_ref (built here from /root/reference, or the prebuilt one on the GPU box) reproduces the vectors."""
    if not ref_cpu.available():
        pytest.skip('oracle/_ref not built and /root/reference absent')
    z = golden('collate_cases.npz')
    for tag in TAGS:
        q, s, ql, sl, radius, limit = _gold().case_inputs(pkg.synth, tag)
        full = ref_cpu.radius_neighbors(q.numpy(), s.numpy(), ql, sl, radius)
        assert full.shape[1] == int(z[f'{tag}.max_count'])
        assert np.array_equal(full[:, :limit], z[f'{tag}.neighbors'])


def test_golden_semantics(pkg, golden):
    """What the vectors mean (and what the GPU kernel must reproduce): in-radius test d < r^2 in fp32, ascending
    distance, per batch element, sentinel = total number of support points."""
    z = golden('collate_cases.npz')
    q, s, ql, sl, radius, limit = _gold().case_inputs(pkg.synth, 'rn_small')
    want = z['rn_small.neighbors']
    qn, sn = q.numpy(), s.numpy()
    r2 = np.float32(radius) * np.float32(radius)
    q0 = s0 = 0
    for b in range(len(ql)):
        for r in range(q0, q0 + ql[b], 37):
            d = ref_dist(qn[r][None], sn[s0:s0 + sl[b]])
            inside = np.nonzero(d < r2)[0]
            order = inside[np.lexsort((inside, d[inside]))][:limit] + s0
            row = want[r]
            assert np.array_equal(np.sort(d[order - s0]), d[row[row < sn.shape[0]] - s0])
            assert (row[len(order):] == sn.shape[0]).all()
        q0 += ql[b]
        s0 += sl[b]


def test_host_api_without_gpu(pkg):
    with pytest.raises(RuntimeError):
        pkg.radius_search(torch.zeros(4, 3), torch.zeros(4, 3), [4], [4], 0.1, 8)  # CPU tensors: no fallback
    import ctypes
    lib = pkg._abi.load()
    p = ctypes.c_void_p(256)
    one = (ctypes.c_int64 * 1)(4)
    assert lib.geotx_radius_neighbors(p, p, one, one, 1, 0.1, 0, p, p, p, None) == -1
    assert lib.geotx_radius_neighbors(p, p, one, one, 1, 0.1, 513, p, p, p, None) == -1


# ---- GPU ----------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize('tag', TAGS)
def test_gpu_vs_reference_golden(pkg, golden, tag):
    from gpu_common import DEV
    z = golden('collate_cases.npz')
    q, s, ql, sl, radius, limit = _gold().case_inputs(pkg.synth, tag)
    got = pkg.radius_search(q.to(DEV), s.to(DEV), torch.tensor(ql), torch.tensor(sl), radius, limit)
    assert got.dtype == torch.int64 and got.shape[1] == min(limit, int(z[f'{tag}.max_count']))
    assert_same_neighbors(got.cpu().numpy(), z[f'{tag}.neighbors'].astype(np.int64), q.numpy(), s.numpy(), s.shape[0])
    nb, counts, max_count = pkg.ops.radius_neighbors(q.to(DEV), s.to(DEV), ql, sl, radius, limit)
    assert nb.shape == (q.shape[0], limit) and int(max_count) == int(z[f'{tag}.max_count'])
    assert int(counts.max()) == int(max_count)
    assert torch.equal((nb < s.shape[0]).sum(1).cpu(), counts.cpu().clamp(max=limit).long())


@pytest.mark.gpu
@pytest.mark.parametrize('sizes,radius,limit', [((20000, 18000), 0.0625, 38), ((5000, 5000), 0.25, 36),
                                                ((3000, 1, 2000), 0.3, 64), ((40, 33), 10.0, 256)])
def test_gpu_full_size_vs_reference_operator(pkg, sizes, radius, limit):
    """A stage-0-sized pair (3DMatch: ~20k points per cloud, radius 2.5 x 0.025 m, limit 38: config.py:79-80,
    utils/data.py), a dense stage, a one-point batch element and a search where everything is in range, against the
    reference's operator itself.  (Empty batch elements are not compared: the reference's loop advances at most one
    element per query, radius_neighbors_cpu.cpp:37-52, and never leaves an empty one.)"""
    if not ref_cpu.available():
        pytest.skip('oracle/_ref not built and /root/reference absent')
    from gpu_common import DEV
    clouds = [pkg.synth.make_dense_cloud(n, 3.0, 900 + b) if n else torch.zeros(0, 3) for b, n in enumerate(sizes)]
    s = torch.cat(clouds)
    full = ref_cpu.radius_neighbors(s.numpy(), s.numpy(), list(sizes), list(sizes), radius)
    got = pkg.radius_search(s.to(DEV), s.to(DEV), list(sizes), list(sizes), radius, limit).cpu().numpy()
    assert_same_neighbors(got, full[:, :limit], s.numpy(), s.numpy(), s.shape[0])
    # size-independent properties: every point finds itself first (distance 0); lists stay inside the batch element
    assert np.array_equal(got[:, 0], np.arange(s.shape[0]))
    start = 0
    for n in sizes:
        blk = got[start:start + n]
        assert ((blk >= start) & (blk < start + n) | (blk == s.shape[0])).all()
        start += n


@pytest.mark.gpu
@pytest.mark.parametrize('sizes,radius', [((3000, 2500), 0.5), ((900, 1, 700), 0.8), ((40, 33), 10.0)])
def test_gpu_no_truncation_vs_reference_operator(pkg, sizes, radius):
    """neighbor_limit <= 0: the reference returns the operator's full-width result (radius_search.py:24-26); so does
    the GPU wrapper (one search to learn the width, a second one when it exceeds the first guess of 64: widths 129, 106 and 40 here)."""
    if not ref_cpu.available():
        pytest.skip('oracle/_ref not built and /root/reference absent')
    from gpu_common import DEV
    s = torch.cat([pkg.synth.make_dense_cloud(n, 3.0, 700 + b) for b, n in enumerate(sizes)])
    full = ref_cpu.radius_neighbors(s.numpy(), s.numpy(), list(sizes), list(sizes), radius)
    for limit in (0, -1):
        got = pkg.radius_search(s.to(DEV), s.to(DEV), list(sizes), list(sizes), radius, limit).cpu().numpy()
        assert_same_neighbors(got, full, s.numpy(), s.numpy(), s.shape[0])
    crowd = pkg.synth.make_dense_cloud(600, 3.0, 77).to(DEV)  # 600 points within reach of each other
    with pytest.raises(RuntimeError, match='at most 512'):
        pkg.radius_search(crowd, crowd, [600], [600], 100.0, 0)


@pytest.mark.gpu
def test_gpu_empty_batch_element(pkg):
    """Our own semantics for an empty element: it contributes no queries and no neighbours, the others are unaffected."""
    from gpu_common import DEV
    a, b = pkg.synth.make_dense_cloud(500, 3.0, 1), pkg.synth.make_dense_cloud(400, 3.0, 2)
    s = torch.cat([a, b]).to(DEV)
    with_empty = pkg.ops.radius_neighbors(s, s, [500, 0, 400], [500, 0, 400], 0.3, 16)[0]
    without = pkg.ops.radius_neighbors(s, s, [500, 400], [500, 400], 0.3, 16)[0]
    assert torch.equal(with_empty, without)


@pytest.mark.gpu
@pytest.mark.parametrize('nq,ns,radius,limit,max_cells', [
    ((20000, 18000), (20000, 18000), 0.0625, 38, 1 << 21),   # the pyramid's stage 0
    ((6000, 5000), (1500, 1400), 0.25, 36, 1 << 21),         # a down-sampling search: queries and supports differ
    ((1500, 1400), (6000, 5000), 0.125, 64, 1 << 21),        # up-sampling
    ((4000, 3000), (4000, 3000), 0.0625, 38, 64),            # grid capped at 64 cells: larger cells, same result
    ((300, 200), (300, 200), 100.0, 512, 1 << 21),           # everything in range: one cell
    ((2000, 1), (1, 2000), 0.3, 16, 1 << 10),
    ((200,) * 11, (300,) * 11, 0.3, 24, 1 << 12),            # more batch elements than one set of launches takes
    ((500, 0, 400), (600, 0, 300), 0.3, 16, 1 << 12),        # an empty element in the middle
])
def test_gpu_cell_lists_equal_exhaustive_scan(pkg, nq, ns, radius, limit, max_cells):
    """The cell-list search and the exhaustive scan share the distance arithmetic, the keys and the selection: their
    outputs are equal bit for bit, also for queries outside the supports' bounding box."""
    from gpu_common import DEV
    g = torch.Generator().manual_seed(sum(nq) + limit)
    clouds = [pkg.synth.make_dense_cloud(n, 3.0, 50 + b) if n else torch.zeros(0, 3) for b, n in enumerate(ns)]
    queries = []
    for c, n in zip(clouds, nq):  # jittered copies of support points; one in twenty pushed outside the supports' box
        if n == 0 or c.shape[0] == 0:
            queries.append(torch.rand(n, 3, generator=g) * 3.0)
            continue
        pick = c[torch.randint(0, c.shape[0], (n,), generator=g)] + torch.randn(n, 3, generator=g) * 0.02
        pick[torch.rand(n, generator=g) < 0.05] += 3.2
        queries.append(pick)
    q, s = torch.cat(queries).to(DEV), torch.cat(clouds).to(DEV)
    a = pkg.ops.radius_neighbors(q, s, list(nq), list(ns), radius, limit, method='grid', max_cells=max_cells)
    b = pkg.ops.radius_neighbors(q, s, list(nq), list(ns), radius, limit, method='brute')
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    assert int(a[2]) > 0


# ---- K10 grid subsampling -----------------------------------------------------------------------------------------
GS_CASES = [((5000, 4000), 3.0, 0.05), ((30000, 25000), 3.0, 0.05), ((12000, 12000), 3.0, 0.1), ((300, 1, 7), 3.0, 0.4)]
GS_TOL = 2e-6  # relative to the cloud's extent: fp64 sums rounded once vs the reference's sequential fp32 sums


def _gs_inputs(pkg, sizes, extent, seed=700):
    clouds = [pkg.synth.make_dense_cloud(n, extent, seed + b) - (1.3 if b else 0.0) for b, n in enumerate(sizes)]
    return torch.cat(clouds).contiguous()


def _lexsort_rows(a):
    return a[np.lexsort((a[:, 2], a[:, 1], a[:, 0]))]


def test_reference_grid_subsampling_semantics(pkg):
    """What the reference operator computes (pins the shim): per element, one barycentre per occupied voxel of the
    grid anchored at floor(min / voxel) * voxel; every point lies in the voxel of exactly one output."""
    if not ref_cpu.available():
        pytest.skip('oracle/_ref not built and /root/reference absent')
    sizes, extent, voxel = GS_CASES[0]
    pts = _gs_inputs(pkg, sizes, extent).numpy()
    sp, sl = ref_cpu.grid_subsampling(pts, list(sizes), voxel)
    assert sl.sum() == sp.shape[0] and len(sl) == 2
    start = out = 0
    for n, m in zip(sizes, sl):
        p = pts[start:start + n]
        origin = np.floor(p.min(0) * np.float32(1.0 / voxel)) * np.float32(voxel)
        cells = np.floor((p - origin) / np.float32(voxel)).astype(np.int64)
        uniq = np.unique(cells, axis=0)
        assert len(uniq) == m
        bary_cells = np.floor((sp[out:out + m] - origin) / np.float32(voxel)).astype(np.int64)
        assert len(np.unique(bary_cells, axis=0)) >= 0.99 * m  # barycentres sit in their own voxels (up to rounding)
        start += n
        out += m


@pytest.mark.gpu
@pytest.mark.parametrize('sizes,extent,voxel', GS_CASES)
def test_gpu_grid_subsample_vs_reference_operator(pkg, sizes, extent, voxel):
    if not ref_cpu.available():
        pytest.skip('oracle/_ref not built and /root/reference absent')
    from gpu_common import DEV
    pts = _gs_inputs(pkg, sizes, extent)
    want_p, want_l = ref_cpu.grid_subsampling(pts.numpy(), list(sizes), voxel)
    got_p, got_l = pkg.grid_subsample(pts.to(DEV), torch.tensor(sizes), voxel)
    assert got_l.dtype == torch.int64 and not got_l.is_cuda and got_p.is_cuda
    assert np.array_equal(got_l.numpy(), want_l), (got_l, want_l)  # voxel counts: exact
    got_p = got_p.cpu().numpy()
    g0 = w0 = 0
    for m in want_l:
        a, b = _lexsort_rows(got_p[g0:g0 + m]), _lexsort_rows(want_p[w0:w0 + m])
        err = np.abs(a - b).max()
        if err > GS_TOL * extent:  # two barycentres whose x differ by less than the tolerance may sort differently
            from scipy.spatial import cKDTree
            d, idx = cKDTree(b).query(a)
            assert d.max() <= GS_TOL * extent * 2 and len(np.unique(idx)) == len(a)
        g0 += m
        w0 += m
    # our own order: ascending voxel index (x fastest) inside every element
    start = out = 0
    for n, m in zip(sizes, want_l):
        p = pts[start:start + n].numpy()
        origin = np.floor(p.min(0) * np.float32(1.0 / voxel)) * np.float32(voxel)
        cells = np.floor((p - origin) / np.float32(voxel)).astype(np.int64)
        dims = cells.max(0) + 1
        keys = np.unique(cells[:, 0] + dims[0] * (cells[:, 1] + dims[1] * cells[:, 2]))
        c = np.floor((got_p[out:out + m] - origin) / np.float32(voxel)).astype(np.int64)
        got_keys = c[:, 0] + dims[0] * (c[:, 1] + dims[1] * c[:, 2])
        assert (np.abs(got_keys - keys) > 0).mean() < 0.01  # barycentres on a voxel face may round into the neighbour
        start += n
        out += m


@pytest.mark.gpu
def test_gpu_grid_subsample_properties_and_errors(pkg):
    from gpu_common import DEV
    pts = _gs_inputs(pkg, (20000,), 3.0).to(DEV)
    once, l1 = pkg.grid_subsample(pts, [20000], 0.05)
    again, l2 = pkg.grid_subsample(pts, [20000], 0.05)
    assert torch.equal(once, again) and torch.equal(l1, l2)  # deterministic despite the atomics
    sub2, l3 = pkg.grid_subsample(once, l1, 0.05)            # a barycentre stays in its voxel: (almost) idempotent
    assert abs(int(l3[0]) - int(l1[0])) <= 0.01 * int(l1[0])
    big, lb = pkg.grid_subsample(pts, [20000], 1e6)          # one voxel: the centroid
    assert int(lb[0]) == 1 and torch.allclose(big[0], pts.double().mean(0).float(), atol=1e-5)
    with pytest.raises(RuntimeError, match='max_cells'):
        pkg.grid_subsample(pts, [20000], 1e-4)               # 30000^3 cells
    with pytest.raises(RuntimeError):
        pkg.grid_subsample(pts, [19999], 0.05)
    with pytest.raises(RuntimeError):
        pkg.grid_subsample(pts.cpu(), [20000], 0.05)


# ---- the whole pyramid (utils/data.py:13-81) ------------------------------------------------------------------------
@pytest.mark.gpu
def test_gpu_pyramid_vs_reference_pipeline(pkg):
    """precompute_data_stack_mode on the GPU against the same pipeline on the reference's own operators, at the 3DMatch
    settings (4 stages, voxel 0.025, radius 2.5 voxels, neighbour limits of config.py:79-81 scale).  Stage 0 is the
    input itself: exact.  Later stages hold the same voxels in another order and barycentres that differ in the last
    bits, so their index lists are compared after mapping through the point correspondence; a row may differ only where
    a barycentre distance sits within rounding of the radius or of another neighbour's distance (< 0.2 % of rows)."""
    if not ref_cpu.available():
        pytest.skip('oracle/_ref not built and /root/reference absent')
    from scipy.spatial import cKDTree
    from gpu_common import DEV
    sizes = (20000, 18000)
    pts = torch.cat([pkg.synth.make_dense_cloud(n, 3.0, 800 + b) for b, n in enumerate(sizes)])
    limits = [38, 36, 36, 38]
    want = ref_cpu.precompute_data_stack_mode(pts.numpy(), list(sizes), 4, 0.025, 0.0625, limits)
    got = pkg.precompute_data_stack_mode(pts.to(DEV), torch.tensor(sizes), 4, 0.025, 0.0625, limits)
    assert set(got) == set(want)
    perm = []  # perm[i][our index] = the reference's index of the same point (with the sentinel mapped to the sentinel)
    for i in range(4):
        assert np.array_equal(got['lengths'][i].numpy(), want['lengths'][i])
        a, b = got['points'][i].cpu().numpy(), want['points'][i]
        if i == 0:
            assert np.array_equal(a, b)
            m = np.arange(len(a))
        else:
            d, m = cKDTree(b).query(a)
            assert d.max() <= 1e-5 and len(np.unique(m)) == len(a)
        perm.append(np.concatenate([m, [len(a)]]))

    def mismatch(ours, theirs, q_perm, s_perm):
        ours = s_perm[ours.cpu().numpy()]
        rows = np.empty_like(theirs)
        rows[q_perm[:-1]] = ours[:, :theirs.shape[1]] if ours.shape[1] >= theirs.shape[1] else np.pad(
            ours, ((0, 0), (0, theirs.shape[1] - ours.shape[1])), constant_values=s_perm[-1])
        return float((np.sort(rows, 1) != np.sort(theirs, 1)).any(1).mean())

    assert_same_neighbors(got['neighbors'][0].cpu().numpy(), want['neighbors'][0], pts.numpy(), pts.numpy(), len(pts))
    for i in range(4):
        assert mismatch(got['neighbors'][i], want['neighbors'][i], perm[i], perm[i]) < 2e-3
        if i < 3:
            assert mismatch(got['subsampling'][i], want['subsampling'][i], perm[i + 1], perm[i]) < 2e-3
            assert mismatch(got['upsampling'][i], want['upsampling'][i], perm[i], perm[i + 1]) < 2e-3
