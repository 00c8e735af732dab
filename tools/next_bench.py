"""Times the 'next' rows of SURVEY.md 8(f) alone on the GPU: K6 point-to-node partition and K7 optimal transport, at the
3DMatch and KITTI shapes (and beside them the same step written with stock torch ops on the same GPU, i.e. the
reference's formulation).  python tools/next_bench.py [--no-torch] [--once]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
once_only = '--once' in sys.argv  # one call per kernel: for ncu


def timed(fn, iters=50):
    fn()
    torch.cuda.synchronize()
    if once_only:
        return 0.0
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(10):
            keep = fn()
    g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters // 10):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    del keep
    return a.elapsed_time(b) * 1e3 / (iters // 10 * 10)


def eager(fn, iters=10):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / iters


def torch_partition(points, nodes, K):
    """pointcloud_partition.py:82-98 with stock torch CUDA ops."""
    x2 = (nodes ** 2).sum(-1)[:, None]
    y2 = (points ** 2).sum(-1)[None, :]
    d = (x2 - 2 * nodes @ points.t() + y2).clamp(min=0.0)
    p2n = d.min(dim=0)[1]
    masks = torch.zeros(nodes.shape[0], dtype=torch.bool, device=points.device)
    masks.index_fill_(0, p2n, True)
    match = torch.zeros_like(d, dtype=torch.bool)
    match[p2n, torch.arange(points.shape[0], device=points.device)] = True
    d.masked_fill_(~match, 1e12)
    idx = d.topk(k=K, dim=1, largest=False)[1]
    kmask = p2n[idx] == torch.arange(nodes.shape[0], device=points.device)[:, None]
    idx.masked_fill_(~kmask, points.shape[0])
    return p2n, masks, idx, kmask


def torch_transport(scores, alpha, iters, rm, cm, inf=1e12):
    """learnable_sinkhorn.py:21-63 with stock torch CUDA ops."""
    B, M, N = scores.shape
    prm = torch.zeros(B, M + 1, dtype=torch.bool, device=scores.device)
    prm[:, :M] = ~rm
    pcm = torch.zeros(B, N + 1, dtype=torch.bool, device=scores.device)
    pcm[:, :N] = ~cm
    S = torch.cat([torch.cat([scores, alpha.expand(B, M, 1)], -1), alpha.expand(B, 1, N + 1)], 1)
    S.masked_fill_(prm[:, :, None] | pcm[:, None, :], -inf)
    nr, nc = rm.float().sum(1), cm.float().sum(1)
    norm = -torch.log(nr + nc)
    mu = norm[:, None].expand(B, M + 1).clone()
    mu[:, M] = torch.log(nc) + norm
    mu[prm] = -inf
    nu = norm[:, None].expand(B, N + 1).clone()
    nu[:, N] = torch.log(nr) + norm
    nu[pcm] = -inf
    u, v = torch.zeros_like(mu), torch.zeros_like(nu)
    for _ in range(iters):
        u = mu - torch.logsumexp(S + v[:, None, :], dim=2)
        v = nu - torch.logsumexp(S + u[:, :, None], dim=1)
    return S + u[:, :, None] + v[:, None, :] - norm[:, None, None]


def torch_procrustes(src, ref, w, eps=1e-5):
    """procrustes.py:40-67 with stock torch ops, including the reference's host round trip for the SVD."""
    w = (w / (w.sum(1, keepdim=True) + eps))[:, :, None]
    cs, cr = (src * w).sum(1, keepdim=True), (ref * w).sum(1, keepdim=True)
    H = (src - cs).permute(0, 2, 1) @ (w * (ref - cr))
    U, _, V = torch.svd(H.cpu())
    Ut, V = U.transpose(1, 2).cuda(), V.cuda()
    eye = torch.eye(3).unsqueeze(0).repeat(src.shape[0], 1, 1).cuda()
    eye[:, -1, -1] = torch.sign(torch.det(V @ Ut))
    R = V @ eye @ Ut
    return R, (cr.permute(0, 2, 1) - R @ cs.permute(0, 2, 1)).squeeze(2)


def main():
    with_torch = '--no-torch' not in sys.argv and not once_only
    res = {}
    for shape, nodes_n, ppn, K in (('3dmatch', 1024, 16, 64), ('kitti', 2048, 24, 128), ('3dmatch', 4096, 16, 64)):
        points, nodes = (t.cuda() for t in pkg.synth.make_partition_inputs(shape, nodes_n, ppn))
        us = timed(lambda: pkg.ops.point_to_node_partition(points, nodes, K, return_points=True))
        row = {'us': us, 'points': points.shape[0], 'nodes': nodes_n, 'point_limit': K,
               'pair_evals_per_s': points.shape[0] * nodes_n / max(us, 1e-9) * 1e6}
        if with_torch:
            row['torch_eager_us'] = eager(lambda: torch_partition(points, nodes, K))
        res[f'partition_{shape}_{nodes_n}'] = row
    for shape, K in (('3dmatch', 64), ('kitti', 128)):
        scores, rm, cm = (t.cuda() for t in pkg.synth.make_patch_scores(256, K, K, 256))
        alpha = torch.ones((), device='cuda')
        us = timed(lambda: pkg.ops.log_optimal_transport(scores, alpha, 100, rm, cm))
        row = {'us': us, 'problems': 256, 'size': K + 1, 'iterations': 100,
               'ns_per_half_iteration': us * 1e3 / 200}
        if with_torch:
            row['torch_eager_us'] = eager(lambda: torch_transport(scores, alpha, 100, rm, cm), iters=3)
        res[f'transport_{shape}'] = row
    for batch, n in ((256, 64), (256, 128), (1, 5000)):
        src, ref, w = (t.cuda() for t in pkg.synth.make_procrustes_inputs(batch, n)[:3])
        row = {'us': timed(lambda: pkg.ops.weighted_procrustes(src, ref, w)), 'problems': batch, 'points': n}
        if with_torch:
            row['torch_eager_host_svd_us'] = eager(lambda: torch_procrustes(src, ref, w))
        res[f'procrustes_{batch}x{n}'] = row
    import time
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    import ref_cpu
    for sizes, radius, limit in (((20000, 18000), 0.0625, 38), ((5000, 4500), 0.25, 36)):
        clouds = [pkg.synth.make_dense_cloud(n, 3.0, 900 + b) for b, n in enumerate(sizes)]
        s_cpu = torch.cat(clouds)
        s_gpu = s_cpu.cuda()
        row = {'us': timed(lambda: pkg.ops.radius_neighbors(s_gpu, s_gpu, list(sizes), list(sizes), radius, limit)),
               'brute_us': timed(lambda: pkg.ops.radius_neighbors(s_gpu, s_gpu, list(sizes), list(sizes), radius, limit, method='brute')),
               'points': list(sizes), 'radius': radius, 'limit': limit,
               'pair_evals': sum(n * n for n in sizes)}
        if with_torch and ref_cpu.available():  # the reference's own operator (nanoflann KD-tree, one host thread)
            t0 = time.perf_counter()
            ref_cpu.radius_neighbors(s_cpu.numpy(), s_cpu.numpy(), list(sizes), list(sizes), radius)
            row['reference_cpu_operator_us'] = (time.perf_counter() - t0) * 1e6
        res[f'radius_neighbors_{sizes[0]}'] = row
    for sizes, voxel in (((30000, 25000), 0.05), ((120000, 100000), 0.05)):
        clouds = [pkg.synth.make_dense_cloud(n, 3.0, 700 + b) for b, n in enumerate(sizes)]
        p_cpu = torch.cat(clouds)
        p_gpu = p_cpu.cuda()
        row = {'us': timed(lambda: pkg.ops.grid_subsample(p_gpu, list(sizes), voxel)), 'points': list(sizes), 'voxel': voxel}
        if with_torch and ref_cpu.available():  # the reference's own operator (std::unordered_map, one host thread)
            t0 = time.perf_counter()
            ref_cpu.grid_subsampling(p_cpu.numpy(), list(sizes), voxel)
            row['reference_cpu_operator_us'] = (time.perf_counter() - t0) * 1e6
        res[f'grid_subsample_{sizes[0]}'] = row
    sizes, limits = (20000, 18000), [38, 36, 36, 38]
    p_cpu = torch.cat([pkg.synth.make_dense_cloud(n, 3.0, 800 + b) for b, n in enumerate(sizes)])
    p_gpu = p_cpu.cuda()
    row = {'us': eager(lambda: pkg.precompute_data_stack_mode(p_gpu, list(sizes), 4, 0.025, 0.0625, limits), iters=5),
           'points': list(sizes), 'stages': 4, 'note': 'eager (4 host reads: 3 subsampled lengths, then the 10 result widths together), CUDA events'}
    if with_torch and ref_cpu.available():
        t0 = time.perf_counter()
        ref_cpu.precompute_data_stack_mode(p_cpu.numpy(), list(sizes), 4, 0.025, 0.0625, limits)
        row['reference_cpu_operators_us'] = (time.perf_counter() - t0) * 1e6
    res['pyramid_3dmatch'] = row
    print(json.dumps(res))


if __name__ == '__main__':
    main()
