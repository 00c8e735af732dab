"""Synthetic inputs for the hot path: seeded, shaped like the reference's three
experiment configs (experiments/*/config.py:103-113).  Used by tests, bench.py and smoke()."""
import math

import numpy as np
import torch

SEED = 7351  # experiments/*/config.py:13

CONFIGS = {
    # name: GeometricTransformer ctor args (experiments/*/config.py:103-113, ModelNet :123-133)
    '3dmatch': dict(input_dim=1024, output_dim=256, hidden_dim=256, num_heads=4,
                    blocks=['self', 'cross'] * 3, sigma_d=0.2, sigma_a=15, angle_k=3, reduction_a='max'),
    'kitti': dict(input_dim=2048, output_dim=256, hidden_dim=128, num_heads=4,
                  blocks=['self', 'cross'] * 3, sigma_d=4.8, sigma_a=15, angle_k=3, reduction_a='max'),
    'modelnet': dict(input_dim=512, output_dim=256, hidden_dim=256, num_heads=4,
                     blocks=['self', 'cross'] * 3, sigma_d=0.2, sigma_a=15, angle_k=3, reduction_a='max'),
}


def _rotation(gen, max_deg):
    axis = torch.randn(3, generator=gen, dtype=torch.float64)
    axis = axis / axis.norm()
    ang = (torch.rand(1, generator=gen, dtype=torch.float64).item()) * math.radians(max_deg)
    kx = torch.tensor([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]], dtype=torch.float64)
    return torch.eye(3, dtype=torch.float64) + math.sin(ang) * kx + (1 - math.cos(ang)) * (kx @ kx)


def make_cloud(shape, n, gen):
    if shape == '3dmatch':
        s = 3.0 * (n / 1024.0) ** (1.0 / 3.0)  # ~0.3 m superpoint spacing
        return torch.rand(n, 3, generator=gen) * s
    if shape == 'kitti':
        return torch.rand(n, 3, generator=gen) * torch.tensor([160.0, 160.0, 8.0]) - torch.tensor([80.0, 80.0, 4.0])
    if shape == 'modelnet':
        return torch.rand(n, 3, generator=gen)
    raise ValueError(shape)


def make_pair(shape='3dmatch', n=1024, m=None, seed=SEED, input_dim=None):
    """Returns ref_points (1,n,3), src_points (1,m,3), ref_feats (1,n,Din), src_feats (1,m,Din), fp32 CPU."""
    m = n if m is None else m
    din = CONFIGS[shape]['input_dim'] if input_dim is None else input_dim
    gen = torch.Generator().manual_seed(seed)
    ref = make_cloud(shape, n, gen)
    perm = torch.randperm(n, generator=gen)[:m]
    rot = _rotation(gen, 45.0)
    trans = (torch.rand(3, generator=gen, dtype=torch.float64) - 0.5)
    src = (ref[perm].double() @ rot.T + trans).float() + torch.randn(m, 3, generator=gen) * 0.01
    ref_feats = torch.randn(n, din, generator=gen)
    src_feats = torch.randn(m, din, generator=gen)
    return ref[None].contiguous(), src[None].contiguous(), ref_feats[None].contiguous(), src_feats[None].contiguous()


def tie_free(points, k=3):
    """Precondition for bit-exact k-NN parity.  torch.topk leaves the order of exact ties
    unspecified, so parity with the reference is only defined when the reference's own fp32 distances
    (sqrt(max(0, x2 - 2xy + y2)), pairwise_distance.py:22-30) are strictly increasing over the first k+2
    sorted neighbours of every row.  points: (n,3) fp32 CPU tensor."""
    x = points.float()
    n = x.shape[0]
    sq = torch.sum(x ** 2, dim=-1)
    d = torch.sqrt((sq.unsqueeze(-1) - 2 * torch.matmul(x, x.t()) + sq.unsqueeze(-2)).clamp(min=0.0))
    kk = min(k + 2, n)
    part = d.topk(kk, dim=1, largest=False)[0]
    return bool((part[:, 1:] > part[:, :-1]).all())


def make_tie_free_pair(shape='3dmatch', n=1024, m=None, seed=SEED, input_dim=None, k=3, max_tries=64):
    """make_pair, bumping the seed by 1000 until both clouds satisfy tie_free ("else bump the
    seed and record it").  Returns (ref_points, src_points, ref_feats, src_feats, seed_used)."""
    for t in range(max_tries):
        s = seed + 1000 * t
        out = make_pair(shape, n, m, s, input_dim)
        if tie_free(out[0][0], k) and tie_free(out[1][0], k):
            return out + (s,)
    raise RuntimeError('no tie-free seed found')


def make_fine_cloud(shape, nodes, points_per_node, gen):
    """Fine-level points for the point-to-node partition: `points_per_node` points scattered around
    every node (sigma = half the node spacing, so patches overlap and sizes vary), shuffled.  nodes (M,3) -> (N,3).
    3DMatch: stage-2 voxels 0.05 m under stage-4 nodes 0.2 m => about 16 points per node (config.py:79)."""
    m = nodes.shape[0]
    spacing = {'3dmatch': 0.3, 'kitti': 4.8, 'modelnet': 0.1}[shape]
    n = m * points_per_node
    owner = torch.arange(n) % m
    pts = nodes[owner] + torch.randn(n, 3, generator=gen) * (0.5 * spacing)
    return pts[torch.randperm(n, generator=gen)].contiguous()


def make_partition_inputs(shape='3dmatch', num_nodes=1024, points_per_node=16, seed=SEED):
    """Returns (points (N,3), nodes (M,3)) fp32 CPU for point_to_node_partition."""
    gen = torch.Generator().manual_seed(seed)
    nodes = make_cloud(shape, num_nodes, gen)
    return make_fine_cloud(shape, nodes, points_per_node, gen), nodes.contiguous()


def make_patch_scores(num_patches=256, rows=64, cols=64, dim=256, seed=SEED, valid_frac=0.8):
    """Inputs of the optimal-transport step (experiments/*/model.py:184-189): patch feature products / sqrt(dim) and
    the patches' validity masks (every patch keeps at least one valid point, as the partition guarantees).
    Returns scores (P,rows,cols) fp32, row_masks (P,rows) bool, col_masks (P,cols) bool."""
    gen = torch.Generator().manual_seed(seed)
    a = torch.randn(num_patches, rows, dim, generator=gen)
    b = a[:, torch.randperm(rows, generator=gen)][:, :cols] if cols <= rows else torch.randn(num_patches, cols, dim, generator=gen)
    b = b + 0.5 * torch.randn(num_patches, cols, dim, generator=gen)
    scores = torch.einsum('bnd,bmd->bnm', a, b) / dim ** 0.5
    nr = (torch.rand(num_patches, generator=gen) * valid_frac * rows).long() + max(1, int((1 - valid_frac) * rows))
    nc = (torch.rand(num_patches, generator=gen) * valid_frac * cols).long() + max(1, int((1 - valid_frac) * cols))
    row_masks = torch.arange(rows)[None, :] < nr.clamp(1, rows)[:, None]  # valid points come first (sorted patches)
    col_masks = torch.arange(cols)[None, :] < nc.clamp(1, cols)[:, None]
    return scores.contiguous(), row_masks.contiguous(), col_masks.contiguous()


def make_procrustes_inputs(batch=256, n=64, seed=SEED, noise=0.02, outliers=0.2, planar=False, reflect=False):
    """Inputs of weighted_procrustes as the fine matching feeds it (local_global_registration.py:169): per problem a
    random rigid motion of `n` points plus noise, a fraction of gross outliers with small weights, zero-padded tails
    (weights 0).  planar=True puts the points on a plane (rank-2 covariance); reflect=True mirrors ref so that the
    determinant correction of the Kabsch solution is exercised.
    Returns src (B,n,3), ref (B,n,3), weights (B,n), true transforms (B,4,4), fp32 CPU."""
    gen = torch.Generator().manual_seed(seed)
    src = torch.rand(batch, n, 3, generator=gen) * 2 - 1
    if planar:
        src[:, :, 2] = 0.3 * src[:, :, 0] - 0.2 * src[:, :, 1]
    T = torch.eye(4).repeat(batch, 1, 1)
    for b in range(batch):
        T[b, :3, :3] = _rotation(gen, 180.0).float()
        T[b, :3, 3] = torch.rand(3, generator=gen) * 2 - 1
    ref = src @ T[:, :3, :3].transpose(1, 2) + T[:, None, :3, 3] + noise * torch.randn(batch, n, 3, generator=gen)
    if reflect:
        ref[:, :, 0] = -ref[:, :, 0]
    weights = torch.rand(batch, n, generator=gen) * 0.9 + 0.1
    bad = torch.rand(batch, n, generator=gen) < outliers
    ref = torch.where(bad[:, :, None], torch.rand(batch, n, 3, generator=gen) * 4 - 2, ref)
    weights = torch.where(bad, weights * 0.05, weights)
    valid = (torch.rand(batch, generator=gen) * 0.5 + 0.5) * n
    pad = torch.arange(n)[None, :] >= valid[:, None].clamp(min=4)
    weights = torch.where(pad, torch.zeros_like(weights), weights)
    src = torch.where(pad[:, :, None], torch.zeros_like(src), src)
    ref = torch.where(pad[:, :, None], torch.zeros_like(ref), ref)
    return src.contiguous(), ref.contiguous(), weights.contiguous(), T


def make_dense_cloud(n, extent=3.0, seed=SEED, jitter=0.005):
    """A raw-scan-like cloud for the collate operators: n points on three orthogonal walls of a
    room of size `extent` (surfaces, not a volume, so that neighbour counts per radius look like a depth scan's), with
    a little off-surface jitter.  Returns (n,3) fp32 CPU."""
    gen = torch.Generator().manual_seed(seed)
    uv = torch.rand(n, 2, generator=gen) * extent
    wall = torch.randint(0, 3, (n,), generator=gen)
    off = torch.randn(n, generator=gen) * jitter
    pts = torch.empty(n, 3)
    for w in range(3):
        m = wall == w
        a, b = [(1, 2), (0, 2), (0, 1)][w]
        pts[m, a], pts[m, b], pts[m, w] = uv[m, 0], uv[m, 1], off[m]
    return pts.contiguous()


def make_lgr_inputs(num_patches=256, K=64, seed=SEED, outlier_patches=0.3, noise=0.01, valid_frac=0.7):
    """Inputs of LocalGlobalRegistration.forward as model.py:198-205 feeds it: per patch correspondence K reference and
    K source points (src = the reference points under ONE global rigid motion, permuted, plus noise; in a fraction of
    the patches the source points are unrelated: wrong superpoint matches), validity masks, and a (P,K,K) matrix of
    log matching scores that is confident (0.3 ... 0.9) on the true point matches and near zero elsewhere.
    Returns ref_knn_points, src_knn_points, ref_knn_masks, src_knn_masks, score_mat, global_scores, transform."""
    gen = torch.Generator().manual_seed(seed)
    centers = torch.rand(num_patches, 1, 3, generator=gen) * 3.0
    ref = centers + (torch.rand(num_patches, K, 3, generator=gen) - 0.5) * 0.3
    T = torch.eye(4)
    T[:3, :3] = _rotation(gen, 60.0).float()
    T[:3, 3] = torch.rand(3, generator=gen) - 0.5
    # ref = R src + t  =>  src = R^T (ref - t)
    src_true = (ref - T[:3, 3]) @ T[:3, :3]
    perm = torch.stack([torch.randperm(K, generator=gen) for _ in range(num_patches)])  # src slot j holds ref perm[j]
    src = torch.gather(src_true, 1, perm[:, :, None].expand(-1, -1, 3)) + noise * torch.randn(num_patches, K, 3, generator=gen)
    bad = torch.rand(num_patches, generator=gen) < outlier_patches
    src[bad] = torch.rand(int(bad.sum()), K, 3, generator=gen) * 3.0
    nr = ((torch.rand(num_patches, generator=gen) * (1 - valid_frac) + valid_frac) * K).long().clamp(3, K)
    ns = ((torch.rand(num_patches, generator=gen) * (1 - valid_frac) + valid_frac) * K).long().clamp(3, K)
    ref_masks = torch.arange(K)[None, :] < nr[:, None]
    src_masks = torch.arange(K)[None, :] < ns[:, None]
    prob = torch.rand(num_patches, K, K, generator=gen) * 0.02  # background
    conf = torch.rand(num_patches, K, generator=gen) * 0.6 + 0.3
    weak = torch.rand(num_patches, K, generator=gen) < 0.25      # a quarter of the true matches stay below the threshold
    conf = torch.where(weak, conf * 0.1, conf)
    prob[torch.arange(num_patches)[:, None], perm, torch.arange(K)[None, :].expand(num_patches, -1)] = conf
    global_scores = torch.rand(num_patches, generator=gen) * 0.5 + 0.5
    return (ref.contiguous(), src.contiguous(), ref_masks.contiguous(), src_masks.contiguous(),
            torch.log(prob).contiguous(), global_scores, T)


def make_tail_inputs(num_nodes=512, points_per_node=16, dim=256, seed=SEED, keep=0.85, noise=0.003, feat_noise=0.05):
    """A registration pair for the stages AFTER the transformer (experiments/*/model.py:98-108,141-205: partition,
    coarse matching, patch gather, optimal transport, local-to-global registration): two overlapping fine clouds with
    their coarse nodes and stand-ins for the backbone's features -- random Fourier features of a point's position in
    the REFERENCE frame (coarse: wavelength ~ node spacing, fine: ~ point spacing), so that corresponding points have
    near-equal features whatever the motion.
    Returns dict(ref_points_f, src_points_f, ref_points_c, src_points_c, ref_feats_c, src_feats_c, ref_feats_f,
    src_feats_f, transform) -- all fp32 CPU; transform (4,4) maps src onto ref."""
    gen = torch.Generator().manual_seed(seed)
    ref_c = make_cloud('3dmatch', num_nodes, gen)
    ref_f = make_fine_cloud('3dmatch', ref_c, points_per_node, gen)
    R = _rotation(gen, 45.0).float()
    t = (torch.rand(3, generator=gen) - 0.5)
    T = torch.eye(4)
    T[:3, :3], T[:3, 3] = R, t

    def to_src(p):  # ref frame -> src frame: p_ref = R p_src + t
        return (p - t) @ R  # rows: R^T (p - t)

    def take(p, frac, jitter):
        idx = torch.randperm(p.shape[0], generator=gen)[:max(4, int(frac * p.shape[0]))]
        q = p[idx] + torch.randn(idx.shape[0], 3, generator=gen) * jitter
        return q.contiguous()

    src_f_ref = take(ref_f, keep, noise)        # src points, still in the ref frame
    src_c_ref = take(ref_c, keep, 0.02)

    def fourier(p, wavelength, g):
        omega = torch.randn(3, dim, generator=g) * (2.0 * math.pi / wavelength) / math.sqrt(3.0)
        phase = torch.rand(dim, generator=g) * 2.0 * math.pi
        return omega, phase

    gc = torch.Generator().manual_seed(seed + 1)
    oc, pc = fourier(None, 1.2, gc)
    of, pf = fourier(None, 0.25, gc)

    def feats(p, omega, phase):
        f = torch.cos(p @ omega + phase)
        return (f + torch.randn(f.shape, generator=gen) * feat_noise).contiguous()

    return dict(ref_points_f=ref_f, src_points_f=to_src(src_f_ref).contiguous(), ref_points_c=ref_c.contiguous(),
                src_points_c=to_src(src_c_ref).contiguous(), ref_feats_c=feats(ref_c, oc, pc),
                src_feats_c=feats(src_c_ref, oc, pc), ref_feats_f=feats(ref_f, of, pf),
                src_feats_f=feats(src_f_ref, of, pf), transform=T)
