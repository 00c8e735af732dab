"""Tensor-level wrappers over the C ABI (include/geotx_b200.h).  torch is used for device memory and the
current CUDA stream only; every function launches hand-written sm_100a kernels and raises RuntimeError on any
non-CUDA / wrongly-typed input (no CPU fallback, same policy as the reference's CHECK_* macros,
/root/reference/geotransformer/extensions/common/torch_helper.h:6-35)."""
import ctypes

import torch

from . import _abi

GEOTX_F32 = _abi.GEOTX_F32
GEOTX_BF16 = _abi.GEOTX_BF16
GEOTX_F16 = _abi.GEOTX_F16
_DT = {torch.float32: GEOTX_F32, torch.bfloat16: GEOTX_BF16, torch.float16: GEOTX_F16}


def _check(t, name, dtype=None, dims=None):
    if not isinstance(t, torch.Tensor):
        raise RuntimeError(f'{name} must be a torch.Tensor')
    if not t.is_cuda:
        raise RuntimeError(f'{name} must be a CUDA tensor (ref-geotransformer_b200 has no CPU path)')
    _same_device(t, name)
    if not t.is_contiguous():
        raise RuntimeError(f'{name} must be contiguous')
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f'{name} must be {dtype}, got {t.dtype}')
    if dims is not None and t.dim() != dims:
        raise RuntimeError(f'{name} must have {dims} dims, got shape {tuple(t.shape)}')
    return t


def _same_device(t, name):
    """The C ABI launches on the CURRENT device's current stream: a tensor that lives on another GPU would be addressed
    from the wrong context.  Fail loudly; the caller wraps the call in `with torch.cuda.device(t.device):`."""
    cur = torch.cuda.current_device()
    if t.device.index is not None and t.device.index != cur:
        raise RuntimeError(f'{name} lives on cuda:{t.device.index} but the current device is cuda:{cur}; call under '
                           f'`with torch.cuda.device({name}.device):`')


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class Geometry:
    """Packed points + k-NN reference vectors of B clouds (the scratch geotx_gse_knn fills)."""
    __slots__ = ('buf', 'B', 'n', 'k', 'knn')

    def __init__(self, buf, B, n, k, knn):
        self.buf, self.B, self.n, self.k, self.knn = buf, B, n, k, knn


def gse_knn(points, k):
    """points (B,n,3) fp32 -> Geometry (with .knn (B,n,k) int64).  geotransformer.py:38-42."""
    _check(points, 'points', torch.float32, 3)
    B, n, three = points.shape
    if three != 3:
        raise RuntimeError(f'points must be (B, n, 3), got {tuple(points.shape)}')
    lib = _abi.load()
    nbytes = lib.geotx_gse_geometry_bytes(B, n, k)
    buf = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=points.device)
    knn = torch.empty(B, n, k, dtype=torch.int64, device=points.device)
    _abi.call('geotx_gse_knn', _p(points), B, n, k, _p(knn), _p(buf), _stream())
    return Geometry(buf, B, n, k, knn)


def gse_indices(geom, sigma_d, factor_a):
    """-> d_idx (B,n,n), a_idx (B,n,n,k) fp32.  geotransformer.py:25-55."""
    dev = geom.buf.device
    d_idx = torch.empty(geom.B, geom.n, geom.n, dtype=torch.float32, device=dev)
    a_idx = torch.empty(geom.B, geom.n, geom.n, geom.k, dtype=torch.float32, device=dev)
    _abi.call('geotx_gse_indices', _p(geom.buf), geom.B, geom.n, geom.k, float(sigma_d), float(factor_a),
              _p(d_idx), _p(a_idx), _stream())
    return d_idx, a_idx


def gse_pack_weights(w_d, b_d, w_a, b_a):
    for t, nme in ((w_d, 'proj_d.weight'), (b_d, 'proj_d.bias'), (w_a, 'proj_a.weight'), (b_a, 'proj_a.bias')):
        _check(t, nme, torch.float32)
    C = w_d.shape[0]
    lib = _abi.load()
    nbytes = lib.geotx_gse_packed_weight_bytes(C)
    if nbytes <= 0:
        raise RuntimeError(f'hidden_dim {C} is not supported by the tcgen05 embedding kernel')
    packed = torch.empty(nbytes, dtype=torch.uint8, device=w_d.device)
    _abi.call('geotx_gse_pack_weights', _p(w_d), _p(b_d), _p(w_a), _p(b_a), C, _p(packed), _stream())
    return packed


def _slab(geom, row_begin, row_count):
    row_count = geom.n - row_begin if row_count is None else row_count
    return row_begin, row_count


def gse_embed(geom, div_term, packed, C, sigma_d, factor_a, reduce_mean=False, out_dtype=torch.float16,
              row_begin=0, row_count=None, out=None, qt=None, scores_out=None):
    """tcgen05 path.  -> E (B,row_count,n,C).  geotransformer.py:57-72.
    With `qt` (folded query of the first 'self' block, logical (B,H,n,C), see rpe_scores) the kernel also writes
    scores_p (B,H,n,n) fp32 for the slab's rows into `scores_out` (allocated if None) and returns (E, scores_p)."""
    _check(div_term, 'div_term', torch.float32, 1)
    row_begin, row_count = _slab(geom, row_begin, row_count)
    if out is None:
        out = torch.empty(geom.B, row_count, geom.n, C, dtype=out_dtype, device=geom.buf.device)
    else:
        _check(out, 'out', out_dtype)
    head = (_p(geom.buf), geom.B, geom.n, geom.k, row_begin, row_count, _p(div_term), _p(packed), C, float(sigma_d),
            float(factor_a), int(bool(reduce_mean)), _DT[out_dtype], _p(out))
    if qt is None:
        _abi.call('geotx_gse_embed', *head, _stream())
        return out
    qt, H = _folded_query(qt, geom.B, geom.n, C)
    if scores_out is None:
        scores_out = torch.empty(geom.B, H, geom.n, geom.n, dtype=torch.float32, device=out.device)
    elif tuple(_check(scores_out, 'scores_out', torch.float32, 4).shape) != (geom.B, H, geom.n, geom.n):
        raise RuntimeError(f'scores_out must be {(geom.B, H, geom.n, geom.n)}, got {tuple(scores_out.shape)}')
    _abi.call('geotx_gse_embed_scores', *head, _p(qt), qt.stride(0), qt.stride(1), qt.stride(2), H, _p(scores_out),
              _stream())
    return out, scores_out


def _folded_query(qt, B, n, C):
    """(B,H,n,C) fp32 CUDA view with unit channel stride and 16-byte aligned rows (made contiguous otherwise)."""
    if not (isinstance(qt, torch.Tensor) and qt.is_cuda and qt.dtype == torch.float32 and qt.dim() == 4):
        raise RuntimeError('qt must be a 4-d fp32 CUDA tensor')
    if qt.shape[0] != B or qt.shape[2] != n or qt.shape[3] != C:
        raise RuntimeError(f'qt must be (B={B}, H, n={n}, C={C}), got {tuple(qt.shape)}')
    if qt.stride(3) != 1 or any(st % 4 for st in qt.stride()[:3]) or qt.data_ptr() % 16:
        qt = qt.contiguous()
    return qt, qt.shape[1]


def gse_embed_fp32(geom, div_term, w_d, b_d, w_a, b_a, sigma_d, factor_a, reduce_mean=False,
                   out_dtype=torch.float32, row_begin=0, row_count=None, out=None):
    """CUDA-core fp32 path, any even hidden size."""
    for t, nme in ((w_d, 'proj_d.weight'), (b_d, 'proj_d.bias'), (w_a, 'proj_a.weight'), (b_a, 'proj_a.bias'),
                   (div_term, 'div_term')):
        _check(t, nme, torch.float32)
    C = w_d.shape[0]
    row_begin, row_count = _slab(geom, row_begin, row_count)
    if out is None:
        out = torch.empty(geom.B, row_count, geom.n, C, dtype=out_dtype, device=geom.buf.device)
    _abi.call('geotx_gse_embed_fp32', _p(geom.buf), geom.B, geom.n, geom.k, row_begin, row_count, _p(div_term),
              _p(w_d), _p(b_d), _p(w_a), _p(b_a), C, float(sigma_d), float(factor_a), int(bool(reduce_mean)),
              _DT[out_dtype], _p(out), _stream())
    return out


def _row_strided(t, name, dtype=torch.float32):
    """(B, tokens, D) fp32 (or half) CUDA view with unit channel stride, 16-byte aligned rows: returns (tensor, row
    stride).  Column blocks of a fused projection output qualify; anything else is made contiguous."""
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == dtype and t.dim() == 3):
        raise RuntimeError(f'{name} must be a 3-d {dtype} CUDA tensor (ref-geotransformer_b200 has no CPU path)')
    _same_device(t, name)
    B, n, D = t.shape
    unit = 16 // t.element_size()
    ok = t.stride(2) == 1 and t.stride(1) % unit == 0 and t.stride(1) >= D and t.data_ptr() % 16 == 0 and \
        (B == 1 or t.stride(0) == n * t.stride(1))
    if not ok:
        t = t.contiguous()
    return t, t.stride(1)


def rpe_scores(qt, E, n=None, row_begin=0, out=None):
    """qt: folded query, logical shape (B,H,n,C) fp32 (any view with unit channel stride);
    E (B,row_count,m,C) fp16|bf16|fp32  ->  scores_p (B,H,n,m) fp32 (slab rows written).
    rpe_transformer.py:54-56 with proj_p folded."""
    if not (isinstance(qt, torch.Tensor) and qt.is_cuda and qt.dtype == torch.float32 and qt.dim() == 4):
        raise RuntimeError('qt must be a 4-d fp32 CUDA tensor')
    B, H, nq, C = qt.shape
    if qt.stride(3) != 1 or any(st % 4 for st in qt.stride()[:3]) or qt.data_ptr() % 16:
        qt = qt.contiguous()
    _check(E, 'E', None, 4)
    if E.dtype not in _DT:
        raise RuntimeError(f'E must be fp32, fp16 or bf16, got {E.dtype}')
    Bq, row_count, m, Ce = E.shape
    if Bq != B or Ce != C:
        raise RuntimeError(f'shape mismatch: qt {tuple(qt.shape)} vs E {tuple(E.shape)}')
    n = nq if n is None else n
    if out is None:
        out = torch.empty(B, H, n, m, dtype=torch.float32, device=E.device)
    _abi.call('geotx_rpe_scores', _p(qt), qt.stride(0), qt.stride(1), qt.stride(2), _p(E), _DT[E.dtype], B, H, n, m, C,
              row_begin, row_count, _p(out), _stream())
    return out


def attention(q, k, v, num_heads, scores_p=None, key_mask=None, scale=None, split_keys=True, out=None):
    """q (B,n,H*c), k,v (B,m,H*c) fp32 token-major (dense or column blocks of a wider row-major matrix)
    -> (B,n,H*c) dense.  softmax((q k^T + scores_p)*scale) v.  rpe_transformer.py:57-70, vanilla_transformer.py:54-68.
    q, k, v may also be float16 (all three: what `linear_f16(..., half_cols=)` writes): the fp32 kernel (K3b) rounds them
    to half itself, the half kernel (K3h, geotx_attention_f16) takes them as they are -- same operands, same result up to
    fp32 summation order."""
    half = isinstance(q, torch.Tensor) and q.dtype == torch.float16
    dt = torch.float16 if half else torch.float32
    q, ldq = _row_strided(q, 'q', dt)
    k, ldk = _row_strided(k, 'k', dt)
    v, ldv = _row_strided(v, 'v', dt)
    B, n, D = q.shape
    m = k.shape[1]
    c = D // num_heads
    if scores_p is not None:
        _check(scores_p, 'scores_p', torch.float32, 4)
        if tuple(scores_p.shape) != (B, num_heads, n, m):
            raise RuntimeError(f'scores_p must be {(B, num_heads, n, m)}, got {tuple(scores_p.shape)}')
    mask_u8 = None
    if key_mask is not None:
        if tuple(key_mask.shape) != (B, m):
            raise RuntimeError(f'key_mask must be {(B, m)}, got {tuple(key_mask.shape)}')
        mask_u8 = _check(key_mask.to(torch.uint8).contiguous(), 'key_mask')
    scale = 1.0 / (c ** 0.5) if scale is None else scale
    if out is None:
        out = torch.empty(B, n, D, dtype=torch.float32, device=q.device)
    elif tuple(_check(out, 'out', torch.float32, 3).shape) != (B, n, D):
        raise RuntimeError(f'out must be {(B, n, D)}, got {tuple(out.shape)}')
    _abi.call('geotx_attention_f16' if half else 'geotx_attention', _p(q), _p(k), _p(v), _p(scores_p), _p(mask_u8), B,
              num_heads, n, m, c, ldq, ldk, ldv, float(scale), _p(out), int(bool(split_keys)), _stream())
    return out


def add_layernorm(x, residual, weight, bias, eps=1e-5, out=None):
    """LayerNorm(x + residual) over the last dim in one pass (residual may be None).  fp32, C % 4 == 0, C <= 1024."""
    _check(x, 'x', torch.float32)
    if residual is not None:
        _check(residual, 'residual', torch.float32)
        if residual.shape != x.shape:
            raise RuntimeError(f'residual shape {tuple(residual.shape)} != x shape {tuple(x.shape)}')
    _check(weight, 'weight', torch.float32, 1)
    _check(bias, 'bias', torch.float32, 1)
    C = x.shape[-1]
    if out is None:
        out = torch.empty_like(x)
    elif _check(out, 'out', torch.float32).shape != x.shape:
        raise RuntimeError(f'out shape {tuple(out.shape)} != x shape {tuple(x.shape)}')
    _abi.call('geotx_add_layernorm', _p(x), _p(residual), _p(weight), _p(bias), x.numel() // C, C, float(eps), _p(out),
              _stream())
    return out


def superpoint_matching(ref_feats, src_feats, num_correspondences, dual_normalization=True, ref_masks=None,
                        src_masks=None, normalize=False, return_matrix=False):
    """Coarse matching (model.py:141-157, superpoint_matching.py:13-51) on the hand-written fp32 path.

    ref_feats (n,d), src_feats (m,d) fp32; masks bool (True = keep) or None.  Returns a dict with
      ref_idx, src_idx int64 (num_correspondences,), scores fp32 (num_correspondences,), count int32 (1,)  [device;
      the first `count` entries are valid], ref_unit / src_unit (the F.normalize'd rows, only with normalize=True),
      matrix (n,m) (only with return_matrix=True).  No host synchronisation."""
    _check(ref_feats, 'ref_feats', torch.float32, 2)
    _check(src_feats, 'src_feats', torch.float32, 2)
    n, d = ref_feats.shape
    m, d2 = src_feats.shape
    if d != d2:
        raise RuntimeError(f'feature widths differ: {d} vs {d2}')
    dev = ref_feats.device
    keep = []
    for t, name, rows in ((ref_masks, 'ref_masks', n), (src_masks, 'src_masks', m)):
        if t is None:
            keep.append(None)
            continue
        _check(t, name, torch.bool, 1)
        if t.shape[0] != rows:
            raise RuntimeError(f'{name} must have {rows} entries, got {t.shape[0]}')
        keep.append(t.view(torch.uint8))
    K = int(num_correspondences)
    lib = _abi.load()
    nbytes = lib.geotx_spm_workspace_bytes(n, m, d, K)
    if nbytes <= 0:
        raise RuntimeError(f'superpoint_matching: bad sizes n={n} m={m} d={d} num_correspondences={K}')
    work = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    out = dict(ref_idx=torch.empty(K, dtype=torch.int64, device=dev), src_idx=torch.empty(K, dtype=torch.int64, device=dev),
               scores=torch.empty(K, dtype=torch.float32, device=dev), count=torch.empty(1, dtype=torch.int32, device=dev))
    if normalize:
        out['ref_unit'] = torch.empty_like(ref_feats)
        out['src_unit'] = torch.empty_like(src_feats)
    if return_matrix:
        out['matrix'] = torch.empty(n, m, dtype=torch.float32, device=dev)
    _abi.call('geotx_superpoint_matching', _p(ref_feats), _p(src_feats), n, m, d, _p(keep[0]), _p(keep[1]),
              1 if normalize else 0, _p(out.get('ref_unit')), _p(out.get('src_unit')), 1 if dual_normalization else 0, K,
              _p(out['ref_idx']), _p(out['src_idx']), _p(out['scores']), _p(out['count']), _p(out.get('matrix')),
              _p(work), _stream())
    return out


def point_to_node_partition(points, nodes, point_limit, return_count=False, return_points=False):
    """modules/ops/pointcloud_partition.py:61-107 on the hand-written path (no (M,N) matrix in HBM).

    points (N,3), nodes (M,3) fp32.  Returns the reference's tuple
      (point_to_node int64 (N,), [node_sizes int64 (M,) if return_count,] node_masks bool (M,),
       node_knn_indices int64 (M,K), node_knn_masks bool (M,K))
    and, with return_points=True (extension), node_knn_points (M,K,3) = padded_points[node_knn_indices]
    (model.py:105-108) as a last element.  No host synchronisation."""
    _check(points, 'points', torch.float32, 2)
    _check(nodes, 'nodes', torch.float32, 2)
    if points.shape[1] != 3 or nodes.shape[1] != 3:
        raise RuntimeError(f'points and nodes must be (*, 3), got {tuple(points.shape)} and {tuple(nodes.shape)}')
    N, M, K = points.shape[0], nodes.shape[0], int(point_limit)
    if K > N:  # torch.topk's own message (pointcloud_partition.py:93)
        raise RuntimeError(f'selected index k out of range (point_limit={K} > {N} points)')
    lib = _abi.load()
    nbytes = lib.geotx_p2n_workspace_bytes(N, M, K)
    if nbytes <= 0:
        raise RuntimeError(f'point_to_node_partition: bad sizes points={N} nodes={M} point_limit={K}')
    dev = points.device
    work = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    point_to_node = torch.empty(N, dtype=torch.int64, device=dev)
    node_sizes = torch.empty(M, dtype=torch.int64, device=dev) if return_count else None
    node_masks = torch.empty(M, dtype=torch.bool, device=dev)
    knn_indices = torch.empty(M, K, dtype=torch.int64, device=dev)
    knn_masks = torch.empty(M, K, dtype=torch.bool, device=dev)
    knn_points = torch.empty(M, K, 3, dtype=torch.float32, device=dev) if return_points else None
    _abi.call('geotx_point_to_node_partition', _p(points), N, _p(nodes), M, K, _p(point_to_node), _p(node_sizes),
              _p(node_masks), _p(knn_indices), _p(knn_masks), _p(knn_points), _p(work), _stream())
    out = (point_to_node,) + ((node_sizes,) if return_count else ()) + (node_masks, knn_indices, knn_masks)
    return out + ((knn_points,) if return_points else ())


def log_optimal_transport(scores, alpha, num_iterations, row_masks=None, col_masks=None, inf=1e12):
    """modules/sinkhorn/learnable_sinkhorn.py:21-63 in one launch: scores (B,M,N) fp32, alpha a 1-element fp32 CUDA
    tensor (the module's parameter), masks bool (True = valid) or None -> (B,M+1,N+1) fp32."""
    _check(scores, 'scores', torch.float32, 3)
    _check(alpha, 'alpha', torch.float32)
    if alpha.numel() != 1:
        raise RuntimeError(f'alpha must hold one element, got shape {tuple(alpha.shape)}')
    B, M, N = scores.shape
    masks = []
    for t, name, cols in ((row_masks, 'row_masks', M), (col_masks, 'col_masks', N)):
        if t is None:
            masks.append(None)
            continue
        _check(t, name, torch.bool, 2)
        if tuple(t.shape) != (B, cols):
            raise RuntimeError(f'{name} must be {(B, cols)}, got {tuple(t.shape)}')
        masks.append(t.view(torch.uint8))
    out = torch.empty(B, M + 1, N + 1, dtype=torch.float32, device=scores.device)
    if B == 0:
        return out
    _abi.call('geotx_log_optimal_transport', _p(scores), _p(masks[0]), _p(masks[1]), _p(alpha), int(num_iterations),
              float(inf), B, M, N, _p(out), _stream())
    return out


def weighted_procrustes(src_points, ref_points, weights=None, weight_thresh=0.0, eps=1e-5):
    """modules/registration/procrustes.py:31-59 on the device (no host SVD): src_points, ref_points (B,N,3) fp32,
    weights (B,N) fp32 or None -> transform (B,4,4) fp32 mapping src onto ref.  No host synchronisation."""
    _check(src_points, 'src_points', torch.float32, 3)
    _check(ref_points, 'ref_points', torch.float32, 3)
    if src_points.shape != ref_points.shape or src_points.shape[2] != 3:
        raise RuntimeError(f'src_points and ref_points must both be (B, N, 3), got {tuple(src_points.shape)} and '
                           f'{tuple(ref_points.shape)}')
    B, N, _ = src_points.shape
    if weights is not None:
        _check(weights, 'weights', torch.float32, 2)
        if tuple(weights.shape) != (B, N):
            raise RuntimeError(f'weights must be {(B, N)}, got {tuple(weights.shape)}')
    out = torch.empty(B, 4, 4, dtype=torch.float32, device=src_points.device)
    if B == 0:
        return out
    _abi.call('geotx_weighted_procrustes', _p(src_points), _p(ref_points), _p(weights), B, N, float(weight_thresh),
              float(eps), _p(out), _stream())
    return out


def _host_lengths(lengths, name):
    """Batch lengths as a ctypes int64 array (they live on the host in the reference: utils/data.py builds them in the
    collate function).  A CUDA tensor is accepted but costs a synchronising copy."""
    vals = [int(x) for x in (lengths.tolist() if isinstance(lengths, torch.Tensor) else lengths)]
    if any(v < 0 for v in vals):
        raise RuntimeError(f'{name} must be non-negative, got {vals}')
    return (ctypes.c_int64 * len(vals))(*vals), vals


_RN_WORK = {}
MAX_NEIGHBOR_LIMIT = 512  # GEOTX_MAX_NEIGHBOR_LIMIT (include/geotx_b200.h): neighbours K9 keeps per query


def radius_neighbors(q_points, s_points, q_lengths, s_lengths, radius, neighbor_limit, method='grid', max_cells=1 << 20):
    """extensions/cpu/radius_neighbors (ext.radius_neighbors) + modules/ops/radius_search.py:24-27 on the device.
    q_points (Nq,3), s_points (Ns,3) fp32 CUDA; q_lengths, s_lengths: the batch lengths (host).  Returns
    (neighbors int64 (Nq, neighbor_limit) padded with Ns, counts int32 (Nq,), max_count int32 (1,)); nothing
    synchronises.  method 'grid' (default) searches per-element cell lists (geotx_radius_neighbors_grid), 'brute' scans
    every support point (geotx_radius_neighbors); the results are identical."""
    _check(q_points, 'q_points', torch.float32, 2)
    _check(s_points, 's_points', torch.float32, 2)
    if q_points.shape[1] != 3 or s_points.shape[1] != 3:
        raise RuntimeError(f'points must be (*, 3), got {tuple(q_points.shape)} and {tuple(s_points.shape)}')
    ql, qv = _host_lengths(q_lengths, 'q_lengths')
    sl, sv = _host_lengths(s_lengths, 's_lengths')
    if len(qv) != len(sv) or sum(qv) != q_points.shape[0] or sum(sv) != s_points.shape[0]:
        raise RuntimeError(f'lengths {qv} / {sv} do not match {q_points.shape[0]} query and {s_points.shape[0]} support points')
    limit = int(neighbor_limit)
    dev = q_points.device
    neighbors = torch.empty(q_points.shape[0], limit, dtype=torch.int64, device=dev)
    counts = torch.empty(q_points.shape[0], dtype=torch.int32, device=dev)
    max_count = torch.empty(1, dtype=torch.int32, device=dev)
    if method == 'brute' or not float(radius) > 0.0:
        _abi.call('geotx_radius_neighbors', _p(q_points), _p(s_points), ql, sl, len(qv), float(radius), limit,
                  _p(neighbors), _p(counts), _p(max_count), _stream())
        return neighbors, counts, max_count
    if method != 'grid':
        raise RuntimeError(f"radius_neighbors: method must be 'grid' or 'brute', got {method!r}")
    cap = 1024
    while cap < sum(sv):
        cap <<= 1
    key = (dev, cap, int(max_cells))
    if key not in _RN_WORK:
        nbytes = _abi.load().geotx_radius_neighbors_workspace_bytes(cap, int(max_cells))
        if nbytes <= 0:
            raise RuntimeError(f'radius_neighbors: bad max_cells={max_cells}')
        _RN_WORK[key] = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    _abi.call('geotx_radius_neighbors_grid', _p(q_points), _p(s_points), ql, sl, len(qv), float(radius), limit,
              _p(neighbors), _p(counts), _p(max_count), _p(_RN_WORK[key]), cap, int(max_cells), _stream())
    return neighbors, counts, max_count


_GS_WORK = {}


def grid_subsample(points, lengths, voxel_size, max_cells=1 << 24):
    """extensions/cpu/grid_subsampling (ext.grid_subsampling) on the device.  points (N,3) fp32 CUDA, lengths: the batch
    lengths (host).  Returns (s_points (N,3) buffer whose first sum(s_lengths) rows are valid, s_lengths int64 (B,) on
    the device; -1 marks an element whose bounding grid exceeds max_cells).  Nothing synchronises.  The workspace
    (32 B per cell) is kept per device and size."""
    _check(points, 'points', torch.float32, 2)
    if points.shape[1] != 3:
        raise RuntimeError(f'points must be (N, 3), got {tuple(points.shape)}')
    ln, vals = _host_lengths(lengths, 'lengths')
    if sum(vals) != points.shape[0] or any(v < 1 for v in vals):
        raise RuntimeError(f'lengths {vals} must be positive and sum to {points.shape[0]}')
    lib = _abi.load()
    key = (points.device, int(max_cells))
    if key not in _GS_WORK:
        nbytes = lib.geotx_grid_subsample_workspace_bytes(int(max_cells))
        if nbytes <= 0:
            raise RuntimeError(f'grid_subsample: bad max_cells={max_cells}')
        _GS_WORK[key] = torch.empty(nbytes, dtype=torch.uint8, device=points.device)
    s_points = torch.empty_like(points)
    s_lengths = torch.empty(len(vals), dtype=torch.int64, device=points.device)
    _abi.call('geotx_grid_subsample', _p(points), ln, len(vals), float(voxel_size), _p(s_points), _p(s_lengths),
              _p(_GS_WORK[key]), int(max_cells), _stream())
    return s_points, s_lengths


def local_global_registration(ref_knn_points, src_knn_points, ref_knn_masks, src_knn_masks, score_mat, global_scores,
                              k, acceptance_radius, mutual=True, confidence_threshold=0.05,
                              correspondence_threshold=3, num_refinement_steps=5):
    """modules/geotransformer/local_global_registration.py:145-235 on the device.  ref_knn_points, src_knn_points
    (P,K,3) fp32; ref_knn_masks, src_knn_masks (P,K) bool; score_mat (P,K,K) fp32 log-likelihoods whose last dimension
    is dense (the reference's matching_scores[:, :-1, :-1] view is taken as it is); global_scores (P,) fp32 or None.
    Returns (ref_corr_points (cap,3), src_corr_points (cap,3), corr_scores (cap,), count int32 (1,), transform (4,4))
    with cap = 2*P*K*k; the first count[0] rows are the correspondences in torch.nonzero order.  Nothing
    synchronises."""
    _check(ref_knn_points, 'ref_knn_points', torch.float32, 3)
    _check(src_knn_points, 'src_knn_points', torch.float32, 3)
    P, K, _ = ref_knn_points.shape
    if tuple(ref_knn_points.shape) != (P, K, 3) or tuple(src_knn_points.shape) != (P, K, 3):
        raise RuntimeError(f'knn points must both be (P, K, 3), got {tuple(ref_knn_points.shape)} and '
                           f'{tuple(src_knn_points.shape)}')
    masks = []
    for name, m in (('ref_knn_masks', ref_knn_masks), ('src_knn_masks', src_knn_masks)):
        _check(m, name, None, 2)
        if m.dtype not in (torch.bool, torch.uint8) or tuple(m.shape) != (P, K):
            raise RuntimeError(f'{name} must be bool {(P, K)}, got {m.dtype} {tuple(m.shape)}')
        masks.append(m.view(torch.uint8) if m.dtype == torch.bool else m)
    if not isinstance(score_mat, torch.Tensor) or not score_mat.is_cuda or score_mat.dtype != torch.float32:
        raise RuntimeError('score_mat must be a float32 CUDA tensor')
    if tuple(score_mat.shape) != (P, K, K) or (K > 1 and score_mat.stride(2) != 1):
        raise RuntimeError(f'score_mat must be {(P, K, K)} with a dense last dimension, got {tuple(score_mat.shape)} '
                           f'strides {score_mat.stride()}')
    if global_scores is not None:
        _check(global_scores, 'global_scores', torch.float32, 1)
        if global_scores.shape[0] != P:
            raise RuntimeError(f'global_scores must be ({P},), got {tuple(global_scores.shape)}')
    dev = ref_knn_points.device
    cap = 2 * P * K * int(k)
    ref_corr = torch.empty(cap, 3, dtype=torch.float32, device=dev)
    src_corr = torch.empty(cap, 3, dtype=torch.float32, device=dev)
    scores = torch.empty(cap, dtype=torch.float32, device=dev)
    count = torch.empty(1, dtype=torch.int32, device=dev)
    transform = torch.empty(4, 4, dtype=torch.float32, device=dev)
    if P == 0:
        raise RuntimeError('local_global_registration: no patch pairs')
    nbytes = _abi.load().geotx_lgr_workspace_bytes(P, K)
    work = torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=dev)
    _abi.call('geotx_local_global_registration', _p(ref_knn_points), _p(src_knn_points), _p(masks[0]), _p(masks[1]),
              _p(score_mat), int(score_mat.stride(0)), int(score_mat.stride(1)), _p(global_scores), P, K, int(k),
              float(acceptance_radius), int(bool(mutual)), float(confidence_threshold), int(correspondence_threshold),
              int(num_refinement_steps), _p(ref_corr), _p(src_corr), _p(scores), _p(count), _p(transform), _p(work),
              _stream())
    return ref_corr, src_corr, scores, count, transform


def patch_scores(ref_feats, src_feats, ref_node_knn_indices, src_node_knn_indices, ref_node_corr_indices,
                 src_node_corr_indices):
    """experiments/*/model.py:169-170,176-179,187-188 in one kernel: ref_feats (Nr,C), src_feats (Ns,C) fp32;
    *_node_knn_indices (M,K) int64 (Nr / Ns = padded slot); *_node_corr_indices (P,) int64 -> scores (P,K,K) fp32 =
    gathered patch feature products / sqrt(C).  Nothing synchronises."""
    _check(ref_feats, 'ref_feats', torch.float32, 2)
    _check(src_feats, 'src_feats', torch.float32, 2)
    C = ref_feats.shape[1]
    if src_feats.shape[1] != C or C % 4:
        raise RuntimeError(f'feature widths must be equal and a multiple of 4, got {C} and {src_feats.shape[1]}')
    for name, t, dims in (('ref_node_knn_indices', ref_node_knn_indices, 2), ('src_node_knn_indices', src_node_knn_indices, 2),
                          ('ref_node_corr_indices', ref_node_corr_indices, 1), ('src_node_corr_indices', src_node_corr_indices, 1)):
        _check(t, name, torch.int64, dims)
    K = ref_node_knn_indices.shape[1]
    P = ref_node_corr_indices.shape[0]
    if src_node_knn_indices.shape[1] != K or src_node_corr_indices.shape[0] != P:
        raise RuntimeError('ref and src sides must have the same patch size and the same number of correspondences')
    out = torch.empty(P, K, K, dtype=torch.float32, device=ref_feats.device)
    if P == 0:
        return out
    _abi.call('geotx_patch_scores', _p(ref_feats), int(ref_feats.shape[0]), _p(src_feats), int(src_feats.shape[0]), C,
              _p(ref_node_knn_indices), _p(src_node_knn_indices), K, _p(ref_node_corr_indices),
              _p(src_node_corr_indices), P, _p(out), _stream())
    return out


class EmbeddingTables:
    """Sampled proj_d(sinus(x)) + b_d and proj_a(sinus(x)) + b_a for the tabulated embedding (geotx_gse_embed_table)."""
    __slots__ = ('buf', 'C', 'h_a', 'h_d', 'rows_a', 'rows_d', 'overflow', 'weights')

    def __init__(self, buf, C, h_a, h_d, rows_a, rows_d, overflow, weights):
        self.buf, self.C, self.h_a, self.h_d, self.weights = buf, C, h_a, h_d, weights
        self.rows_a, self.rows_d, self.overflow = rows_a, rows_d, overflow


def gse_build_tables(div_term, w_d, b_d, w_a, b_a, factor_a, d_index_max, h_a=1.0 / 16, h_d=0.25):
    """Tables for `gse_embed_table` (K2t: round 1's default embedding, now the `GEOTX_EMBED=table` variant): angle arguments lie in [0, pi * factor_a],
    distance arguments in [0, d_index_max] are read from the table, larger ones are evaluated exactly (slowly)."""
    import math
    for name, t in (('div_term', div_term), ('w_d', w_d), ('b_d', b_d), ('w_a', w_a), ('b_a', b_a)):
        _check(t, name, torch.float32)
    C = w_d.shape[0]
    rows_a = int(math.ceil(math.pi * float(factor_a) / h_a)) + 3
    rows_d = int(math.ceil(float(d_index_max) / h_d)) + 5
    nbytes = _abi.load().geotx_gse_table_bytes(C, rows_a, rows_d)
    if nbytes <= 0:
        raise RuntimeError(f'gse_build_tables: hidden_dim {C} is not supported by the tabulated embedding')
    buf = torch.empty(nbytes, dtype=torch.uint8, device=w_d.device)
    overflow = torch.zeros(1, dtype=torch.int32, device=w_d.device)
    _abi.call('geotx_gse_build_tables', _p(div_term), _p(w_d), _p(b_d), _p(w_a), _p(b_a), C, float(h_a), float(h_d),
              rows_a, rows_d, _p(buf), _stream())
    return EmbeddingTables(buf, C, float(h_a), float(h_d), rows_a, rows_d, overflow,
                           (div_term, w_d.contiguous(), b_d.contiguous()))


def gse_embed_table(geom, tables, sigma_d, factor_a, row_begin=0, row_count=None, out=None):
    """Tabulated embedding (K2t, variant): -> E (B,row_count,n,C) f16; tables.overflow becomes 1 if an argument fell
    outside a table.  geotransformer.py:57-72."""
    row_begin, row_count = _slab(geom, row_begin, row_count)
    if out is None:
        out = torch.empty(geom.B, row_count, geom.n, tables.C, dtype=torch.float16, device=geom.buf.device)
    else:
        _check(out, 'out', torch.float16)
    _abi.call('geotx_gse_embed_table', _p(geom.buf), geom.B, geom.n, geom.k, row_begin, row_count, _p(tables.buf),
              tables.C, tables.rows_a, tables.rows_d, tables.h_a, tables.h_d, float(sigma_d), float(factor_a),
              _p(tables.weights[0]), _p(tables.weights[1]), _p(tables.weights[2]), _p(out), _p(tables.overflow), _stream())
    return out


class ChebCoefficients:
    """Chebyshev coefficients of proj_a(sinus(x)) and proj_d(sinus(x)) + b_d + b_a for geotx_gse_embed_cheb (K2c)."""
    __slots__ = ('buf', 'C', 'angle_index_max', 'ka_tiles', 'stats', 'overflow', 'plain', '_report')
    SPLIT_COEFFICIENTS = 80  # distance sets 1 (32 coefficients) and 2 (48): what the distance split of K3r consumes

    def __init__(self, buf, C, angle_index_max, ka_tiles, stats, overflow, plain):
        self.buf, self.C, self.angle_index_max, self.ka_tiles = buf, C, angle_index_max, ka_tiles
        self.stats, self.overflow, self.plain, self._report = stats, overflow, plain, None

    @staticmethod
    def split_columns(max_d_set):
        """Columns of `u` K3r expects for a call prepared for distance sets up to `max_d_set`: sets 1 and 2 (32 + 48), and
        set 3 (64 more) when the call may need it."""
        return 80 if max_d_set <= 2 else 144

    def distance_split_rows(self, max_d_set=2):
        """(80 | 144, C) fp32: the unrounded Chebyshev coefficients G_d of distance sets 1, 2 (and 3), row m = coefficient m."""
        r0 = 16 * self.ka_tiles + 16  # behind the angle set and distance set 0
        return self.plain[r0:r0 + self.split_columns(max_d_set)]

    def report(self):
        """One host read per weight set: [{'range', 'coefficients', 'max_error', 'max_abs_f', 'max_abs_g'}, ...] for the
        angle set and the six distance sets (errors measured by the build kernel in fp64 against the module's own
        weights, at 256 points per channel between the quadrature nodes)."""
        if self._report is None:
            lib = _abi.load()
            st = self.stats.cpu().view(7, 4).tolist()
            rows = [dict(range=self.angle_index_max, coefficients=16 * self.ka_tiles)]
            rows += [dict(range=float(lib.geotx_gse_cheb_d_range(s)), coefficients=16 * (1, 2, 3, 4, 6, 8)[s])
                     for s in range(6)]
            for r, (err, fmax, gmax, _) in zip(rows, st):
                r.update(max_error=err, max_abs_f=fmax, max_abs_g=gmax)
            self._report = rows
        return self._report


def gse_cheb_ka_tiles(angle_index_max):
    """16-coefficient tiles the angle term needs for arguments in [0, angle_index_max]; 0 = not covered by K2c."""
    return int(_abi.load().geotx_gse_cheb_ka_tiles(float(angle_index_max)))


def gse_cheb_max_d_set(d_index_max):
    """Smallest prebuilt distance set whose range reaches d_index_max (5 = the largest, 210 index units)."""
    lib = _abi.load()
    for s in range(6):
        if float(d_index_max) <= float(lib.geotx_gse_cheb_d_range(s)):
            return s
    return 5


def gse_build_cheb(div_term, w_d, b_d, w_a, b_a, factor_a):
    """Coefficients for `gse_embed_cheb`, built on the device in fp64 from the module's weights (once per weight set).
    geotransformer.py:60-64 + positional_embedding.py:28-32."""
    import math
    for name, t in (('div_term', div_term), ('w_d', w_d), ('b_d', b_d), ('w_a', w_a), ('b_a', b_a)):
        _check(t, name, torch.float32)
    C = w_d.shape[0]
    angle_index_max = math.pi * float(factor_a) * (1.0 + 1e-6)
    lib = _abi.load()
    ka = int(lib.geotx_gse_cheb_ka_tiles(angle_index_max))
    nbytes = lib.geotx_gse_cheb_bytes(C, ka) if ka > 0 else 0
    if nbytes <= 0:
        raise RuntimeError(f'gse_build_cheb: hidden_dim {C} / angle range {angle_index_max:.1f} is not supported by the '
                           f'low-rank embedding kernel')
    dev = w_d.device
    buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    stats = torch.empty(28, dtype=torch.float32, device=dev)
    overflow = torch.zeros(1, dtype=torch.int32, device=dev)
    plain = torch.empty(int(lib.geotx_gse_cheb_plain_rows(ka)), C, dtype=torch.float32, device=dev)
    _abi.call('geotx_gse_build_cheb', _p(div_term), _p(w_d.contiguous()), _p(b_d.contiguous()), _p(w_a.contiguous()),
              _p(b_a.contiguous()), C, angle_index_max, ka, _p(buf), _p(stats), _p(plain), _stream())
    return ChebCoefficients(buf, C, angle_index_max, ka, stats, overflow, plain)


def gse_embed_cheb(geom, coef, sigma_d, factor_a, reduce_mean, out_dtype, max_d_set=5, row_begin=0, row_count=None,
                   out=None):
    """Low-rank embedding K2c: -> E (B,row_count,n,C) f16 / bf16; coef.overflow becomes 1 if a distance fell beyond the
    range of set `max_d_set`.  geotransformer.py:57-72."""
    row_begin, row_count = _slab(geom, row_begin, row_count)
    if out_dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError(f'gse_embed_cheb: out_dtype must be float16 or bfloat16, got {out_dtype}')
    if out is None:
        out = torch.empty(geom.B, row_count, geom.n, coef.C, dtype=out_dtype, device=geom.buf.device)
    else:
        _check(out, 'out', out_dtype)
    _abi.call('geotx_gse_embed_cheb', _p(geom.buf), geom.B, geom.n, geom.k, row_begin, row_count, _p(coef.buf), coef.C,
              coef.angle_index_max, coef.ka_tiles, int(max_d_set), float(sigma_d), float(factor_a), int(bool(reduce_mean)),
              _abi.GEOTX_F16 if out_dtype == torch.float16 else _abi.GEOTX_BF16, _p(out), _p(coef.overflow), _stream())
    return out


def rpe_scores_cheb(geom, coef, qt, sigma_d, factor_a, reduce_mean=False, max_d_set=5, row_begin=0, row_count=None,
                    out=None, u=None):
    """K3r: rpe_transformer.py:54-56 (proj_p folded into qt, see rpe_scores) fused with the embedding
    (geotransformer.py:57-72): scores_p (B,H,n,n) fp32 straight from the geometry; E is rebuilt tile by tile in
    registers and never stored.  qt: logical (B,H,n,C) fp32 with a dense last dimension (any strides over b, h, i).
    u (optional): logical (B,H,n,M) fp32, u[b,h,i,:] = coef.distance_split_rows(max_d_set) @ qt[b,h,i,:], M = 80 for
    max_d_set <= 2, else 144 (the distance split: the distance term's share of the scores becomes a 32 / 48 / 64-term
    product instead of a C-term one; used for H <= 4, k = 3, 'max' and distance sets <= 3, ignored otherwise)."""
    row_begin, row_count = _slab(geom, row_begin, row_count)
    if not isinstance(qt, torch.Tensor) or not qt.is_cuda or qt.dtype != torch.float32 or qt.dim() != 4:
        raise RuntimeError('qt must be a 4-d float32 CUDA tensor (B,H,n,C)')
    B, H, n, C = qt.shape
    if (B, n, C) != (geom.B, geom.n, coef.C) or qt.stride(3) != 1:
        raise RuntimeError(f'qt {tuple(qt.shape)} (strides {qt.stride()}) does not match the geometry '
                           f'(B={geom.B}, n={geom.n}) / hidden_dim {coef.C}, or its last dimension is not dense')
    if out is None:
        out = torch.empty(B, H, n, n, dtype=torch.float32, device=qt.device)
    else:
        _check(out, 'out', torch.float32, 4)
        if tuple(out.shape) != (B, H, n, n):
            raise RuntimeError(f'out must be {(B, H, n, n)}, got {tuple(out.shape)}')
    us = (0, 0, 0)
    if u is not None:
        if not isinstance(u, torch.Tensor) or not u.is_cuda or u.dtype != torch.float32 or \
                tuple(u.shape) != (B, H, n, coef.split_columns(int(max_d_set))) or u.stride(3) != 1:
            raise RuntimeError(f'u must be a float32 CUDA tensor of logical shape {(B, H, n, coef.split_columns(int(max_d_set)))} with a '
                               f'dense last dimension')
        us = (int(u.stride(0)), int(u.stride(1)), int(u.stride(2)))
    _abi.call('geotx_rpe_scores_cheb', _p(geom.buf), B, n, geom.k, row_begin, row_count, _p(coef.buf), C,
              coef.angle_index_max, coef.ka_tiles, int(max_d_set), float(sigma_d), float(factor_a), int(bool(reduce_mean)),
              _p(qt), int(qt.stride(0)), int(qt.stride(1)), int(qt.stride(2)), H, _p(u), us[0], us[1], us[2], _p(out),
              _p(coef.overflow), _stream())
    return out


def to_half(t):
    """fp32 CUDA tensor -> contiguous IEEE-half copy (the weight format of linear_f16 / post_attention)."""
    _check(t, 't', torch.float32)
    t = t.contiguous()
    out = torch.empty(t.shape, dtype=torch.float16, device=t.device)
    _abi.call('geotx_f32_to_f16', _p(t), _p(out), t.numel(), _stream())
    return out


def pack_linear(w):
    """nn.Linear weight (N, K) fp32 CUDA -> its packed half image for linear_f16 / post_attention: same shape and dtype
    float16, but the element order is geotx_pack_linear_weight's (K chunks of 32, chunk-major, swizzled 16-byte units;
    csrc/dense_tile.cuh) -- an opaque operand, not a matrix to index."""
    _check(w, 'w', torch.float32, 2)
    w = w.contiguous()
    N, K = w.shape
    if N % 8 or K % 32:
        raise RuntimeError(f'pack_linear: weight {tuple(w.shape)} needs N % 8 == 0 and K % 32 == 0')
    out = torch.empty(N, K, dtype=torch.float16, device=w.device)
    _abi.call('geotx_pack_linear_weight', _p(w), N, K, _p(out), _stream())
    return out


def _rows2d(t, name, cols):
    """(…, cols) fp32 CUDA tensor whose leading dimensions flatten into equally strided rows -> (rows, row stride)."""
    if not isinstance(t, torch.Tensor) or not t.is_cuda or t.dtype != torch.float32 or t.dim() < 1:
        raise RuntimeError(f'{name} must be a float32 CUDA tensor (ref-geotransformer_b200 has no CPU path)')
    _same_device(t, name)
    if t.shape[-1] != cols or (cols > 1 and t.stride(-1) != 1):
        raise RuntimeError(f'{name}: last dimension must be dense and of size {cols}, got {tuple(t.shape)} strides {t.stride()}')
    rows = t.numel() // cols
    lead = [(s, st) for s, st in zip(t.shape[:-1], t.stride()[:-1]) if s != 1]
    ld = lead[-1][1] if lead else cols
    for (s0, st0), (s1, st1) in zip(lead[:-1], lead[1:]):
        if st0 != s1 * st1:
            raise RuntimeError(f'{name}: rows are not equally strided ({tuple(t.shape)}, strides {t.stride()})')
    return rows, ld


def linear_supported(K, N):
    return K >= 32 and K % 32 == 0 and K <= 2048 and N >= 64 and N % 64 == 0


def linear_f16(x, w_half, bias, out=None, relu=False, half_cols=0):
    """y = x W^T + b (nn.Linear semantics) on the hand-written mma.sync kernel K13: x (..., K) fp32 with equally strided
    rows, w_half (N, K) the PACKED half image of the weight (`pack_linear(weight)`), bias (N,) fp32 or None -> (..., N)
    fp32 (into `out` if given; rows of `out` may be strided).  K % 32 == 0, K <= 2048, N % 64 == 0.
    half_cols > 0 (a multiple of 64; geotx_linear_f16_mixed): returns (y_half (..., half_cols) float16, y (..., N -
    half_cols) fp32 or None when nothing is left) -- the first output columns rounded to half (saturating), for
    `attention`, the rest as usual."""
    if not isinstance(w_half, torch.Tensor) or not w_half.is_cuda or w_half.dtype != torch.float16 or w_half.dim() != 2 \
            or not w_half.is_contiguous():
        raise RuntimeError('w_half must be a contiguous 2-d float16 CUDA tensor (N, K)')
    N, K = w_half.shape
    if not linear_supported(K, N):
        raise RuntimeError(f'linear_f16: K={K} must be a multiple of 32 (<= 2048) and N={N} a multiple of 64')
    T, ld_x = _rows2d(x, 'x', K)
    if bias is not None:
        _check(bias, 'bias', torch.float32, 1)
    if half_cols:
        if out is not None or half_cols % 64 or not 0 < half_cols <= N:
            raise RuntimeError(f'linear_f16: half_cols={half_cols} must be a multiple of 64 in (0, {N}] and excludes out=')
        y_half = torch.empty(*x.shape[:-1], half_cols, dtype=torch.float16, device=x.device)
        rest = torch.empty(*x.shape[:-1], N - half_cols, dtype=torch.float32, device=x.device) if half_cols < N else None
        _abi.call('geotx_linear_f16_mixed', _p(x), ld_x, T, K, _p(w_half), _p(bias), N, int(bool(relu)), _p(y_half),
                  half_cols, half_cols, _p(rest), N - half_cols, _stream())
        return y_half, rest
    if out is None:
        out = torch.empty(*x.shape[:-1], N, dtype=torch.float32, device=x.device)
    T_out, ld_y = _rows2d(out, 'out', N)
    if T_out != T:
        raise RuntimeError(f'out has {T_out} rows, x has {T}')
    _abi.call('geotx_linear_f16', _p(x), ld_x, T, K, _p(w_half), _p(bias), N, int(bool(relu)), _p(out), ld_y, _stream())
    return out


def post_attention(hidden, x_in, w_lin, b_lin, g1, be1, eps1, w_exp, b_exp, w_sq, b_sq, g2, be2, eps2, out=None):
    """K4f: LayerNorm(x1 + squeeze(ReLU(expand(x1)))) with x1 = LayerNorm(linear(hidden) + x_in) in one kernel
    (rpe_transformer.py:101-103 / vanilla_transformer.py:99-101 + output_layer.py:15-21).  hidden, x_in (..., C) fp32
    with equally strided rows, C in {128, 256}; w_* packed half images (`pack_linear`) of the nn.Linear weights; the rest fp32."""
    C = hidden.shape[-1]
    T, ld_h = _rows2d(hidden, 'hidden', C)
    T2, ld_x = _rows2d(x_in, 'x_in', C)
    if T2 != T:
        raise RuntimeError(f'x_in has {T2} rows, hidden has {T}')
    for name, t, shape in (('w_lin', w_lin, (C, C)), ('w_exp', w_exp, (2 * C, C)), ('w_sq', w_sq, (C, 2 * C))):
        if not isinstance(t, torch.Tensor) or not t.is_cuda or t.dtype != torch.float16 or tuple(t.shape) != shape \
                or not t.is_contiguous():
            raise RuntimeError(f'{name} must be a contiguous float16 CUDA tensor of shape {shape}')
    for name, t, size in (('b_lin', b_lin, C), ('g1', g1, C), ('be1', be1, C), ('b_exp', b_exp, 2 * C), ('b_sq', b_sq, C),
                          ('g2', g2, C), ('be2', be2, C)):
        _check(t, name, torch.float32, 1)
        if t.numel() != size:
            raise RuntimeError(f'{name} must hold {size} elements, got {t.numel()}')
    if out is None:
        out = torch.empty(hidden.shape, dtype=torch.float32, device=hidden.device)
    T3, ld_o = _rows2d(out, 'out', C)
    if T3 != T:
        raise RuntimeError(f'out has {T3} rows, hidden has {T}')
    _abi.call('geotx_post_attention', _p(hidden), ld_h, _p(x_in), ld_x, T, C, _p(w_lin), _p(b_lin), _p(g1), _p(be1),
              float(eps1), _p(w_exp), _p(b_exp), _p(w_sq), _p(b_sq), _p(g2), _p(be2), float(eps2), _p(out), ld_o, _stream())
    return out
