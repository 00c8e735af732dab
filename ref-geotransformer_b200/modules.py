"""Host-side mirror of the reference's module API for the hot path.

Same class names, constructor signatures, forward signatures, error behaviour and state_dict keys as
/root/reference/geotransformer/modules/geotransformer/geotransformer.py and
/root/reference/geotransformer/modules/transformer/{rpe_transformer,vanilla_transformer,output_layer,
conditional_transformer,positional_embedding}.py, so `load_state_dict(strict=True)` of a reference checkpoint
works and the classes drop into `GeoTransformer.forward` (experiments/*/model.py:35-45,135-140).

The heavy math runs in the sm_100a kernels behind include/geotx_b200.h (see ops.py).  The small dense layers
(C x C Linears, LayerNorm, ReLU: <0.5 % of the path's flops) are issued through torch on
the same stream.  Forward-only: calling with autograd enabled on anything that requires grad raises."""
import math
import warnings
import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops

# slab of E (bytes) written by the embedding kernel and consumed from L2 by the first block's scores kernel; the
# B200 L2 is 126 MB.  Measured sweep in DESIGN.md section 5.
DEFAULT_L2_SLAB_BYTES = 0


def _joint_view(a, b):
    """(1, n+m, C) view over two (1, n, C) / (1, m, C) tensors when b starts where a ends INSIDE ONE storage (two slices
    of one token buffer); None otherwise.  Comparing data pointers alone is not enough: two separate allocations that
    the caching allocator happened to place back to back would pass, and as_strided over the first would then run past
    its storage."""
    if a.dim() != 3 or b.dim() != 3 or a.shape[0] != 1 or b.shape[0] != 1 or a.shape[2] != b.shape[2]:
        return None
    if a.dtype != b.dtype or not a.is_contiguous() or not b.is_contiguous():
        return None
    sa, sb = a.untyped_storage(), b.untyped_storage()
    if sa.data_ptr() != sb.data_ptr() or a.storage_offset() + a.numel() != b.storage_offset():
        return None
    if (b.storage_offset() + b.numel()) * a.element_size() > sa.nbytes():
        return None
    return torch.as_strided(a, (1, a.shape[1] + b.shape[1], a.shape[2]), a.stride())

_ACTIVATIONS = {  # geotransformer/modules/layers/factory.py:18-27,71-80
    'ReLU': nn.ReLU, 'LeakyReLU': nn.LeakyReLU, 'ELU': nn.ELU, 'GELU': nn.GELU, 'Sigmoid': nn.Sigmoid,
    'Softplus': nn.Softplus, 'Tanh': nn.Tanh, 'Identity': nn.Identity,
}


def build_act_layer(act_cfg):
    if act_cfg is None:
        return nn.Identity()
    if isinstance(act_cfg, str):
        name, kwargs = act_cfg, {}
    else:
        kwargs = dict(act_cfg)
        name = kwargs.pop('type')
    assert name in _ACTIVATIONS, f'Illegal activation: {name}.'
    if name == 'LeakyReLU' and 'negative_slope' not in kwargs:
        kwargs['negative_slope'] = 0.2
    return _ACTIVATIONS[name](**kwargs)


def build_dropout_layer(p, **kwargs):
    # factory.py:48-53.  Dropout only acts in training mode, and this implementation is forward/inference-only,
    # so a non-zero p is kept as an nn.Dropout module (identity under eval()) for API parity.
    if p is None or p == 0:
        return nn.Identity()
    return nn.Dropout(p=p, **kwargs)


class _dense_precision:
    """The small dense layers (C x C Linears; <0.5 % of the path's flops) go through cuBLASLt.  On the bf16 / tensor-core
    path they run as TF32 tensor-core GEMMs (10-bit mantissa, fp32 accumulate; the module's outputs stay within 1e-4 of
    the reference); with precision='fp32' they stay exact fp32.  Restores the caller's global setting on exit."""

    def __init__(self, allow_tf32):
        self.allow = allow_tf32

    def __enter__(self):
        self.prev = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = self.allow
        return self

    def __exit__(self, *exc):
        torch.backends.cuda.matmul.allow_tf32 = self.prev
        return False


def _forward_only_guard(module, *tensors):
    if not torch.is_grad_enabled():
        return
    needs = any(t is not None and t.requires_grad for t in tensors) or \
        any(p.requires_grad for p in module.parameters())
    if needs:
        raise RuntimeError(
            f'{type(module).__name__} (ref-geotransformer_b200) is forward-only: its CUDA kernels define no backward. '
            'Call it under torch.no_grad() (as the reference tester does, engine/single_tester.py:42-43) or '
            'freeze the parameters and inputs.')


def _dense_own():
    """GEOTX_DENSE: 'own' (default) routes the token-wise Linears / LayerNorms of the fused schedule through the
    hand-written kernels K13 (geotx_linear_f16) and K4f (geotx_post_attention); 'cublas' keeps round 1's cuBLASLt TF32
    GEMMs + K4 (also what shapes outside the kernels' tile constraints take)."""
    mode = os.environ.get('GEOTX_DENSE', 'own')
    if mode not in ('own', 'cublas'):
        raise ValueError(f"GEOTX_DENSE must be 'own' or 'cublas', got {mode!r}")
    return mode == 'own'


def _half_of(owner, name, t):
    """Packed half image (ops.pack_linear: the operand format of K13 / K4f) of weight `t`, cached on `owner` and rebuilt
    when `t` changes (new storage or in-place update)."""
    cache = owner.__dict__.setdefault('_half_cache', {})
    key = (t.data_ptr(), t._version)
    ent = cache.get(name)
    if ent is None or ent[0] != key:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError('a weight changed since its packed half copy was made; run one forward outside CUDA graph '
                               'capture first')
        ent = (key, ops.pack_linear(t.detach()))
        # The copy is made on whatever stream is current (the second cloud's in_proj runs on a side stream) and then
        # used from every stream: wait for it once, here, instead of ordering every later use.
        torch.cuda.current_stream(t.device).synchronize()
        cache[name] = ent
    return ent[1]


def _linear(x, owner, name, weight, bias, out=None, half_cols=0):
    """nn.Linear on K13 when the shape fits its tiles and GEOTX_DENSE allows it, else F.linear.  half_cols > 0: returns
    (the first half_cols output columns as float16 -- the q | k | v operands of the attention kernel --, the remaining
    columns as fp32 or None), see ops.linear_f16."""
    N, K = weight.shape
    if _dense_own() and x.is_cuda and x.dtype == torch.float32 and ops.linear_supported(K, N) and x.stride(-1) == 1 \
            and half_cols % 64 == 0:
        return ops.linear_f16(x, _half_of(owner, name, weight), bias, out=out, half_cols=half_cols)
    y = F.linear(x, weight, bias)
    if half_cols:
        return y[..., :half_cols].clamp(-65504.0, 65504.0).to(torch.float16), (y[..., half_cols:] if half_cols < N else None)
    if out is not None:
        out.copy_(y)
        return out
    return y


def _attention_f16():
    """GEOTX_ATT: 'f16' (default) has the fused schedule's projections write q | k | v as half for the pipelined
    attention kernel K3h (geotx_attention_f16); 'f32' keeps fp32 q | k | v and K3b, which rounds them itself."""
    mode = os.environ.get('GEOTX_ATT', 'f16')
    if mode not in ('f16', 'f32'):
        raise ValueError(f"GEOTX_ATT must be 'f16' or 'f32', got {mode!r}")
    return mode == 'f16'


class SinusoidalPositionalEmbedding(nn.Module):
    """positional_embedding.py:8-34.  Only the `div_term` buffer matters on the fused path (the kernels read it
    instead of recomputing it); forward() is kept for API parity and runs as plain device ops."""

    def __init__(self, d_model):
        super().__init__()
        if d_model % 2 != 0:
            raise ValueError(f'Sinusoidal positional encoding with odd d_model: {d_model}')
        self.d_model = d_model
        div_indices = torch.arange(0, d_model, 2).float()
        div_term = torch.exp(div_indices * (-math.log(10000.0) / d_model))
        self.register_buffer('div_term', div_term)

    def forward(self, emb_indices):
        shape = emb_indices.shape
        omegas = emb_indices.reshape(-1, 1) * self.div_term.view(1, -1)
        return torch.stack([torch.sin(omegas), torch.cos(omegas)], dim=2).view(*shape, self.d_model).detach()


class GeometricStructureEmbedding(nn.Module):
    """geotransformer.py:9-72.  `precision`: 'bf16' = 16-bit storage of E through the tensor-core kernels (hidden_dim
    128/256), 'fp32' = CUDA-core kernel (any even hidden_dim); 'auto' picks the former when the hidden size allows it.
    `embed_kernel` (attribute; default from GEOTX_EMBED, 'cheb') selects the kernel behind a 16-bit E:
      'cheb'    K2c, the low-rank (Chebyshev) contraction on mma.sync: any angle_k, 'max' and 'mean', f16 / bf16;
                its truncation + rounding error is MEASURED against this module's weights when the coefficients are
                built (`cheb_report()`), and a weight set that misses `cheb_error_budget` falls back to 'tensor' with a
                warning (never to a CPU path);
      'table'   K2t, the sampled tables of round 1 (hidden_dim 128/256, angle_k 3, 'max', f16 only);
      'tensor'  K2, the full-rank tcgen05 contraction (also used for fp32 output and the fused first-block scores).
    `cheb_d_index_max` is the largest distance / sigma_d K2c / K3r are prepared for: it sizes their shared memory (64
    index units -- 12.8 m at sigma_d = 0.2, 307 m at KITTI's 4.8 -- keeps two CTAs per SM; at most 210); below it the
    kernel itself picks the shortest expansion that covers the cloud at hand.  A cloud that exceeds it raises the
    kernel's overflow flag: the first eager call of every shape reads the flag, widens `cheb_d_index_max` to 210 and
    repeats the call; CUDA-graph replays check the flag of the PREVIOUS replay (GraphedGeometricTransformer).
    `table_d_index_max`: reach of K2t's table."""

    def __init__(self, hidden_dim, sigma_d, sigma_a, angle_k, reduction_a='max', precision='auto'):
        super().__init__()
        self.sigma_d = sigma_d
        self.sigma_a = sigma_a
        self.factor_a = 180.0 / (self.sigma_a * math.pi)
        self.angle_k = angle_k
        self.embedding = SinusoidalPositionalEmbedding(hidden_dim)
        self.proj_d = nn.Linear(hidden_dim, hidden_dim)
        self.proj_a = nn.Linear(hidden_dim, hidden_dim)
        self.reduction_a = reduction_a
        if self.reduction_a not in ['max', 'mean']:
            raise ValueError(f'Unsupported reduction mode: {self.reduction_a}.')
        if precision not in ('auto', 'bf16', 'fp32'):
            raise ValueError(f'Unsupported precision: {precision}.')
        self.hidden_dim = hidden_dim
        self.precision = precision
        self._packed = None
        self._packed_key = None
        self._e_bound = math.inf
        self.embed_kernel = os.environ.get('GEOTX_EMBED', 'cheb')
        if self.embed_kernel not in ('tensor', 'table', 'cheb'):
            raise ValueError(f"GEOTX_EMBED must be 'cheb', 'table' or 'tensor', got {self.embed_kernel!r}")
        self.table_d_index_max = 96.0
        self._tables_cache = None
        self._tables_key = None
        self.cheb_d_index_max = 64.0
        self.cheb_error_budget = 2e-3  # largest |expansion - f| over largest |f|, per coefficient set
        self._cheb_cache = None
        self._cheb_key = None
        self._cheb_ok = False
        self._overflow_checked = set()

    def _use_tensor_cores(self):
        ok = self.hidden_dim in (128, 256)
        if self.precision == 'bf16' and not ok:
            raise RuntimeError(f"precision='bf16' needs hidden_dim 128 or 256, got {self.hidden_dim}")
        return ok and self.precision != 'fp32'

    def _packed_weights(self):
        ts = (self.proj_d.weight, self.proj_d.bias, self.proj_a.weight, self.proj_a.bias)
        key = tuple((t.data_ptr(), t._version) for t in ts)
        if self._packed is None or key != self._packed_key:
            w_d, b_d, w_a, b_a = [t.detach().contiguous() for t in ts]
            self._packed = ops.gse_pack_weights(w_d, b_d, w_a, b_a)
            # |sin|, |cos| <= 1, so |E_c| <= ||W_d[c,:]||_1 + ||W_a[c,:]||_1 + |b_d[c] + b_a[c]| whatever the points are
            self._e_bound = float((w_d.abs().sum(1) + w_a.abs().sum(1) + (b_d + b_a).abs()).max())
            self._packed_key = key
        return self._packed

    def _tables(self):
        ts = (self.proj_d.weight, self.proj_d.bias, self.proj_a.weight, self.proj_a.bias)
        key = tuple((t.data_ptr(), t._version) for t in ts) + (self.table_d_index_max,)
        if self._tables_cache is None or key != self._tables_key:
            w_d, b_d, w_a, b_a = [t.detach().contiguous() for t in ts]
            self._tables_cache = ops.gse_build_tables(self.embedding.div_term, w_d, b_d, w_a, b_a, self.factor_a,
                                                      self.table_d_index_max, h_a=0.125, h_d=0.5)
            self._tables_key = key
        return self._tables_cache

    def _weight_key(self):
        ts = (self.proj_d.weight, self.proj_d.bias, self.proj_a.weight, self.proj_a.bias, self.embedding.div_term)
        return tuple((t.data_ptr(), t._version) for t in ts)

    def _cheb(self):
        """Chebyshev coefficients of the current weights (rebuilt when a weight changes), or None when K2c's measured
        error on this weight set exceeds `cheb_error_budget` (the caller then takes the tcgen05 kernel)."""
        key = self._weight_key()
        if self._cheb_cache is None or key != self._cheb_key:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError('the embedding weights changed since the coefficients of the low-rank embedding '
                                   'kernel were built; run one forward outside CUDA graph capture first')
            ts = (self.proj_d.weight, self.proj_d.bias, self.proj_a.weight, self.proj_a.bias)
            w_d, b_d, w_a, b_a = [t.detach().contiguous() for t in ts]
            coef = ops.gse_build_cheb(self.embedding.div_term, w_d, b_d, w_a, b_a, self.factor_a)
            used = range(7)  # the angle set and all six distance sets (cheb_d_index_max may be widened later)
            rep = coef.report()
            bad = [rep[i] for i in used
                   if not rep[i]['max_error'] <= self.cheb_error_budget * max(rep[i]['max_abs_f'], 1e-30)]
            self._cheb_ok = not bad
            if bad:
                warnings.warn('ref-geotransformer_b200: the low-rank embedding kernel misses its error budget on these '
                              f'weights ({bad[0]}); using the tcgen05 embedding kernel instead')
            self._cheb_cache, self._cheb_key = coef, key
        return self._cheb_cache if self._cheb_ok else None

    def cheb_report(self):
        """Measured accuracy of K2c's expansions on the current weights (see ops.ChebCoefficients.report)."""
        self._cheb()
        return self._cheb_cache.report()

    def _use_cheb(self, out_dtype, qt):
        return (self.embed_kernel == 'cheb' and self.hidden_dim in (128, 256) and qt is None and
                out_dtype in (torch.float16, torch.bfloat16) and self.precision != 'fp32' and
                ops.gse_cheb_ka_tiles(math.pi * self.factor_a * (1.0 + 1e-6)) > 0 and self._cheb() is not None)

    def check_overflow(self):
        """Raises if an embedding argument ever fell outside the range its kernel was prepared for (synchronises)."""
        for cache, what in ((self._cheb_cache, 'cheb_d_index_max'), (self._tables_cache, 'the angle table')):
            if cache is not None and int(cache.overflow.item()) != 0:
                raise RuntimeError(f'GeometricStructureEmbedding: an argument fell outside {what}; raise '
                                   f"cheb_d_index_max (<= 210) or set embed_kernel = 'tensor'")

    def _use_table(self, out_dtype, qt):
        return (self.embed_kernel == 'table' and self.hidden_dim in (128, 256) and self.angle_k == 3 and
                self.reduction_a == 'max' and out_dtype == torch.float16 and qt is None and self.precision != 'fp32')

    def storage_dtype(self):
        """Format E is stored in between the embedding kernel and the 'self' blocks: IEEE half (11-bit significand)
        when the weights cannot push |E| out of its range (the bound above, checked when the weights are packed),
        else bfloat16; fp32 on the CUDA-core path."""
        if not self._use_tensor_cores():
            return torch.float32
        self._packed_weights()
        return torch.float16 if self._e_bound < 3.0e4 else torch.bfloat16

    @torch.no_grad()
    def geometry(self, points):
        return ops.gse_knn(points.contiguous(), self.angle_k)

    @torch.no_grad()
    def get_embedding_indices(self, points):
        geom = self.geometry(points)
        return ops.gse_indices(geom, self.sigma_d, self.factor_a)

    def embed(self, geom, out_dtype, row_begin=0, row_count=None, out=None, qt=None):
        """E of the rows [row_begin, row_begin + row_count).  With `qt` (tensor-core path, <= 4 heads) returns
        (E, scores_p): the first 'self' block's additive scores come out of the same kernel."""
        mean = self.reduction_a == 'mean'
        if self._use_cheb(out_dtype, qt):
            return self._cheb_call(geom, lambda: ops.gse_embed_cheb(
                geom, self._cheb_cache, self.sigma_d, self.factor_a, mean, out_dtype,
                ops.gse_cheb_max_d_set(self.cheb_d_index_max), row_begin, row_count, out))
        if self._use_table(out_dtype, qt):
            return ops.gse_embed_table(geom, self._tables(), self.sigma_d, self.factor_a, row_begin, row_count, out)
        if self._use_tensor_cores():
            return ops.gse_embed(geom, self.embedding.div_term, self._packed_weights(), self.hidden_dim, self.sigma_d,
                                 self.factor_a, mean, out_dtype, row_begin, row_count, out, qt=qt)
        if qt is not None:
            raise RuntimeError('fused first-block scores need the tensor-core embedding kernel')
        return ops.gse_embed_fp32(geom, self.embedding.div_term, self.proj_d.weight.detach(),
                                  self.proj_d.bias.detach(), self.proj_a.weight.detach(), self.proj_a.bias.detach(),
                                  self.sigma_d, self.factor_a, mean, out_dtype, row_begin, row_count, out)

    def distance_split_rows(self):
        """(80 | 144, C) Chebyshev coefficients of the distance term (sets 1, 2 and, when `cheb_d_index_max` reaches beyond
        64, 3) for folding into the query projection."""
        if self._cheb() is None:
            raise RuntimeError('the low-rank embedding kernel is not in use for these weights')
        return self._cheb_cache.distance_split_rows(ops.gse_cheb_max_d_set(self.cheb_d_index_max))

    def scores(self, geom, qt, row_begin=0, row_count=None, out=None, u=None):
        """K3r: scores_p[b,h,i,j] = qt[b,h,i,:] . E[b,i,j,:] (rpe_transformer.py:54-56 with proj_p folded into qt) without E
        ever being formed in memory; E enters the contraction rounded to half, exactly like a stored half E."""
        if not self._use_cheb(torch.float16, None):
            raise RuntimeError('GeometricStructureEmbedding.scores needs the low-rank embedding kernel (embed_kernel = '
                               "'cheb', hidden_dim 128 / 256)")
        return self._cheb_call(geom, lambda: ops.rpe_scores_cheb(
            geom, self._cheb_cache, qt, self.sigma_d, self.factor_a, self.reduction_a == 'mean',
            ops.gse_cheb_max_d_set(self.cheb_d_index_max), row_begin, row_count, out, u=u))

    def _cheb_call(self, geom, launch):
        """Runs a K2c / K3r launch; on the first eager call of a shape the overflow flag is read (one synchronisation) and
        a cloud larger than `cheb_d_index_max` widens the range to the kernels' maximum and repeats the launch."""
        res = launch()
        key = (geom.B, geom.n, self.cheb_d_index_max)
        if key not in self._overflow_checked and not torch.cuda.is_current_stream_capturing():
            self._overflow_checked.add(key)
            if int(self._cheb_cache.overflow.item()) != 0:
                if self.cheb_d_index_max >= 210.0:
                    self.check_overflow()  # raises
                warnings.warn(f'ref-geotransformer_b200: a cloud exceeds cheb_d_index_max = {self.cheb_d_index_max:g} index '
                              'units; widening it to 210 (one CTA per SM for the embedding kernels from now on)')
                self.cheb_d_index_max = 210.0
                self._cheb_cache.overflow.zero_()
                return self._cheb_call(geom, launch)  # the launch reads cheb_d_index_max when it runs
        return res

    def forward(self, points, out_dtype=torch.float32):
        """(B,n,3) -> (B,n,n,C).  Standalone calls return fp32 like the reference; GeometricTransformer asks for
        the 16-bit storage format (storage_dtype()) internally."""
        _forward_only_guard(self, points)
        return self.embed(self.geometry(points), out_dtype)


class AttentionOutput(nn.Module):
    """output_layer.py:6-21."""

    def __init__(self, d_model, dropout=None, activation_fn='ReLU'):
        super().__init__()
        self.expand = nn.Linear(d_model, d_model * 2)
        self.activation = build_act_layer(activation_fn)
        self.squeeze = nn.Linear(d_model * 2, d_model)
        self.dropout = build_dropout_layer(dropout)
        self.norm = nn.LayerNorm(d_model)

    def forward(self, input_states):
        hidden_states = self.squeeze(self.activation(self.expand(input_states)))
        hidden_states = self.dropout(hidden_states)
        return self.norm(input_states + hidden_states)


def _dense_scores(q, k, num_heads, scores_p, key_weights, key_masks, attention_factors, attention_masks):
    """Materialised attention matrix for the optional modifiers the fused kernel does not cover
    (rpe_transformer.py:59-66 / vanilla_transformer.py:55-64); unreachable from GeometricTransformer."""
    B, n, D = q.shape
    c = D // num_heads
    qh = q.view(B, n, num_heads, c).permute(0, 2, 1, 3)
    kh = k.view(B, -1, num_heads, c).permute(0, 2, 1, 3)
    scores = torch.matmul(qh, kh.transpose(-1, -2))
    if scores_p is not None:
        scores = scores + scores_p
    scores = scores / c ** 0.5
    if attention_factors is not None:
        scores = attention_factors.unsqueeze(1) * scores
    if key_weights is not None:
        scores = scores * key_weights.unsqueeze(1).unsqueeze(1)
    if key_masks is not None:
        scores = scores.masked_fill(key_masks.unsqueeze(1).unsqueeze(1), float('-inf'))
    if attention_masks is not None:
        scores = scores.masked_fill(attention_masks, float('-inf'))
    return F.softmax(scores, dim=-1)


class _FusedAttentionBase(nn.Module):
    need_scores = False  # set True to also return the (B,H,N,M) attention matrix like the reference does

    def _attend(self, q, k, v, scores_p, key_weights, key_masks, attention_factors, attention_masks=None):
        dense = self.need_scores or key_weights is not None or attention_factors is not None or \
            attention_masks is not None or (self.training and isinstance(self.dropout, nn.Dropout))
        if not dense:
            return ops.attention(q, k, v, self.num_heads, scores_p, key_masks), None
        scores = _dense_scores(q, k, self.num_heads, scores_p, key_weights, key_masks, attention_factors,
                               attention_masks)
        B, m, D = v.shape
        vh = v.view(B, m, self.num_heads, D // self.num_heads).permute(0, 2, 1, 3)
        hidden = torch.matmul(self.dropout(scores), vh)
        return hidden.permute(0, 2, 1, 3).reshape(B, -1, D), scores


class RPEMultiHeadAttention(_FusedAttentionBase):
    """rpe_transformer.py:18-72.  proj_p is folded into the query (exact):
    sum_c q_c (W_p e + b_p)_c = (W_p^T q) . e + q . b_p, and the second term is constant over keys so it cancels
    in the softmax.  proj_p.bias therefore has no effect on any output (as in exact arithmetic for the reference) unless
    `key_weights` / `attention_factors` rescale the scores before the softmax; forward() restores the term then."""

    def __init__(self, d_model, num_heads, dropout=None):
        super().__init__()
        if d_model % num_heads != 0:
            raise ValueError('`d_model` ({}) must be a multiple of `num_heads` ({}).'.format(d_model, num_heads))
        self.d_model = d_model
        self.num_heads = num_heads
        self.d_model_per_head = d_model // num_heads
        self.proj_q = nn.Linear(self.d_model, self.d_model)
        self.proj_k = nn.Linear(self.d_model, self.d_model)
        self.proj_v = nn.Linear(self.d_model, self.d_model)
        self.proj_p = nn.Linear(self.d_model, self.d_model)
        self.dropout = build_dropout_layer(dropout)

    def fused_qkvp(self, gd=None):
        """One projection for the whole layer: rows [W_q; W_k; W_v; W_qp] with W_qp[h] = W_p[h]^T W_q[h], so that
        x -> (q | k | v | folded query of every head) is a single (C -> 3C + H*C) GEMM.  The fold is exact algebra
        (formed in fp64, stored fp32); rebuilt whenever one of the source parameters changes.
        gd (M, C): Chebyshev coefficients G_d of the embedding's distance term (ops.ChebCoefficients.distance_split_rows);
        then H*M more rows W_u[h] = G_d W_qp[h] follow, so that the same GEMM also yields u = G_d qt, the distance term's
        share of the scores in coefficient space (K3r's distance split)."""
        ps = (self.proj_q.weight, self.proj_q.bias, self.proj_k.weight, self.proj_k.bias, self.proj_v.weight,
              self.proj_v.bias, self.proj_p.weight) + ((gd,) if gd is not None else ())
        key = tuple((t.data_ptr(), t._version, tuple(t.shape)) for t in ps)
        if getattr(self, '_fused_key', None) != key:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError('a weight changed since the fused projection was built; run one forward outside CUDA '
                                   'graph capture first')
            H, c, C = self.num_heads, self.d_model_per_head, self.d_model
            wp = self.proj_p.weight.detach().double().view(H, c, C)
            wq = self.proj_q.weight.detach().double().view(H, c, C)
            bq = self.proj_q.bias.detach().double().view(H, c)
            w_qp = torch.einsum('hcd,hci->hdi', wp, wq).reshape(H * C, C)
            b_qp = torch.einsum('hcd,hc->hd', wp, bq).reshape(H * C)
            ws = [self.proj_q.weight.detach().double(), self.proj_k.weight.detach().double(),
                  self.proj_v.weight.detach().double(), w_qp]
            bs = [self.proj_q.bias.detach().double(), self.proj_k.bias.detach().double(),
                  self.proj_v.bias.detach().double(), b_qp]
            if gd is not None:
                g64 = gd.detach().double()                                           # (M, C)
                ws.append(torch.einsum('mc,hci->hmi', g64, w_qp.view(H, C, C)).reshape(-1, C))
                bs.append(torch.einsum('mc,hc->hm', g64, b_qp.view(H, C)).reshape(-1))
            self._fused, self._fused_key = (torch.cat(ws).float().contiguous(), torch.cat(bs).float().contiguous()), key
        return self._fused

    def folded_query(self, q):
        """q (B,N,H*c) -> qt view of logical shape (B,H,N,C): qt[b,h,i,:] = W_p[h]^T q[b,h,i,:]."""
        B, n, D = q.shape
        H, c = self.num_heads, self.d_model_per_head
        wp = self.proj_p.weight.view(H, c, D)                          # rows h*c..h*c+c of proj_p.weight
        qh = q.view(B * n, H, c).transpose(0, 1)                       # (H, B*n, c)
        qt = torch.bmm(qh, wp)                                         # (H, B*n, C)
        return qt.view(H, B, n, D).permute(1, 0, 2, 3)                 # logical (B,H,N,C), no copy

    def forward(self, input_q, input_k, input_v, embed_qk, key_weights=None, key_masks=None, attention_factors=None):
        """embed_qk: (B,N,M,C) fp16, bf16 or fp32 embedding E (from GeometricStructureEmbedding)."""
        _forward_only_guard(self, input_q, input_k, input_v, embed_qk)
        q = self.proj_q(input_q)
        k = self.proj_k(input_k)
        v = self.proj_v(input_v)
        scores_p = ops.rpe_scores(self.folded_query(q), embed_qk.contiguous())
        if key_weights is not None or attention_factors is not None:
            # rpe_transformer.py:58-62 multiplies the summed scores by attention_factors[b,i,j] / key_weights[b,j] BEFORE
            # the softmax: the folded-away term q . b_p is then no longer constant over the keys and must be restored
            B, n, _ = q.shape
            H, c = self.num_heads, self.d_model_per_head
            qb = torch.einsum('bnhc,hc->bhn', q.view(B, n, H, c), self.proj_p.bias.view(H, c))
            scores_p = scores_p + qb.unsqueeze(-1)
        return self._attend(q, k, v, scores_p, key_weights, key_masks, attention_factors)


class MultiHeadAttention(_FusedAttentionBase):
    """vanilla_transformer.py:15-70."""

    def __init__(self, d_model, num_heads, dropout=None):
        super().__init__()
        if d_model % num_heads != 0:
            raise ValueError('`d_model` ({}) must be a multiple of `num_heads` ({}).'.format(d_model, num_heads))
        self.d_model = d_model
        self.num_heads = num_heads
        self.d_model_per_head = d_model // num_heads
        self.proj_q = nn.Linear(self.d_model, self.d_model)
        self.proj_k = nn.Linear(self.d_model, self.d_model)
        self.proj_v = nn.Linear(self.d_model, self.d_model)
        self.dropout = build_dropout_layer(dropout)

    def fused_kv(self):
        """[W_k; W_v] as one (C -> 2C) projection of the memory tokens."""
        ps = (self.proj_k.weight, self.proj_k.bias, self.proj_v.weight, self.proj_v.bias)
        key = tuple((t.data_ptr(), t._version, tuple(t.shape)) for t in ps)
        if getattr(self, '_fused_key', None) != key:
            w = torch.cat([self.proj_k.weight.detach(), self.proj_v.weight.detach()]).contiguous()
            b = torch.cat([self.proj_k.bias.detach(), self.proj_v.bias.detach()]).contiguous()
            self._fused, self._fused_key = (w, b), key
        return self._fused

    def fused_qkv(self):
        """[W_q; W_k; W_v] as one (C -> 3C) projection (a 'cross' block projects every token of both clouds once)."""
        ps = (self.proj_q.weight, self.proj_q.bias) + tuple(self.fused_kv())
        key = tuple((t.data_ptr(), t._version) for t in ps)
        if getattr(self, '_fused3_key', None) != key:
            w = torch.cat([self.proj_q.weight.detach(), ps[2]]).contiguous()
            b = torch.cat([self.proj_q.bias.detach(), ps[3]]).contiguous()
            self._fused3, self._fused3_key = (w, b), key
        return self._fused3

    def forward(self, input_q, input_k, input_v, key_weights=None, key_masks=None, attention_factors=None,
                attention_masks=None):
        _forward_only_guard(self, input_q, input_k, input_v)
        q = self.proj_q(input_q)
        k = self.proj_k(input_k)
        v = self.proj_v(input_v)
        return self._attend(q, k, v, None, key_weights, key_masks, attention_factors, attention_masks)


class RPEAttentionLayer(nn.Module):
    """rpe_transformer.py:75-104."""

    def __init__(self, d_model, num_heads, dropout=None):
        super().__init__()
        self.attention = RPEMultiHeadAttention(d_model, num_heads, dropout=dropout)
        self.linear = nn.Linear(d_model, d_model)
        self.dropout = build_dropout_layer(dropout)
        self.norm = nn.LayerNorm(d_model)

    def forward(self, input_states, memory_states, position_states, memory_weights=None, memory_masks=None,
                attention_factors=None):
        hidden_states, attention_scores = self.attention(
            input_states, memory_states, memory_states, position_states,
            key_weights=memory_weights, key_masks=memory_masks, attention_factors=attention_factors)
        hidden_states = self.dropout(self.linear(hidden_states))
        return self.norm(hidden_states + input_states), attention_scores


class RPETransformerLayer(nn.Module):
    """rpe_transformer.py:107-131."""

    def __init__(self, d_model, num_heads, dropout=None, activation_fn='ReLU'):
        super().__init__()
        self.attention = RPEAttentionLayer(d_model, num_heads, dropout=dropout)
        self.output = AttentionOutput(d_model, dropout=dropout, activation_fn=activation_fn)

    def forward(self, input_states, memory_states, position_states, memory_weights=None, memory_masks=None,
                attention_factors=None):
        hidden_states, attention_scores = self.attention(
            input_states, memory_states, position_states,
            memory_weights=memory_weights, memory_masks=memory_masks, attention_factors=attention_factors)
        return self.output(hidden_states), attention_scores


class AttentionLayer(nn.Module):
    """vanilla_transformer.py:73-102."""

    def __init__(self, d_model, num_heads, dropout=None):
        super().__init__()
        self.attention = MultiHeadAttention(d_model, num_heads, dropout=dropout)
        self.linear = nn.Linear(d_model, d_model)
        self.dropout = build_dropout_layer(dropout)
        self.norm = nn.LayerNorm(d_model)

    def forward(self, input_states, memory_states, memory_weights=None, memory_masks=None, attention_factors=None,
                attention_masks=None):
        hidden_states, attention_scores = self.attention(
            input_states, memory_states, memory_states, key_weights=memory_weights, key_masks=memory_masks,
            attention_factors=attention_factors, attention_masks=attention_masks)
        hidden_states = self.dropout(self.linear(hidden_states))
        return self.norm(hidden_states + input_states), attention_scores


class TransformerLayer(nn.Module):
    """vanilla_transformer.py:105-129."""

    def __init__(self, d_model, num_heads, dropout=None, activation_fn='ReLU'):
        super().__init__()
        self.attention = AttentionLayer(d_model, num_heads, dropout=dropout)
        self.output = AttentionOutput(d_model, dropout=dropout, activation_fn=activation_fn)

    def forward(self, input_states, memory_states, memory_weights=None, memory_masks=None, attention_factors=None,
                attention_masks=None):
        hidden_states, attention_scores = self.attention(
            input_states, memory_states, memory_weights=memory_weights, memory_masks=memory_masks,
            attention_factors=attention_factors, attention_masks=attention_masks)
        return self.output(hidden_states), attention_scores


def _check_block_type(block):
    if block not in ['self', 'cross']:
        raise ValueError('Unsupported block type "{}".'.format(block))


_SIDE_STREAMS = {}


def _side_stream(device):
    """One auxiliary stream per device for the small launches that are overlapped with a bandwidth-bound kernel."""
    device = torch.device(device)
    if device not in _SIDE_STREAMS:
        _SIDE_STREAMS[device] = torch.cuda.Stream(device)
    return _SIDE_STREAMS[device]


class StreamedEmbedding:
    """E of one cloud, produced slab by slab together with the FIRST 'self' block's scores_p.

    The first 'self' block's folded query depends only on in_proj(feats), not on E, so E need not exist before the
    transformer starts: the embedding kernel writes a slab of query rows (sized to stay inside the L2 cache) and the
    scores kernel consumes that slab straight away, while its lines are still L2-resident.  That removes one of the
    three full HBM reads of E (one per 'self' block); the slab still drains to HBM for the later blocks.
    `slab_bytes` = 0 produces E in one launch (the scores kernel then reads it from HBM as in the later blocks)."""

    def __init__(self, embedding, geom, dtype, slab_bytes, fuse_scores=False, regenerate_bytes=0, direct=False):
        self.embedding, self.geom, self.dtype, self.slab_bytes = embedding, geom, dtype, slab_bytes
        self.fuse_scores = fuse_scores
        self.E = None
        # direct: E is never stored AND never formed in memory: every 'self' block gets its scores_p from K3r
        # (geotx_rpe_scores_cheb), which rebuilds E tile by tile in registers from the geometry
        self.direct = direct
        # > 0: E is NEVER stored.  Every 'self' block regenerates it in query-row slabs of this many bytes, each slab
        # consumed at once by the scores kernel (for clouds whose n^2 C bytes do not fit the GPU next to the other
        # cloud's: 137 GB per cloud at n = 16384).  Costs (number of 'self' blocks - 1) extra embedding passes.
        self.regenerate_bytes = regenerate_bytes
        self._slab = None

    def direct_scores(self, qt, u=None):
        """qt (1,H,n,C) [, u (1,H,n,80)] -> scores_p (1,H,n,n) straight from the geometry (K3r); E never exists."""
        return self.embedding.scores(self.geom, qt, u=u)

    def regenerated_scores(self, qt):
        """qt (1,H,n,C) -> scores_p (1,H,n,n); E exists one slab at a time."""
        n, C = self.geom.n, self.embedding.hidden_dim
        row_bytes = n * C * (4 if self.dtype == torch.float32 else 2)
        rows = max(1, min(n, self.regenerate_bytes // row_bytes))
        if self._slab is None:
            self._slab = torch.empty(1, rows, n, C, dtype=self.dtype, device=qt.device)
        scores_p = torch.empty(1, qt.shape[1], n, n, dtype=torch.float32, device=qt.device)
        for r0 in range(0, n, rows):
            rc = min(rows, n - r0)
            self.embedding.embed(self.geom, self.dtype, r0, rc, out=self._slab[:, :rc])
            ops.rpe_scores(qt, self._slab[:, :rc], n=n, row_begin=r0, out=scores_p)
        return scores_p

    def materialise(self):
        if self.direct:
            self.direct = False  # a consumer wants E itself (layer-by-layer path): store it after all
        if self.regenerate_bytes > 0:
            raise RuntimeError('this embedding is in regenerate mode (it does not fit in memory); only the fused '
                               'single-pair schedule can consume it')
        if self.E is None:
            self.E = self.embedding.embed(self.geom, self.dtype)
        return self.E

    def first_scores(self, qt):
        """qt (1,H,n,C) -> scores_p (1,H,n,n) of the first 'self' block; E is complete afterwards."""
        if self.E is not None:
            return ops.rpe_scores(qt, self.E)
        if self.fuse_scores and self.slab_bytes <= 0 and qt.shape[1] <= 4 and self.embedding._use_tensor_cores():
            # the embedding kernel emits these scores from its accumulators: one HBM read of E less
            self.E, scores_p = self.embedding.embed(self.geom, self.dtype, qt=qt)
            return scores_p
        n, C = self.geom.n, self.embedding.hidden_dim
        row_bytes = n * C * (4 if self.dtype == torch.float32 else 2)
        rows = n if self.slab_bytes <= 0 else max(1, min(n, self.slab_bytes // row_bytes))
        if rows >= n or self.geom.B != 1:
            return ops.rpe_scores(qt, self.materialise())
        E = torch.empty(1, n, n, C, dtype=self.dtype, device=qt.device)
        scores_p = torch.empty(1, qt.shape[1], n, n, dtype=torch.float32, device=qt.device)
        for r0 in range(0, n, rows):
            rc = min(rows, n - r0)
            self.embedding.embed(self.geom, self.dtype, r0, rc, out=E[:, r0:r0 + rc])
            ops.rpe_scores(qt, E[:, r0:r0 + rc], n=n, row_begin=r0, out=scores_p)
        self.E = E
        return scores_p


class RPEConditionalTransformer(nn.Module):
    """conditional_transformer.py:73-117."""

    def __init__(self, blocks, d_model, num_heads, dropout=None, activation_fn='ReLU', return_attention_scores=False,
                 parallel=False):
        super().__init__()
        self.blocks = blocks
        layers = []
        for block in self.blocks:
            _check_block_type(block)
            if block == 'self':
                layers.append(RPETransformerLayer(d_model, num_heads, dropout=dropout, activation_fn=activation_fn))
            else:
                layers.append(TransformerLayer(d_model, num_heads, dropout=dropout, activation_fn=activation_fn))
        self.layers = nn.ModuleList(layers)
        self.return_attention_scores = return_attention_scores
        self.parallel = parallel
        self.overlap_attention = os.environ.get('GEOTX_OVERLAP_ATTENTION', '1') != '0'
        # False (default): one forward at a time -- a small attention grid is spread over the GPU by splitting the keys
        # over thread-block clusters (shortest latency).  True: several forwards are in flight (GraphedGeometricTransformer
        # (throughput=True), PipelinedGeometricTransformer) -- SM-time counts, not latency: one lean CTA per 64 queries
        # walks all keys (no split / merge overhead).  Same results up to fp32 summation order.
        self.throughput = False

    # ---- fused schedule (same math, fewer and larger launches) ------------------------------------------------
    # Both clouds share every layer's weights, so all token-wise work of a 'self' block (one fused q|k|v|folded-query
    # projection, the output Linear, the two residual LayerNorms, the C->2C->C FFN) runs once over the concatenated
    # tokens of both clouds; only the two attention cores see the clouds separately.  'cross' blocks keep the
    # reference's sequential dependency (conditional_transformer.py:110-111).  Used when nothing asks for the
    # materialised attention matrix; otherwise the layer-by-layer path below runs.
    def _fused_ok(self, feats0, feats1):
        if self.return_attention_scores or feats0.shape[0] != 1 or feats0.dim() != 3:
            return False
        C = feats0.shape[-1]
        if C % 4 or C > 1024:
            return False
        for layer in self.layers:
            att = layer.attention.attention
            if att.d_model_per_head not in (16, 32, 64):
                return False
            if self.training and any(isinstance(m, nn.Dropout) for m in layer.modules()):
                return False
        return True

    @staticmethod
    def _post_attention(layer, hidden, x_in, out=None):
        """rpe_transformer.py:101-103 / vanilla_transformer.py:99-101 + output_layer.py:15-21 on (1, T, C) tokens."""
        al, ao = layer.attention, layer.output
        C = hidden.shape[-1]
        if _dense_own() and C in (128, 256) and isinstance(ao.activation, nn.ReLU) and hidden.dtype == torch.float32:
            return ops.post_attention(hidden, x_in, _half_of(al.linear, 'w', al.linear.weight), al.linear.bias,
                                      al.norm.weight, al.norm.bias, al.norm.eps,
                                      _half_of(ao.expand, 'w', ao.expand.weight), ao.expand.bias,
                                      _half_of(ao.squeeze, 'w', ao.squeeze.weight), ao.squeeze.bias,
                                      ao.norm.weight, ao.norm.bias, ao.norm.eps, out=out)
        h = F.linear(hidden, al.linear.weight, al.linear.bias)
        x1 = ops.add_layernorm(h, x_in, al.norm.weight, al.norm.bias, al.norm.eps)
        T, C = x1.shape[1], x1.shape[2]
        if isinstance(ao.activation, nn.ReLU):
            e = torch._addmm_activation(ao.expand.bias, x1.view(T, C), ao.expand.weight.t())  # bias + ReLU epilogue
        else:
            e = ao.activation(F.linear(x1.view(T, C), ao.expand.weight, ao.expand.bias))
        sq = F.linear(e, ao.squeeze.weight, ao.squeeze.bias).view(1, T, C)
        return ops.add_layernorm(sq, x1, ao.norm.weight, ao.norm.bias, ao.norm.eps, out=out)

    def _forward_fused(self, feats0, feats1, embeddings0, embeddings1, masks0, masks1):
        N, C = feats0.shape[1], feats0.shape[2]
        embs = [embeddings0, embeddings1]
        x = _joint_view(feats0, feats1)  # GeometricTransformer.forward projects both clouds into one buffer: no copy
        if x is None:
            x = torch.cat([feats0, feats1], dim=1)  # (1, N+M, C); cloud slices of it are contiguous views
        T = x.shape[1]
        split = not self.throughput
        for i, block in enumerate(self.blocks):
            layer = self.layers[i]
            att = layer.attention.attention
            H = att.num_heads
            if block == 'self':
                # K3r's distance split: the projection also yields u = G_d qt (80 coefficients per head and token)
                gd = None
                if H % 4 == 0 and all(isinstance(e, StreamedEmbedding) and e.direct for e in embs):
                    gd = embs[0].embedding.distance_split_rows()
                w, b = att.fused_qkvp(gd)
                # one projection: q | k | v | folded query per head [| u].  (1, T, 3C + H*C [+ H*80]) fp32, or the q | k | v
                # block as half (what the attention kernel contracts in) and the rest as fp32
                if _attention_f16():
                    qkv, y = _linear(x, att, 'qkvp', w, b, half_cols=3 * C)
                    off = 0
                else:
                    qkv = y = _linear(x, att, 'qkvp', w, b)
                    off = 3 * C
                M = 0 if gd is None else gd.shape[0]
                hidden = x.new_empty(1, T, C)
                cur = side = None
                streamed = any(isinstance(e, StreamedEmbedding) for e in embs)
                if streamed and not any(isinstance(e, StreamedEmbedding) and
                                        (e.direct or e.fuse_scores or e.slab_bytes > 0 or e.regenerate_bytes > 0)
                                        for e in embs):
                    # plain schedule: both embedding kernels back to back, each with the whole GPU to itself
                    embs = [e.materialise() if isinstance(e, StreamedEmbedding) else e for e in embs]
                    streamed = False
                # E-free mode: the two clouds' chains (K3r -> attention) are independent; the second cloud's runs on a side
                # stream, so that its persistent K3r fills the SMs the first cloud's K3r drains from and the two short
                # attention launches share the GPU (GEOTX_PARALLEL_CLOUDS=0: one stream)
                fork = (self.overlap_attention and os.environ.get('GEOTX_PARALLEL_CLOUDS', '1') != '0' and
                        all(isinstance(e, StreamedEmbedding) and e.direct for e in embs))
                for which, (lo, hi, mask) in enumerate(((0, N, masks0), (N, T, masks1))):
                    emb = embs[which]
                    qt = y[:, lo:hi, off:off + H * C].unflatten(2, (H, C)).permute(0, 2, 1, 3)  # (1, H, n, C) view
                    if isinstance(emb, StreamedEmbedding) and emb.direct:
                        u = y[:, lo:hi, off + H * C:].unflatten(2, (H, M)).permute(0, 2, 1, 3) if M else None
                        if fork and which == 1:
                            cur = torch.cuda.current_stream(x.device)
                            side = _side_stream(x.device)
                            side.wait_stream(cur)
                            with torch.cuda.stream(side):
                                scores_p = emb.direct_scores(qt, u)
                                ops.attention(qkv[:, lo:hi, 0:C], qkv[:, lo:hi, C:2 * C], qkv[:, lo:hi, 2 * C:3 * C], H, scores_p,
                                              mask, split_keys=split, out=hidden[:, lo:hi])
                            del scores_p  # allocated and used on the side stream only
                            continue
                        scores_p = emb.direct_scores(qt, u)  # K3r: E of this cloud is never formed in memory
                    elif isinstance(emb, StreamedEmbedding) and emb.regenerate_bytes > 0:
                        scores_p = emb.regenerated_scores(qt)  # E of this cloud is never stored
                    elif isinstance(emb, StreamedEmbedding):
                        scores_p = emb.first_scores(qt)  # the first block interleaves with the embedding kernel
                        embs[which] = emb.E
                    else:
                        scores_p = ops.rpe_scores(qt, emb.contiguous())
                    att_args = (qkv[:, lo:hi, 0:C], qkv[:, lo:hi, C:2 * C], qkv[:, lo:hi, 2 * C:3 * C], H, scores_p, mask)
                    if which == 0 and self.overlap_attention and not streamed:
                        # the first cloud's attention core (a short, latency-bound launch) runs on a side stream
                        # underneath the second cloud's HBM-bound scores kernel
                        cur = torch.cuda.current_stream(x.device)
                        side = _side_stream(x.device)
                        side.wait_stream(cur)
                        with torch.cuda.stream(side):
                            ops.attention(*att_args, split_keys=split, out=hidden[:, lo:hi])
                        scores_p.record_stream(side)
                    else:
                        ops.attention(*att_args, split_keys=split, out=hidden[:, lo:hi])
                if side is not None:
                    cur.wait_stream(side)
                x = self._post_attention(layer, hidden, x)
            else:
                # One q|k|v projection of ALL tokens serves both steps' queries (each step queries with the OLD tokens
                # of its cloud) and the first step's memory (the old second cloud); only the second step's memory --
                # the UPDATED first cloud, conditional_transformer.py:110-111 -- needs its own k|v projection
                # (parallel=True reads the old one: no second projection at all).
                x_new = torch.empty_like(x)
                w3, b3 = att.fused_qkv()
                f16 = _attention_f16()
                y = _linear(x, att, 'qkv', w3, b3, half_cols=3 * C)[0] if f16 else _linear(x, att, 'qkv', w3, b3)  # (1, T, 3C)
                order = ((0, N, N, T, masks1), (N, T, 0, N, masks0))  # (query cloud, memory cloud, memory mask)
                for step, (qlo, qhi, mlo, mhi, mask) in enumerate(order):
                    if step == 1 and not self.parallel:
                        wkv, bkv = att.fused_kv()
                        kv = _linear(x_new[:, mlo:mhi], att, 'kv', wkv, bkv, half_cols=2 * C)[0] if f16 else \
                            _linear(x_new[:, mlo:mhi], att, 'kv', wkv, bkv)
                        k_mem, v_mem = kv[:, :, :C], kv[:, :, C:]
                    else:
                        k_mem, v_mem = y[:, mlo:mhi, C:2 * C], y[:, mlo:mhi, 2 * C:]
                    hidden = ops.attention(y[:, qlo:qhi, :C], k_mem, v_mem, H, None, mask, split_keys=split)
                    self._post_attention(layer, hidden, x[:, qlo:qhi], out=x_new[:, qlo:qhi])
                x = x_new
        return x[:, :N], x[:, N:]

    def forward(self, feats0, feats1, embeddings0, embeddings1, masks0=None, masks1=None):
        streamed = [e for e in (embeddings0, embeddings1) if isinstance(e, StreamedEmbedding)]
        if self._fused_ok(feats0, feats1):
            _forward_only_guard(self, feats0, feats1, *[e for e in (embeddings0, embeddings1)
                                                         if not isinstance(e, StreamedEmbedding)])
            if 'self' not in self.blocks:
                for e in streamed:
                    e.materialise()
            return self._forward_fused(feats0.contiguous(), feats1.contiguous(), embeddings0, embeddings1, masks0,
                                       masks1)
        if streamed:
            embeddings0 = embeddings0.materialise() if isinstance(embeddings0, StreamedEmbedding) else embeddings0
            embeddings1 = embeddings1.materialise() if isinstance(embeddings1, StreamedEmbedding) else embeddings1
        for layer in self.layers:
            layer.attention.attention.need_scores = bool(self.return_attention_scores)
        attention_scores = []
        for i, block in enumerate(self.blocks):
            if block == 'self':
                feats0, scores0 = self.layers[i](feats0, feats0, embeddings0, memory_masks=masks0)
                feats1, scores1 = self.layers[i](feats1, feats1, embeddings1, memory_masks=masks1)
            elif self.parallel:
                new_feats0, scores0 = self.layers[i](feats0, feats1, memory_masks=masks1)
                new_feats1, scores1 = self.layers[i](feats1, feats0, memory_masks=masks0)
                feats0, feats1 = new_feats0, new_feats1
            else:  # sequential: the second call sees the UPDATED feats0 (conditional_transformer.py:110-111)
                feats0, scores0 = self.layers[i](feats0, feats1, memory_masks=masks1)
                feats1, scores1 = self.layers[i](feats1, feats0, memory_masks=masks0)
            if self.return_attention_scores:
                attention_scores.append([scores0, scores1])
        if self.return_attention_scores:
            return feats0, feats1, attention_scores
        return feats0, feats1


class GeometricTransformer(nn.Module):
    """geotransformer.py:75-155.  Extra keyword `precision` ('auto' | 'bf16' | 'fp32') selects the embedding
    kernel ('bf16' = tcgen05 with bf16 operands, fp32 accumulation; E stored as fp16, or bf16 if the weights could
    overflow it) or the fp32 CUDA-core kernel (E stored as fp32); everything else is the reference signature."""

    def __init__(self, input_dim, output_dim, hidden_dim, num_heads, blocks, sigma_d, sigma_a, angle_k, dropout=None,
                 activation_fn='ReLU', reduction_a='max', precision='auto'):
        super().__init__()
        self.embedding = GeometricStructureEmbedding(hidden_dim, sigma_d, sigma_a, angle_k, reduction_a=reduction_a,
                                                     precision=precision)
        self.in_proj = nn.Linear(input_dim, hidden_dim)
        self.transformer = RPEConditionalTransformer(blocks, hidden_dim, num_heads, dropout=dropout,
                                                     activation_fn=activation_fn)
        self.out_proj = nn.Linear(hidden_dim, output_dim)
        # slab of E kept L2-resident between the embedding kernel and the first block's scores kernel (0 = off)
        self.l2_slab_bytes = int(os.environ.get('GEOTX_L2_SLAB_BYTES', DEFAULT_L2_SLAB_BYTES))
        # first 'self' block's scores from the embedding kernel's accumulators: one HBM pass over E less, but the
        # extra epilogue work costs the embedding kernel more than the pass saves (measured: 0.50 -> 0.64 ms per
        # cloud at n = 1024 against 0.094 ms saved), so it is opt-in
        self.fuse_scores = os.environ.get('GEOTX_FUSE_SCORES', '0') != '0'
        # E storage: 'scores' never forms E in memory at all: each 'self' block's additive scores come from K3r
        # (geotx_rpe_scores_cheb), which rebuilds E tile by tile in registers (needs the low-rank embedding kernel,
        # half-precision E, <= 8 heads, B = 1); 'store' keeps both clouds' E for all 'self' blocks (one embedding
        # pass per cloud, one HBM read per block); 'regenerate' never stores E (one pass per cloud per 'self' block,
        # in slabs of `regenerate_slab_bytes`); 'auto' = 'scores' when possible, else 'store' when both E fit into
        # 80 % of the free device memory, else 'regenerate'.
        self.e_mode = os.environ.get('GEOTX_E_MODE', 'auto')
        self.regenerate_slab_bytes = int(os.environ.get('GEOTX_REGENERATE_SLAB_BYTES', 1 << 30))
        self.last_e_mode = None
        self._auto_mode = {}  # (n, m, dtype) -> decision of 'auto', taken on the first (eager) call of that shape

    def _regenerate_bytes(self, n, m, e_dtype, device, batch):
        if self.e_mode not in ('auto', 'store', 'regenerate', 'scores'):
            raise ValueError(f"GEOTX_E_MODE must be 'auto', 'scores', 'store' or 'regenerate', got {self.e_mode!r}")
        if self.e_mode in ('store', 'scores') or batch != 1 or 'self' not in self.transformer.blocks:
            return 0
        if self.e_mode == 'auto':
            key = (n, m, e_dtype)
            if key not in self._auto_mode:  # never queried while a CUDA graph is being captured (shape seen before)
                item = 4 if e_dtype == torch.float32 else 2
                need = (n * n + m * m) * self.embedding.hidden_dim * item
                free, _ = torch.cuda.mem_get_info(device)
                reusable = torch.cuda.memory_reserved(device) - torch.cuda.memory_allocated(device)
                self._auto_mode[key] = need <= 0.8 * (free + reusable)
            if self._auto_mode[key]:
                return 0
        return self.regenerate_slab_bytes

    def _in_proj_into(self, feats, out, w_t):
        """geotransformer.py:140-141 for one cloud, written into its slice of the joint token buffer."""
        w = self.in_proj.weight
        if _dense_own() and feats.dtype == torch.float32 and ops.linear_supported(w.shape[1], w.shape[0]) and \
                feats.stride(-1) == 1:
            ops.linear_f16(feats, _half_of(self.in_proj, 'w', w), self.in_proj.bias, out=out)
        else:
            torch.addmm(self.in_proj.bias, feats, w_t, out=out)

    def _direct_scores(self, e_dtype, batch, feats):
        """Whether the 'self' blocks take their scores straight from the geometry (K3r) instead of from a stored E."""
        if self.e_mode not in ('auto', 'scores') or 'self' not in self.transformer.blocks:
            return False
        t = self.transformer
        ok = (batch == 1 and e_dtype == torch.float16 and self.embedding._use_cheb(e_dtype, None) and
              t._fused_ok(feats, feats) and all(l.attention.attention.num_heads <= 8 for l in t.layers) and
              not self.fuse_scores and self.l2_slab_bytes <= 0)
        if self.e_mode == 'scores' and not ok:
            raise RuntimeError("GEOTX_E_MODE='scores' needs the low-rank embedding kernel (embed_kernel 'cheb', hidden_dim "
                               '128 / 256), half-precision E, B = 1 and <= 8 heads')
        return ok

    def forward(self, ref_points, src_points, ref_feats, src_feats, ref_masks=None, src_masks=None):
        _forward_only_guard(self, ref_points, src_points, ref_feats, src_feats)
        if not all(isinstance(t, torch.Tensor) and t.is_cuda for t in (ref_points, src_points, ref_feats, src_feats)):
            raise RuntimeError('GeometricTransformer (ref-geotransformer_b200) needs CUDA tensor inputs: there is no '
                               'CPU path')
        tc = self.embedding._use_tensor_cores()
        e_dtype = self.embedding.storage_dtype()
        # The second cloud's k-NN + packing and its in_proj run on a side stream next to the first cloud's: all four
        # are small launches (~25 us of latency-bound work each) that leave most of the GPU idle on their own.
        cur = torch.cuda.current_stream(ref_points.device)
        side = _side_stream(ref_points.device)
        one_buffer = ref_feats.dim() == 3 and ref_feats.shape[0] == 1 and src_feats.shape[0] == 1
        with _dense_precision(allow_tf32=tc):
            if one_buffer:  # both clouds' tokens side by side: the transformer's fused schedule wants exactly this
                N, M = ref_feats.shape[1], src_feats.shape[1]
                x = ref_feats.new_empty(1, N + M, self.in_proj.out_features)
                w_t = self.in_proj.weight.t()
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                src_geom = self.embedding.geometry(src_points)
                if one_buffer:
                    self._in_proj_into(src_feats[0], x[0, N:], w_t)
                else:
                    src_proj = self.in_proj(src_feats)
            ref_geom = self.embedding.geometry(ref_points)
            if one_buffer:
                self._in_proj_into(ref_feats[0], x[0, :N], w_t)
                ref_feats, src_feats = x[:, :N], x[:, N:]
            else:
                ref_feats, src_feats = self.in_proj(ref_feats), src_proj
            cur.wait_stream(side)
            for t in (src_geom.buf, src_geom.knn) + (() if one_buffer else (src_feats,)):
                t.record_stream(cur)  # allocated on the side stream, consumed (and later freed) on this one
            direct = self._direct_scores(e_dtype, ref_points.shape[0], ref_feats)
            regen = 0 if direct else self._regenerate_bytes(ref_points.shape[1], src_points.shape[1], e_dtype,
                                                            ref_points.device, ref_points.shape[0])
            self.last_e_mode = 'scores' if direct else 'regenerate' if regen else 'store'
            ref_embeddings = StreamedEmbedding(self.embedding, ref_geom, e_dtype, self.l2_slab_bytes, self.fuse_scores,
                                               regen, direct)
            src_embeddings = StreamedEmbedding(self.embedding, src_geom, e_dtype, self.l2_slab_bytes, self.fuse_scores,
                                               regen, direct)
            ref_feats, src_feats = self.transformer(ref_feats, src_feats, ref_embeddings, src_embeddings,
                                                    masks0=ref_masks, masks1=src_masks)
            both = _joint_view(ref_feats, src_feats)
            if both is not None:
                # the fused schedule returns two slices of one (1, N+M, C) buffer: project them in one GEMM
                n = ref_feats.shape[1]
                out = _linear(both, self.out_proj, 'w', self.out_proj.weight, self.out_proj.bias)
                return out[:, :n], out[:, n:]
            return self.out_proj(ref_feats), self.out_proj(src_feats)


class GraphedGeometricTransformer:
    """CUDA-graph replay of GeometricTransformer.forward for fixed cloud sizes (inference serving).

    One forward is ~190 kernel launches of a few microseconds each at n ~ 1k; issued eagerly the host cannot keep the
    GPU fed.  The whole forward (our kernels and the cuBLASLt GEMMs between them) is captured once into a CUDA graph on
    static input/output buffers and replayed with one launch.  The graph reads the plain weights from the module's own
    parameter storage, but also DERIVED buffers that exist only as of the capture (the fused q|k|v|folded-query
    projections of every block, the embedding kernel's coefficients / tables / packed weights).  Every call therefore
    compares the version counters of all parameters and buffers with those seen at capture and recaptures when any of
    them moved (load_state_dict, in-place updates): a replay never mixes new and old weights and never reads a derived
    buffer that has been freed.  Replacing a parameter OBJECT (module.weight = nn.Parameter(...)) is not an in-place
    update: call `recapture()` yourself after that.

        fast = GraphedGeometricTransformer(model, ref_points, src_points, ref_feats, src_feats)   # example inputs
        ref_out, src_out = fast(ref_points, src_points, ref_feats, src_feats)    # any inputs of the same shapes
    """

    def __init__(self, model, ref_points, src_points, ref_feats, src_feats, warmup=2, throughput=False):
        self.model = model
        self.throughput = bool(throughput)  # captured with the transformer's throughput schedule (several graphs in flight)
        self.static_in = [torch.empty_like(t) for t in (ref_points, src_points, ref_feats, src_feats)]
        for s, t in zip(self.static_in, (ref_points, src_points, ref_feats, src_feats)):
            s.copy_(t)
        self.warmup = warmup
        self.recapture()

    def _weight_versions(self):
        return [t._version for t in self._weights]

    def stale(self):
        """True when a parameter or buffer of the model changed in place since the graph was captured."""
        return self._weight_versions() != self._captured_versions

    @torch.no_grad()
    def recapture(self):
        self._weights = list(self.model.parameters()) + list(self.model.buffers())
        self._captured_versions = self._weight_versions()
        self.captures = getattr(self, 'captures', 0) + 1
        sched = getattr(self.model, 'transformer', None)
        previous = getattr(sched, 'throughput', None)
        if previous is not None:
            sched.throughput = self.throughput
        try:
            self._capture()
        finally:
            if previous is not None:
                sched.throughput = previous

    def _capture(self):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):  # warm-up off the capture: lazy cuBLAS / function-attribute initialisation
            for _ in range(self.warmup):
                self.model(*self.static_in)
        torch.cuda.current_stream().wait_stream(side)
        # the embedding kernels' overflow flag (a cloud larger than cheb_d_index_max) travels to pinned host memory at the
        # end of every replay; the NEXT call reads it there without synchronising
        emb = getattr(self.model, 'embedding', None)
        cache = getattr(emb, '_cheb_cache', None)
        self._flag_dev = cache.overflow if cache is not None else None
        self._flag_host = torch.zeros(1, dtype=torch.int32).pin_memory() if self._flag_dev is not None else None
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.static_out = self.model(*self.static_in)
            if self._flag_dev is not None:
                self._flag_host.copy_(self._flag_dev, non_blocking=True)

    @torch.no_grad()
    def __call__(self, ref_points, src_points, ref_feats, src_feats):
        for s, t in zip(self.static_in, (ref_points, src_points, ref_feats, src_feats)):
            if s.shape != t.shape:
                raise RuntimeError(f'graphed forward was captured for shape {tuple(s.shape)}, got {tuple(t.shape)}')
            if s.data_ptr() != t.data_ptr():
                s.copy_(t, non_blocking=True)
        self.replay()
        return self.static_out

    def replay(self):
        """graph.replay() on the current stream, after recapturing if the weights changed since the capture (the
        recapture synchronises the device: the graph being replaced may still be running)."""
        if self.stale():
            torch.cuda.synchronize()
            self.recapture()
        if self._flag_host is not None and int(self._flag_host[0]) != 0:
            raise RuntimeError('an earlier replay of this graph saw a cloud larger than the embedding kernels were prepared '
                               f'for (cheb_d_index_max = {self.model.embedding.cheb_d_index_max:g} index units): its outputs '
                               'used clamped distances.  Set model.embedding.cheb_d_index_max (<= 210) and recapture()')
        self.graph.replay()


class PipelinedGeometricTransformer:
    """Registration pairs that ARRIVE IN HOST MEMORY and whose features are wanted back in host memory.

    At n ~ 1k a pair's inputs are 8.4 MB (two (n, 1024) fp32 feature matrices): ~0.35 ms over PCIe against 2.1 ms of
    compute, so a serial copy -> forward -> copy loop leaves the GPU idle for 15 % of the time.  `depth` CUDA-graph
    instances of the forward (each with its own static buffers) are kept in flight instead: the upload of pair p+1 and
    the download of pair p-1 run on copy streams while pair p computes.

        pipe = PipelinedGeometricTransformer(model, ref_points, src_points, ref_feats, src_feats)  # example DEVICE inputs
        for ref_out, src_out in pipe.map(host_pairs):    # iterable of 4-tuples of (pinned) host tensors
            ...                                          # pinned host tensors, valid until the NEXT result is requested
                                                         # (their slot is re-enqueued then): clone what must outlive it

    Results come back in order and are bit-identical to `model(...)` on the same inputs under the same schedule
    (`model.transformer.throughput = True`, which the slots' graphs are captured with when they overlap)."""

    def __init__(self, model, ref_points, src_points, ref_feats, src_feats, depth=2, overlap_compute=None):
        if depth < 1:
            raise ValueError('depth must be >= 1')
        if overlap_compute is None:
            overlap_compute = os.environ.get('GEOTX_PIPE_OVERLAP', '1') != '0'
        # overlap_compute (GEOTX_PIPE_OVERLAP, default on): every slot replays on its own stream, so that the small kernels
        # of one pair (the cross blocks keep few SMs busy) run under the GPU-filling kernels of the next; results are
        # still handed out in order.  n = 1024, host to host: 692 pairs/s without, 703 / 747 / 761 at depth 2 / 3 / 4.
        self.compute_streams = [torch.cuda.Stream() for _ in range(depth)] if overlap_compute and depth > 1 else None
        self.slots = []
        for _ in range(depth):
            g = GraphedGeometricTransformer(model, ref_points, src_points, ref_feats, src_feats,
                                            throughput=self.compute_streams is not None)
            host_out = tuple(torch.empty(o.shape, dtype=o.dtype).pin_memory() for o in g.static_out)
            ev = {k: torch.cuda.Event() for k in ('h2d', 'compute', 'd2h')}
            self.slots.append((g, host_out, ev))
        self.copy_in, self.copy_out = torch.cuda.Stream(), torch.cuda.Stream()

    @torch.no_grad()
    def map(self, pairs):
        caller = torch.cuda.current_stream()
        depth = len(self.slots)
        if self.compute_streams is not None:
            for st in self.compute_streams:
                st.wait_stream(caller)
        pending = []  # slots whose results have not been handed out yet, oldest first
        for p, pair in enumerate(pairs):
            g, host_out, ev = self.slots[p % depth]
            compute = caller if self.compute_streams is None else self.compute_streams[p % depth]
            if len(pending) == depth:  # this slot's previous result must be handed out before its buffers are reused
                done = pending.pop(0)
                done[2]['d2h'].synchronize()
                yield done[1]
            if len(pair) != 4:
                raise RuntimeError('each pair is (ref_points, src_points, ref_feats, src_feats)')
            self.copy_in.wait_event(ev['compute'])  # the slot's last replay no longer reads its static inputs
            with torch.cuda.stream(self.copy_in):
                for s, t in zip(g.static_in, pair):
                    if s.shape != t.shape:
                        raise RuntimeError(f'pipeline was captured for shape {tuple(s.shape)}, got {tuple(t.shape)}')
                    s.copy_(t, non_blocking=True)
                ev['h2d'].record(self.copy_in)
            compute.wait_event(ev['h2d'])
            compute.wait_event(ev['d2h'])  # the slot's last download no longer reads its static outputs
            with torch.cuda.stream(compute):
                g.replay()
            ev['compute'].record(compute)
            self.copy_out.wait_event(ev['compute'])
            with torch.cuda.stream(self.copy_out):
                for h, o in zip(host_out, g.static_out):
                    h.copy_(o, non_blocking=True)
                ev['d2h'].record(self.copy_out)
            pending.append((g, host_out, ev))
        for done in pending:
            done[2]['d2h'].synchronize()
            yield done[1]


class SuperPointMatching(nn.Module):
    """Drop-in for geotransformer/modules/geotransformer/superpoint_matching.py:7-51, on
    `geotx_superpoint_matching` (exact fp32 scores, dual normalisation and global top-K select in hand-written CUDA;
    one (n,m) buffer instead of the reference's five, no full sort).

    forward(ref_feats, src_feats, ref_masks=None, src_masks=None) -> (ref_corr_indices, src_corr_indices, corr_scores)
    with masks True = keep, as in the reference (:19-20).  Without masks nothing synchronises with the host (the
    number of correspondences is min(num_correspondences, n*m), known on the host); with masks the count depends on
    device data and is read back, like the reference's own `torch.nonzero` (:32-33).
    `normalize=True` (extension) fuses model.py:141-142's F.normalize; `match_normalized()` also returns the unit rows."""

    def __init__(self, num_correspondences, dual_normalization=True):
        super().__init__()
        self.num_correspondences = num_correspondences
        self.dual_normalization = dual_normalization

    def _run(self, ref_feats, src_feats, ref_masks, src_masks, normalize):
        out = ops.superpoint_matching(ref_feats.contiguous(), src_feats.contiguous(), self.num_correspondences,
                                      self.dual_normalization, ref_masks, src_masks, normalize=normalize)
        if ref_masks is None and src_masks is None:
            count = min(self.num_correspondences, ref_feats.shape[0] * src_feats.shape[0])
        else:
            count = int(out['count'].item())
        return out, count

    @torch.no_grad()
    def forward(self, ref_feats, src_feats, ref_masks=None, src_masks=None, normalize=False):
        out, count = self._run(ref_feats, src_feats, ref_masks, src_masks, normalize)
        return out['ref_idx'][:count], out['src_idx'][:count], out['scores'][:count]

    @torch.no_grad()
    def match_normalized(self, ref_feats, src_feats, ref_masks=None, src_masks=None):
        """F.normalize + matching in one call: returns (ref_unit, src_unit, ref_idx, src_idx, scores), i.e. what
        model.py:141-157 puts into output_dict['ref_feats_c'], ['src_feats_c'] and the node correspondences."""
        out, count = self._run(ref_feats, src_feats, ref_masks, src_masks, True)
        return out['ref_unit'], out['src_unit'], out['ref_idx'][:count], out['src_idx'][:count], out['scores'][:count]


@torch.no_grad()
def point_to_node_partition(points, nodes, point_limit, return_count=False):
    """Drop-in for geotransformer/modules/ops/pointcloud_partition.py:61-107: same arguments, same
    tuple, on `geotx_point_to_node_partition`."""
    return ops.point_to_node_partition(points.contiguous(), nodes.contiguous(), point_limit, return_count)


class LearnableLogOptimalTransport(nn.Module):
    """Drop-in for geotransformer/modules/sinkhorn/learnable_sinkhorn.py:5-66: same constructor,
    the same single `alpha` parameter (state_dict key `alpha`), same forward; all iterations in one launch of
    `geotx_log_optimal_transport`.  Forward-only, like the rest of this package."""

    def __init__(self, num_iterations, inf=1e12):
        super().__init__()
        self.num_iterations = num_iterations
        self.register_parameter('alpha', torch.nn.Parameter(torch.tensor(1.0)))
        self.inf = inf

    def forward(self, scores, row_masks=None, col_masks=None):
        _forward_only_guard(self, scores)
        return ops.log_optimal_transport(scores.contiguous(), self.alpha.detach(), self.num_iterations,
                                         None if row_masks is None else row_masks.contiguous(),
                                         None if col_masks is None else col_masks.contiguous(), self.inf)

    def __repr__(self):
        return self.__class__.__name__ + '(num_iterations={})'.format(self.num_iterations)


@torch.no_grad()
def weighted_procrustes(src_points, ref_points, weights=None, weight_thresh=0.0, eps=1e-5, return_transform=False):
    """Drop-in for geotransformer/modules/registration/procrustes.py:6-73: same arguments and
    return values ((R, t) or the 4x4 transform; 2-D inputs are treated as a batch of one and squeezed again), with the
    3x3 SVD on the device (`geotx_weighted_procrustes`) instead of the reference's host round trip."""
    squeeze_first = src_points.ndim == 2  # :31-38
    if squeeze_first:
        src_points, ref_points = src_points.unsqueeze(0), ref_points.unsqueeze(0)
        if weights is not None:
            weights = weights.unsqueeze(0)
    transform = ops.weighted_procrustes(src_points.contiguous(), ref_points.contiguous(),
                                        None if weights is None else weights.contiguous(), weight_thresh, eps)
    if return_transform:  # :62-67
        return transform.squeeze(0) if squeeze_first else transform
    R, t = transform[:, :3, :3], transform[:, :3, 3]  # :68-72
    return (R.squeeze(0), t.squeeze(0)) if squeeze_first else (R, t)


class WeightedProcrustes(nn.Module):
    """Drop-in for procrustes.py:76-91."""

    def __init__(self, weight_thresh=0.0, eps=1e-5, return_transform=False):
        super().__init__()
        self.weight_thresh = weight_thresh
        self.eps = eps
        self.return_transform = return_transform

    def forward(self, src_points, tgt_points, weights=None):
        return weighted_procrustes(src_points, tgt_points, weights=weights, weight_thresh=self.weight_thresh,
                                   eps=self.eps, return_transform=self.return_transform)


class LocalGlobalRegistration(nn.Module):
    """Drop-in for geotransformer/modules/geotransformer/local_global_registration.py:11-235: same
    constructor, same forward arguments and results -- (ref_corr_points (C,3), src_corr_points (C,3), corr_scores (C,),
    estimated_transform (4,4)).  One launch sequence of five kernels (K11) replaces the reference's scatter / nonzero /
    host-side chunk loop / six Procrustes calls with a host SVD each; the only synchronisation is the read of C that
    sizes the returned views (`ops.local_global_registration` is the fixed-capacity form without it).
    use_dustbin = True and correspondence_limit are not shipped by any configuration (experiments/*/config.py:116-125)
    and raise; the reference's own dustbin branch cannot run either (local_global_registration.py:78-79 slices the
    correspondence matrix to (B, 1) before the (B, K, K) mask is applied)."""

    def __init__(self, k, acceptance_radius, mutual=True, confidence_threshold=0.05, use_dustbin=False,
                 use_global_score=False, correspondence_threshold=3, correspondence_limit=None,
                 num_refinement_steps=5):
        super().__init__()
        if use_dustbin:
            raise NotImplementedError('LocalGlobalRegistration: use_dustbin=True is not built on the B200 path')
        if correspondence_limit is not None:
            raise NotImplementedError('LocalGlobalRegistration: correspondence_limit is not built on the B200 path')
        self.k = k
        self.acceptance_radius = acceptance_radius
        self.mutual = mutual
        self.confidence_threshold = confidence_threshold
        self.use_dustbin = use_dustbin
        self.use_global_score = use_global_score
        self.correspondence_threshold = correspondence_threshold
        self.correspondence_limit = correspondence_limit
        self.num_refinement_steps = num_refinement_steps

    @torch.no_grad()
    def forward(self, ref_knn_points, src_knn_points, ref_knn_masks, src_knn_masks, score_mat, global_scores):
        ref_corr, src_corr, scores, count, transform = ops.local_global_registration(
            ref_knn_points.contiguous(), src_knn_points.contiguous(), ref_knn_masks.contiguous(),
            src_knn_masks.contiguous(), score_mat if score_mat.stride(-1) == 1 else score_mat.contiguous(),
            global_scores.contiguous() if self.use_global_score else None, self.k, self.acceptance_radius,
            mutual=self.mutual, confidence_threshold=self.confidence_threshold,
            correspondence_threshold=self.correspondence_threshold, num_refinement_steps=self.num_refinement_steps)
        c = int(count.item())
        return ref_corr[:c], src_corr[:c], scores[:c], transform


@torch.no_grad()
def patch_matching_scores(ref_feats_f, src_feats_f, ref_node_knn_indices, src_node_knn_indices, ref_node_corr_indices,
                          src_node_corr_indices):
    """The inline glue of experiments/*/model.py:169-170,176-179,187-188 as one call (K12): the (P, K, K) matching
    scores of the node correspondences' patches, ready for LearnableLogOptimalTransport.  The reference has no function
    of this name; a maintainer replaces those eight lines of `forward` by this call."""
    return ops.patch_scores(ref_feats_f.contiguous(), src_feats_f.contiguous(), ref_node_knn_indices.contiguous(),
                            src_node_knn_indices.contiguous(), ref_node_corr_indices.contiguous(),
                            src_node_corr_indices.contiguous())


@torch.no_grad()
def radius_search(q_points, s_points, q_lengths, s_lengths, radius, neighbor_limit):
    """Drop-in for geotransformer/modules/ops/radius_search.py:7-27 for points that already live on
    the GPU: same arguments, same result -- (N, min(max_count, neighbor_limit)) int64, filled with M.  The reference
    runs nanoflann on the host inside the data loader; here the width of the result depends on device data, so this
    wrapper reads one integer back (the only synchronisation); `ops.radius_neighbors` returns the fixed-width form
    without it.  neighbor_limit <= 0 is the reference's "no truncation" (radius_search.py:25): the search runs once to
    learn the widest neighbourhood and, when that exceeds the first guess, once more at that width; the kernel keeps
    at most GEOTX_MAX_NEIGHBOR_LIMIT = 512 neighbours per query and a wider neighbourhood is an error."""
    q_points, s_points = q_points.contiguous(), s_points.contiguous()
    if neighbor_limit <= 0:
        guess = 64
        neighbors, _, max_count = ops.radius_neighbors(q_points, s_points, q_lengths, s_lengths, radius, guess)
        width = int(max_count.item())
        if width > ops.MAX_NEIGHBOR_LIMIT:
            raise RuntimeError(f'radius_search: a query has {width} neighbours within radius={radius}; the GPU path '
                               f'keeps at most {ops.MAX_NEIGHBOR_LIMIT} per query -- pass a neighbor_limit')
        if width > guess:
            neighbors, _, _ = ops.radius_neighbors(q_points, s_points, q_lengths, s_lengths, radius, width)
        return neighbors[:, :width]
    neighbors, _, max_count = ops.radius_neighbors(q_points, s_points, q_lengths, s_lengths,
                                                   radius, neighbor_limit)
    width = min(int(max_count.item()), int(neighbor_limit))
    return neighbors[:, :width] if width < neighbor_limit else neighbors


@torch.no_grad()
def grid_subsample(points, lengths, voxel_size, max_cells=1 << 24):
    """Drop-in for geotransformer/modules/ops/grid_subsample.py:9-22 (plain path, length_ref = length_src = None;
    SURVEY.md 8(f)-4) for points that already live on the GPU.  Returns (s_points (M,3) CUDA, s_lengths (B,) int64 on
    the HOST, where the reference keeps them); M depends on device data, so the lengths are read back once.  The
    voxels come out in ascending voxel index instead of the reference's hash-table order."""
    buf, s_lengths = ops.grid_subsample(points.contiguous(), lengths, voxel_size, max_cells)
    s_lengths = s_lengths.cpu()
    if bool((s_lengths < 0).any()):
        raise RuntimeError(f'grid_subsample: the voxel grid of a batch element has more than max_cells={max_cells} '
                           f'cells at voxel_size={voxel_size}; pass a larger max_cells')
    return buf[:int(s_lengths.sum())], s_lengths


@torch.no_grad()
def precompute_data_stack_mode(points, lengths, num_stages, voxel_size, radius, neighbor_limits, max_cells=1 << 24):
    """Drop-in for geotransformer/utils/data.py:13-81 (the KPConv pyramid the collate function builds on the host, one
    nanoflann tree per search) for points that already live on the GPU: same arguments, same dictionary
    ('points', 'lengths', 'neighbors', 'subsampling', 'upsampling': lists over the stages), built from
    `grid_subsample` (K10) and `radius_search` (K9).  Point and index tensors are CUDA tensors, lengths stay on the host
    as in the reference.  Subsampled stages list their voxels in ascending voxel index (the reference: hash-table
    order), so indices into stages >= 1 are a permutation of the reference's."""
    assert num_stages == len(neighbor_limits)
    points_list, lengths_list, neighbors_list, subsampling_list, upsampling_list = [], [], [], [], []
    lengths = torch.as_tensor(lengths, dtype=torch.int64).cpu()
    for i in range(num_stages):  # data.py:22-30
        if i > 0:
            points, lengths = grid_subsample(points, lengths, voxel_size=voxel_size, max_cells=max_cells)
        points_list.append(points)
        lengths_list.append(lengths)
        voxel_size *= 2
    # data.py:33-72.  All searches are enqueued with fixed-width results; the widths of the reference's results
    # (min(max_count, limit), radius_search.py:24-27) are read back together afterwards: one synchronisation for the
    # ten searches instead of one each.
    pending = []  # (list to fill, fixed-width neighbors, max_count, limit)

    def search(dst, q_points, s_points, q_lengths, s_lengths, r, limit):
        if limit <= 0:
            raise RuntimeError('precompute_data_stack_mode: neighbor limits must be positive on the GPU path')
        nb, _, max_count = ops.radius_neighbors(q_points.contiguous(), s_points.contiguous(), q_lengths, s_lengths, r,
                                                limit)
        pending.append((dst, nb, max_count, int(limit)))

    for i in range(num_stages):
        cur_points, cur_lengths = points_list[i], lengths_list[i]
        search(neighbors_list, cur_points, cur_points, cur_lengths, cur_lengths, radius, neighbor_limits[i])
        if i < num_stages - 1:
            sub_points, sub_lengths = points_list[i + 1], lengths_list[i + 1]
            search(subsampling_list, sub_points, cur_points, sub_lengths, cur_lengths, radius, neighbor_limits[i])
            search(upsampling_list, cur_points, sub_points, cur_lengths, sub_lengths, radius * 2, neighbor_limits[i + 1])
        radius *= 2
    widths = torch.cat([m for _, _, m, _ in pending]).tolist()
    for (dst, nb, _, limit), width in zip(pending, widths):
        dst.append(nb[:, :width] if width < limit else nb)
    return {'points': points_list, 'lengths': lengths_list, 'neighbors': neighbors_list,
            'subsampling': subsampling_list, 'upsampling': upsampling_list}
