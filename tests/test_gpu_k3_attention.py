"""-m gpu: K3 (folded relative-position scores + flash attention) through the C ABI.  Floating-point kernels: compared
with a plain torch fp32 evaluation of the reference formula (rpe_transformer.py:51-70, vanilla_transformer.py:50-68)
computed on the CPU by the oracle."""
import pytest
import torch

from gpu_common import DEV, rel

pytestmark = pytest.mark.gpu


def _rand(g, *shape):
    return torch.randn(*shape, generator=g)


@pytest.mark.parametrize('C,H,n,m', [(256, 4, 97, 97), (128, 4, 130, 130), (256, 8, 64, 64), (64, 2, 50, 50)])
@pytest.mark.parametrize('e_dtype', [torch.float16, torch.bfloat16, torch.float32])
def test_rpe_scores_folded_query(pkg, C, H, n, m, e_dtype):
    """scores_p[b,h,i,j] = (W_p[h]^T q[b,h,i]) . E[b,i,j]  ==  einsum(q, proj_p(E)) - q.b_p  (SURVEY.md 8a-7)."""
    g = torch.Generator().manual_seed(C + n)
    B, c = 2, C // H
    q = _rand(g, B, n, C)
    E = _rand(g, B, n, m, C)
    wp, bp = _rand(g, C, C) / C ** 0.5, _rand(g, C)
    Eq = E.to(e_dtype).float()  # the kernel sees the stored precision of E
    p = torch.nn.functional.linear(Eq, wp, bp).view(B, n, m, H, c).permute(0, 3, 1, 2, 4)
    qh = q.view(B, n, H, c).permute(0, 2, 1, 3)
    exp = torch.einsum('bhnc,bhnmc->bhnm', qh, p) - torch.einsum('bhnc,hc->bhn', qh, bp.view(H, c)).unsqueeze(-1)
    att = pkg.RPEMultiHeadAttention(C, H).to(DEV)
    with torch.no_grad():
        att.proj_p.weight.copy_(wp)
        att.proj_p.bias.copy_(bp)
        qt = att.folded_query(q.to(DEV))
        got = pkg.ops.rpe_scores(qt, E.to(DEV).to(e_dtype))
        # slab interface: rows [10, 40) only
        part = torch.zeros_like(got)
        pkg.ops.rpe_scores(qt, E[:, 10:40].contiguous().to(DEV).to(e_dtype), n=n, row_begin=10, out=part)
    assert rel(got, exp) <= 5e-5
    assert torch.equal(part[:, :, 10:40], got[:, :, 10:40]) and float(part[:, :, :10].abs().max()) == 0.0


@pytest.mark.parametrize('c,H,n,m', [(64, 4, 200, 160), (32, 4, 77, 300), (64, 4, 64, 64), (16, 2, 33, 65)])
@pytest.mark.parametrize('with_p,with_mask', [(False, False), (True, False), (True, True), (False, True)])
def test_attention_vs_torch(pkg, oracle, c, H, n, m, with_p, with_mask):
    """softmax((q k^T + scores_p)/sqrt(c) [mask]) v ; TF32 tensor-core contractions -> 2e-3."""
    g = torch.Generator().manual_seed(c * 1000 + n)
    B, D = 2, H * c
    q, k, v = _rand(g, B, n, D), _rand(g, B, m, D), _rand(g, B, m, D)
    sp = _rand(g, B, H, n, m) * 3.0 if with_p else None
    mask = (torch.rand(B, m, generator=g) < 0.3) if with_mask else None
    if mask is not None:
        mask[:, 0] = False
    qh, kh, vh = (t.view(B, -1, H, c).permute(0, 2, 1, 3) for t in (q, k, v))
    s = torch.einsum('bhnc,bhmc->bhnm', qh, kh)
    if sp is not None:
        s = s + sp
    s = s / c ** 0.5
    if mask is not None:
        s = s.masked_fill(mask[:, None, None], float('-inf'))
    exp = torch.matmul(torch.softmax(s, -1), vh).permute(0, 2, 1, 3).reshape(B, n, D)
    got = pkg.ops.attention(q.to(DEV), k.to(DEV), v.to(DEV), H, None if sp is None else sp.to(DEV),
                            None if mask is None else mask.to(DEV))
    assert rel(got, exp) <= 2e-3
    # split_keys=False: one CTA walks every key tile: same math, no cluster split/merge
    one = pkg.ops.attention(q.to(DEV), k.to(DEV), v.to(DEV), H, None if sp is None else sp.to(DEV),
                            None if mask is None else mask.to(DEV), split_keys=False)
    assert rel(one, exp) <= 2e-3 and rel(one, got) <= 1e-3  # P is rounded to TF32 against a different running max


@pytest.mark.parametrize('c,H,n,m', [(64, 4, 200, 160), (32, 4, 77, 300), (64, 4, 64, 64), (16, 2, 33, 65), (64, 4, 1024, 1024),
                                     (64, 1, 130, 1000), (32, 4, 2100, 1900)])
@pytest.mark.parametrize('with_p,with_mask', [(False, False), (True, False), (True, True), (False, True)])
def test_attention_half_operands_vs_torch(pkg, c, H, n, m, with_p, with_mask):
    """K3h (geotx_attention_f16): q, k, v handed over as half.  Against torch fp32 on the SAME half-rounded operands the
    only differences are P's rounding to half and the fp32 summation order -> 1e-3; against K3b (fp32 entry, which rounds
    to half itself) likewise.  Shapes: ragged tiles, key counts that are / are not multiples of 4 (16-byte / 4-byte
    score copies), both CTA shapes (the last case has >= 128 query tiles of 128), cluster splits 1..8."""
    g = torch.Generator().manual_seed(c * 1000 + n + m)
    B, D = (2, H * c) if n < 1000 else (1, H * c)
    q, k, v = (_rand(g, B, r, D).to(torch.float16) for r in (n, m, m))
    sp = _rand(g, B, H, n, m) * 3.0 if with_p else None
    mask = (torch.rand(B, m, generator=g) < 0.3) if with_mask else None
    if mask is not None:
        mask[:, 0] = False
    qh, kh, vh = (t.float().view(B, -1, H, c).permute(0, 2, 1, 3) for t in (q, k, v))
    s = torch.einsum('bhnc,bhmc->bhnm', qh, kh)
    if sp is not None:
        s = s + sp
    s = s / c ** 0.5
    if mask is not None:
        s = s.masked_fill(mask[:, None, None], float('-inf'))
    exp = torch.matmul(torch.softmax(s, -1), vh).permute(0, 2, 1, 3).reshape(B, n, D)
    dev = lambda t: None if t is None else t.to(DEV)  # noqa: E731
    got = pkg.ops.attention(dev(q), dev(k), dev(v), H, dev(sp), dev(mask))
    assert got.dtype == torch.float32 and rel(got, exp) <= 1e-3
    one = pkg.ops.attention(dev(q), dev(k), dev(v), H, dev(sp), dev(mask), split_keys=False)
    assert rel(one, exp) <= 1e-3
    old = pkg.ops.attention(dev(q.float()), dev(k.float()), dev(v.float()), H, dev(sp), dev(mask))
    assert rel(got, old) <= 1e-3


def test_attention_half_column_blocks_and_edge_cases(pkg):
    """q | k | v as column blocks of one half matrix (what linear_f16(half_cols=) writes); all keys masked -> NaN like
    torch; mixed dtypes and a non-positive scale are refused."""
    g = torch.Generator().manual_seed(11)
    n, H, c = 300, 4, 64
    C = H * c
    y = torch.randn(1, n, 3 * C, generator=g).to(DEV).to(torch.float16)
    q, k, v = y[:, :, :C], y[:, :, C:2 * C], y[:, :, 2 * C:]
    got = pkg.ops.attention(q, k, v, H)
    assert torch.equal(got, pkg.ops.attention(q.contiguous(), k.contiguous(), v.contiguous(), H))
    nan = pkg.ops.attention(q, k, v, H, None, torch.ones(1, n, dtype=torch.bool, device=DEV))
    assert torch.isnan(nan).all()
    with pytest.raises(RuntimeError, match='float16'):
        pkg.ops.attention(q, k.float(), v, H)
    with pytest.raises(RuntimeError, match='scale must be positive'):
        pkg.ops.attention(q, k, v, H, scale=-1.0)


def test_attention_all_keys_masked_is_nan_like_torch(pkg):
    q = torch.randn(1, 8, 64)
    got = pkg.ops.attention(q.to(DEV), q.to(DEV), q.to(DEV), 2, None, torch.ones(1, 8, dtype=torch.bool).to(DEV))
    assert torch.isnan(got).all()  # torch.softmax over a row of -inf is NaN too (rpe_transformer.py:63-65)


def test_attention_rejects_cpu_and_bad_head_dim(pkg):
    q = torch.randn(1, 8, 96)
    with pytest.raises(RuntimeError, match='CUDA tensor'):
        pkg.ops.attention(q, q, q, 2)
    with pytest.raises(RuntimeError, match='per-head dim'):
        pkg.ops.attention(q.to(DEV), q.to(DEV), q.to(DEV), 2)  # c = 48


@pytest.mark.parametrize('rows,C', [(1, 256), (77, 128), (300, 32), (5, 1024)])
def test_add_layernorm_vs_torch(pkg, rows, C):
    """K4: LayerNorm(x + residual) (rpe_transformer.py:103, output_layer.py:20) against torch fp32 on the CPU."""
    g = torch.Generator().manual_seed(rows + C)
    x, r = torch.randn(1, rows, C, generator=g) * 3, torch.randn(1, rows, C, generator=g)
    w, b = torch.randn(C, generator=g), torch.randn(C, generator=g)
    exp = torch.nn.functional.layer_norm(x + r, (C,), w, b, 1e-5)
    got = pkg.ops.add_layernorm(x.to(DEV), r.to(DEV), w.to(DEV), b.to(DEV), 1e-5)
    assert rel(got, exp) <= 2e-6
    exp1 = torch.nn.functional.layer_norm(x, (C,), w, b, 1e-5)
    assert rel(pkg.ops.add_layernorm(x.to(DEV), None, w.to(DEV), b.to(DEV)), exp1) <= 2e-6
    with pytest.raises(RuntimeError, match='multiple of 4'):
        pkg.ops.add_layernorm(torch.randn(2, 6, device=DEV), None, torch.ones(6, device=DEV), torch.ones(6, device=DEV))


def test_strided_operands_of_a_fused_projection(pkg):
    """q | k | v | folded query taken as column blocks of one wider row-major matrix (no copies)."""
    g = torch.Generator().manual_seed(5)
    B, n, H, c = 1, 150, 4, 64
    C = H * c
    y = torch.randn(B, n, 3 * C + H * C, generator=g).to(DEV)
    E = torch.randn(B, n, n, C, generator=g).to(DEV).to(torch.bfloat16)
    q, k, v = y[:, :, :C], y[:, :, C:2 * C], y[:, :, 2 * C:3 * C]
    qt = y[:, :, 3 * C:].unflatten(2, (H, C)).permute(0, 2, 1, 3)
    sp = pkg.ops.rpe_scores(qt, E)
    assert torch.equal(sp, pkg.ops.rpe_scores(qt.contiguous(), E))
    got = pkg.ops.attention(q, k, v, H, sp)
    assert torch.equal(got, pkg.ops.attention(q.contiguous(), k.contiguous(), v.contiguous(), H, sp))


@pytest.mark.parametrize('with_masks', [False, True])
def test_rpe_attention_with_key_weights_and_attention_factors(pkg, oracle, with_masks):
    """rpe_transformer.py:58-64: attention_factors and key_weights rescale the summed scores BEFORE the softmax, so the
    q . b_p term that the folded query drops is no longer constant over the keys: the module must restore it.
    proj_p.bias is made large so that a missing term cannot hide inside the tolerance."""
    C, H, n, m, B = 128, 4, 70, 70, 2
    g = torch.Generator().manual_seed(11)
    sd = {f'a.proj_{x}.weight': _rand(g, C, C) / C ** 0.5 for x in 'qkvp'}
    sd.update({f'a.proj_{x}.bias': _rand(g, C) * (3.0 if x == 'p' else 0.1) for x in 'qkvp'})
    x = _rand(g, B, n, C)
    E = _rand(g, B, n, m, C)
    kw = torch.rand(B, m, generator=g) + 0.5
    af = torch.rand(B, n, m, generator=g) + 0.5
    masks = (torch.rand(B, m, generator=g) < 0.2) if with_masks else None
    exp, exp_scores = oracle.multi_head_attention(x, x, x, sd, 'a.', H, embed_qk=E, key_weights=kw, key_masks=masks,
                                                  attention_factors=af)
    att = pkg.RPEMultiHeadAttention(C, H).to(DEV).eval()
    att.load_state_dict({k[2:]: v for k, v in sd.items()}, strict=True)
    att.need_scores = True
    with torch.no_grad():
        got, scores = att(x.to(DEV), x.to(DEV), x.to(DEV), E.to(DEV), key_weights=kw.to(DEV),
                          key_masks=None if masks is None else masks.to(DEV), attention_factors=af.to(DEV))
    assert rel(got, exp) <= 1e-4 and rel(scores, exp_scores) <= 1e-4
    # and without the modifiers the bias still has no effect (fast path, folded query)
    with torch.no_grad():
        att.need_scores = False
        fast, _ = att(x.to(DEV), x.to(DEV), x.to(DEV), E.to(DEV))
    plain, _ = oracle.multi_head_attention(x, x, x, sd, 'a.', H, embed_qk=E)
    assert rel(fast, plain) <= 2e-3  # half operands in the flash kernel
