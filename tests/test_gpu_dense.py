"""-m gpu: the hand-written dense glue of the fused schedule, through the C ABI: K13 `geotx_linear_f16` (in_proj,
out_proj, q|k|v|folded-query projections) and K4f `geotx_post_attention` (linear + residual + LayerNorm + FFN + residual
+ LayerNorm in one kernel).  Floating-point kernels: compared with the oracle's plain torch fp32 evaluation of the
reference lines (rpe_transformer.py:101-103, output_layer.py:15-21, geotransformer.py:140-141).
Tolerance: operands enter the contractions as IEEE halves (11-bit significand, like the TF32 library path they replace),
accumulation is fp32: 1e-3 Frobenius-relative per call (measured ~2e-4)."""
import pytest
import torch

from gpu_common import DEV, rel

pytestmark = pytest.mark.gpu
TOL = 1e-3


@pytest.mark.parametrize('T,K,N', [(2048, 256, 1792), (1024, 1024, 256), (717, 512, 256), (50, 128, 896), (4096, 256, 768),
                                   (1, 64, 64)])
def test_linear_vs_torch(pkg, T, K, N):
    g = torch.Generator().manual_seed(T + K + N)
    x = torch.randn(T, K, generator=g)
    w = torch.randn(N, K, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g)
    exp = torch.nn.functional.linear(x, w, b)
    assert torch.equal(pkg.ops.to_half(w.to(DEV)).cpu(), w.half())
    wh = pkg.ops.pack_linear(w.to(DEV))  # K13's operand: same halves, chunk-major swizzled order
    assert wh.dtype == torch.float16 and tuple(wh.shape) == (N, K)
    assert torch.equal(wh.cpu().flatten().sort().values, w.half().flatten().sort().values)
    got = pkg.ops.linear_f16(x.to(DEV), wh, b.to(DEV))
    assert tuple(got.shape) == (T, N) and rel(got, exp) <= TOL
    assert rel(pkg.ops.linear_f16(x.to(DEV), wh, None), exp - b) <= TOL
    assert rel(pkg.ops.linear_f16(x.to(DEV), wh, b.to(DEV), relu=True), exp.clamp_min(0)) <= TOL


@pytest.mark.parametrize('T,K,N,hc', [(2048, 256, 1792, 768), (717, 256, 768, 768), (50, 128, 896, 384), (1, 64, 128, 64)])
def test_linear_mixed_half_and_fp32_columns(pkg, T, K, N, hc):
    """geotx_linear_f16_mixed: the first hc columns of the result as half (exactly the rounding K3b applies to fp32
    q | k | v: round to nearest, saturating), the rest as fp32 -- both bit-identical to the plain call."""
    g = torch.Generator().manual_seed(T + K + N + hc)
    x = torch.randn(T, K, generator=g).to(DEV)
    w = torch.randn(N, K, generator=g) / K ** 0.5
    wh = pkg.ops.pack_linear(w.to(DEV))
    b = torch.randn(N, generator=g)
    b[0], b[1] = 1e6, -1e6  # columns 0 and 1 overflow half: must saturate to +-65504, not inf
    b = b.to(DEV)
    full = pkg.ops.linear_f16(x, wh, b)
    yh, rest = pkg.ops.linear_f16(x, wh, b, half_cols=hc)
    assert yh.dtype == torch.float16 and tuple(yh.shape) == (T, hc)
    assert torch.equal(yh, full[:, :hc].clamp(-65504.0, 65504.0).to(torch.float16))
    assert torch.isfinite(yh).all()
    if hc < N:
        assert rest.dtype == torch.float32 and torch.equal(rest, full[:, hc:])
    else:
        assert rest is None
    with pytest.raises(RuntimeError, match='half_cols'):
        pkg.ops.linear_f16(x, wh, b, half_cols=32)


def test_linear_strided_rows_and_slices(pkg):
    """Inputs / outputs that are column blocks and row slices of larger token buffers (how the fused schedule calls it)."""
    g = torch.Generator().manual_seed(4)
    big = torch.randn(1, 300, 512, generator=g).to(DEV)
    x = big[:, 40:240, 128:384]                     # (1, 200, 256): row stride 512, offset columns
    w = (torch.randn(192, 256, generator=g) / 16).to(DEV)
    b = torch.randn(192, generator=g).to(DEV)
    out_big = torch.full((1, 260, 448), -7.0, device=DEV)
    out = out_big[:, 10:210, 64:256]                # (1, 200, 192) view
    pkg.ops.linear_f16(x, pkg.ops.pack_linear(w), b, out=out)
    exp = torch.nn.functional.linear(x, w, b)
    assert rel(out, exp) <= TOL
    mask = torch.ones_like(out_big, dtype=torch.bool)
    mask[:, 10:210, 64:256] = False
    assert float((out_big[mask] + 7.0).abs().max()) == 0.0, 'wrote outside the output view'
    with pytest.raises(RuntimeError):
        pkg.ops.linear_f16(big[:, :, ::2], pkg.ops.pack_linear(w), b)  # last dimension not dense
    with pytest.raises(RuntimeError):
        pkg.ops.linear_f16(torch.randn(8, 100, device=DEV), torch.randn(64, 100, device=DEV).half(), None)  # K % 32


@pytest.mark.parametrize('C,T', [(256, 2048), (128, 4096), (256, 717), (128, 33), (256, 1)])
def test_post_attention_vs_oracle(pkg, oracle, C, T):
    sd = oracle.init_state_dict(8, 8, C, ['self'], seed=21)
    lp = 'transformer.layers.0.'
    g = torch.Generator().manual_seed(C + T)
    sd[lp + 'attention.norm.weight'] = torch.rand(C, generator=g) + 0.5   # non-trivial LayerNorm parameters
    sd[lp + 'attention.norm.bias'] = torch.randn(C, generator=g) * 0.1
    sd[lp + 'output.norm.weight'] = torch.rand(C, generator=g) + 0.5
    sd[lp + 'output.norm.bias'] = torch.randn(C, generator=g) * 0.1
    hidden = torch.randn(1, T, C, generator=g)
    x_in = torch.randn(1, T, C, generator=g)
    h = torch.nn.functional.linear(hidden, sd[lp + 'attention.linear.weight'], sd[lp + 'attention.linear.bias'])
    x1 = oracle._layer_norm(h + x_in, sd, lp + 'attention.norm.')
    exp = oracle.attention_output(x1, sd, lp + 'output.')
    d = {k: v.to(DEV) for k, v in sd.items()}
    half = {k: pkg.ops.pack_linear(d[lp + k]) for k in ('attention.linear.weight', 'output.expand.weight', 'output.squeeze.weight')}
    args = (half['attention.linear.weight'], d[lp + 'attention.linear.bias'], d[lp + 'attention.norm.weight'],
            d[lp + 'attention.norm.bias'], 1e-5, half['output.expand.weight'], d[lp + 'output.expand.bias'],
            half['output.squeeze.weight'], d[lp + 'output.squeeze.bias'], d[lp + 'output.norm.weight'],
            d[lp + 'output.norm.bias'], 1e-5)
    got = pkg.ops.post_attention(hidden.to(DEV), x_in.to(DEV), *args)
    err = rel(got, exp)
    print(f'C={C} T={T}: post_attention rel err {err:.3e}')
    assert err <= TOL
    if T >= 64:  # row slices of larger buffers, written in place of a slice
        buf = torch.full((1, T + 20, C), -7.0, device=DEV)
        lo, hi = 5, 5 + T // 2
        pkg.ops.post_attention(hidden.to(DEV)[:, lo:hi], x_in.to(DEV)[:, lo:hi], *args, out=buf[:, lo:hi])
        assert torch.equal(buf[:, lo:hi], got[:, lo:hi])
        assert float((buf[:, :lo] + 7.0).abs().max()) == 0.0 and float((buf[:, hi:] + 7.0).abs().max()) == 0.0


def test_fused_schedule_own_kernels_vs_library_gemms(pkg, oracle, monkeypatch):
    """The whole forward with the hand-written dense kernels (default) against the same forward on cuBLASLt TF32 GEMMs +
    K4 (GEOTX_DENSE=cublas, round 1's path) and against the oracle; and no library GEMM is left on the default path
    (checked through the C ABI's launch counter: 2 k-NN calls x 2 kernels + in_proj x 2 + out_proj + per block ...)."""
    cfg = dict(pkg.synth.CONFIGS['3dmatch'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    rp, sp, rf, sf, _ = pkg.synth.make_tie_free_pair('3dmatch', 320, 256, seed=pkg.synth.SEED)
    model = pkg.GeometricTransformer(**cfg).to(DEV).eval()
    model.load_state_dict(sd, strict=True)
    args = [t.to(DEV) for t in (rp, sp, rf, sf)]
    e0, e1 = oracle.geometric_transformer(sd, cfg, rp, sp, rf, sf)
    with torch.no_grad():
        model(*args)  # builds every cached derivative (half weights, coefficients)
        before = pkg._abi.launch_count
        own = [t.clone() for t in model(*args)]
        own_launches = pkg._abi.launch_count - before
        monkeypatch.setenv('GEOTX_DENSE', 'cublas')
        before = pkg._abi.launch_count
        lib = [t.clone() for t in model(*args)]
        lib_launches = pkg._abi.launch_count - before
    print(f'own kernels: {own_launches} launches through the C ABI per forward; with library GEMMs: {lib_launches} (+39 cuBLAS)')
    for o, l, e in zip(own, lib, (e0, e1)):
        assert rel(o, l) <= 1e-3 and rel(o, e) <= 2e-3 and rel(l, e) <= 2e-3
    # 3 self blocks: 1 projection + 2 scores + 2 attention + 1 post = 6; 3 cross blocks: 2 projections + 2 attention +
    # 2 post = 6; + 2 x (pack + k-NN) + 2 in_proj + 1 out_proj
    assert own_launches == 3 * 6 + 3 * 6 + 4 + 2 + 1
