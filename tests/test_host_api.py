"""CPU tests of the host-side mirror of the reference interface and of the C-ABI library surface."""
import ctypes
import os
import re

import pytest
import torch

from load_reference import load_reference, reference_available

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _expected_keys(blocks):
    keys = ['embedding.embedding.div_term', 'embedding.proj_d.weight', 'embedding.proj_d.bias',
            'embedding.proj_a.weight', 'embedding.proj_a.bias', 'in_proj.weight', 'in_proj.bias']
    for i, b in enumerate(blocks):
        p = f'transformer.layers.{i}.'
        for nm in ['proj_q', 'proj_k', 'proj_v'] + (['proj_p'] if b == 'self' else []):
            keys += [p + f'attention.attention.{nm}.weight', p + f'attention.attention.{nm}.bias']
        for nm in ['attention.linear', 'attention.norm', 'output.expand', 'output.squeeze', 'output.norm']:
            keys += [p + nm + '.weight', p + nm + '.bias']
    return keys + ['out_proj.weight', 'out_proj.bias']


@pytest.mark.parametrize('shape', ['3dmatch', 'kitti', 'modelnet'])
def test_state_dict_keys_and_shapes(pkg, oracle, shape):
    cfg = dict(pkg.synth.CONFIGS[shape])
    m = pkg.GeometricTransformer(**cfg)
    sd = m.state_dict()
    assert list(sd.keys()) == _expected_keys(cfg['blocks'])
    assert len(sd) == 111
    ref_sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    assert {k: tuple(v.shape) for k, v in sd.items()} == {k: tuple(v.shape) for k, v in ref_sd.items()}
    m.load_state_dict(ref_sd, strict=True)
    if shape == '3dmatch':
        assert sum(p.numel() for p in m.parameters()) == 3819776  # SURVEY.md 8b [measured]


@pytest.mark.skipif(not reference_available(), reason='/root/reference not present (GPU box)')
def test_state_dict_round_trip_with_reference(pkg):
    GT, GSE = load_reference()
    cfg = dict(pkg.synth.CONFIGS['3dmatch'])
    args = (cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['num_heads'], cfg['blocks'], cfg['sigma_d'],
            cfg['sigma_a'], cfg['angle_k'])
    ref = GT(*args)
    new = pkg.GeometricTransformer(*args)
    new.load_state_dict(ref.state_dict(), strict=True)
    ref.load_state_dict(new.state_dict(), strict=True)
    assert list(ref.state_dict().keys()) == list(new.state_dict().keys())
    # same construction order => same default init under the same seed
    torch.manual_seed(7351)
    a = GT(*args).state_dict()
    torch.manual_seed(7351)
    b = pkg.GeometricTransformer(*args).state_dict()
    assert all(torch.equal(a[k], b[k]) for k in a)


def test_constructor_errors_match_reference(pkg):
    with pytest.raises(ValueError, match='Unsupported reduction mode'):
        pkg.GeometricStructureEmbedding(256, 0.2, 15, 3, reduction_a='sum')
    with pytest.raises(ValueError, match='odd d_model'):
        pkg.SinusoidalPositionalEmbedding(255)
    with pytest.raises(ValueError, match='must be a multiple of `num_heads`'):
        pkg.RPEMultiHeadAttention(256, 3)
    with pytest.raises(ValueError, match='must be a multiple of `num_heads`'):
        pkg.MultiHeadAttention(100, 3)
    with pytest.raises(ValueError, match='Unsupported block type'):
        pkg.RPEConditionalTransformer(['self', 'mixed'], 256, 4)
    with pytest.raises(AssertionError):
        pkg.AttentionOutput(64, activation_fn='Swish')


def test_forward_only_and_no_cpu_fallback(pkg):
    m = pkg.GeometricTransformer(16, 16, 32, 2, ['self', 'cross'], 0.2, 15, 3)
    x = (torch.rand(1, 8, 3), torch.rand(1, 8, 3), torch.rand(1, 8, 16), torch.rand(1, 8, 16))
    with pytest.raises(RuntimeError, match='forward-only'):
        m(*x)
    with torch.no_grad():
        with pytest.raises(RuntimeError, match='CUDA tensor'):
            m(*x)
        with pytest.raises(RuntimeError, match='CUDA tensor'):
            m.embedding.get_embedding_indices(x[0])


def test_folded_query_identity_fp64(pkg, oracle):
    """sum_c q_c (W_p e + b_p)_c = (W_p^T q).e + q.b_p, and the bias term cancels in softmax."""
    torch.manual_seed(0)
    B, n, m, C, H = 2, 7, 9, 32, 4
    att = pkg.RPEMultiHeadAttention(C, H).double()
    sd = {'a.' + k: v for k, v in att.state_dict().items()}
    x = torch.randn(B, n, C, dtype=torch.float64)
    mem = torch.randn(B, m, C, dtype=torch.float64)
    E = torch.randn(B, n, m, C, dtype=torch.float64)
    hidden_ref, scores_ref = oracle.multi_head_attention(x, mem, mem, sd, 'a.', H, embed_qk=E)
    with torch.no_grad():
        q = att.proj_q(x)
        k = att.proj_k(mem)
        qt = att.folded_query(q)  # (B,H,n,C) view
        assert qt.shape == (B, H, n, C) and qt.stride(3) == 1 and qt.stride(2) == C
        scores_p = torch.einsum('bhnc,bnmc->bhnm', qt, E)
        c = C // H
        qh = q.view(B, n, H, c).permute(0, 2, 1, 3)
        kh = k.view(B, m, H, c).permute(0, 2, 1, 3)
        scores = torch.softmax((qh @ kh.transpose(-1, -2) + scores_p) / c ** 0.5, dim=-1)
    assert (scores - scores_ref).abs().max().item() < 1e-13
    # dropping proj_p.bias changes nothing
    sd2 = dict(sd)
    sd2['a.proj_p.bias'] = torch.zeros_like(sd['a.proj_p.bias'])
    _, scores_nb = oracle.multi_head_attention(x, mem, mem, sd2, 'a.', H, embed_qk=E)
    assert (scores_nb - scores_ref).abs().max().item() < 1e-13


def test_abi_library_exports_every_declared_symbol(pkg):
    header = open(os.path.join(ROOT, 'include', 'geotx_b200.h')).read()
    header = re.sub(r'/\*.*?\*/', '', header, flags=re.S)
    declared = sorted(set(re.findall(r'\b(geotx_[a-z0-9_]+)\s*\(', header)))
    assert len(declared) >= 11
    assert sorted(pkg._abi.SIGNATURES.keys()) == declared
    lib = ctypes.CDLL(pkg._abi.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), f'{name} declared in include/geotx_b200.h but not exported'
    loaded = pkg._abi.load()
    assert loaded.geotx_abi_version() == 1
    assert loaded.geotx_gse_geometry_bytes(2, 100, 3) == 16 * 2 * 100 * 4 + 16  # + the largest squared distance
    # K2c: 16 coefficients reach 13 index units, 32 reach 37; six distance sets of 16..128 coefficients
    assert [loaded.geotx_gse_cheb_ka_tiles(x) for x in (12.0, 13.5, 37.0, 38.0, 0.0)] == [1, 2, 2, 0, 0]
    assert [loaded.geotx_gse_cheb_d_range(s) for s in range(7)] == [13.0, 37.0, 64.0, 93.0, 150.0, 210.0, 0.0]
    assert loaded.geotx_gse_cheb_bytes(256, 1) == 2 * 16 * 256 * (1 + 1 + 2 + 3 + 4 + 6 + 8)
    assert loaded.geotx_gse_cheb_bytes(96, 1) == 0 and loaded.geotx_gse_cheb_bytes(128, 3) == 0
    assert loaded.geotx_gse_packed_weight_bytes(256) == 4 * 256 * 256 + 4 * 256
    assert loaded.geotx_gse_packed_weight_bytes(100) == 0


def test_abi_argument_errors_without_gpu(pkg):
    """Argument validation happens before any CUDA call, so it is testable on a CPU-only box."""
    lib = pkg._abi.load()
    buf = ctypes.create_string_buffer(4096)
    p = ctypes.cast(buf, ctypes.c_void_p)
    assert lib.geotx_gse_knn(p, 1, 3, 3, p, p, None) == -1  # n < k+1, as torch.topk raises in the reference
    assert b'out of range' in lib.geotx_last_error()
    assert lib.geotx_gse_knn(p, 1, 10, 9, p, p, None) == -1
    assert lib.geotx_gse_embed(p, 1, 8, 3, 0, 8, p, p, 64, 0.2, 1.0, 0, 1, p, None) == -1
    assert b'hidden_dim' in lib.geotx_last_error()
    assert lib.geotx_attention(p, p, p, None, None, 1, 4, 8, 8, 48, 192, 192, 192, 0.1, p, 1, None) == -1


def test_synthetic_inputs_are_deterministic(pkg):
    a = pkg.synth.make_pair('3dmatch', 64, 50, seed=1)
    b = pkg.synth.make_pair('3dmatch', 64, 50, seed=1)
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    assert a[0].shape == (1, 64, 3) and a[1].shape == (1, 50, 3) and a[2].shape == (1, 64, 1024)
    pts = torch.tensor([[0., 0, 0], [1, 0, 0], [-1, 0, 0], [0, 2, 0], [0, 0, 3], [5, 5, 5]])
    assert not pkg.synth.tie_free(pts)  # two neighbours at exactly distance 1


@pytest.mark.skipif(not reference_available(), reason='/root/reference not present (GPU box)')
def test_install_into_reference_rebinds_and_restores(pkg):
    """Zero-edit drop-in: the reference's own import paths hand out this package's classes after install."""
    import importlib
    GT, GSE = load_reference()
    done = pkg.install_into_reference()
    try:
        g = importlib.import_module('geotransformer.modules.geotransformer')
        t = importlib.import_module('geotransformer.modules.transformer')
        assert g.GeometricTransformer is pkg.GeometricTransformer
        assert g.GeometricStructureEmbedding is pkg.GeometricStructureEmbedding
        assert t.RPEConditionalTransformer is pkg.RPEConditionalTransformer
        assert 'geotransformer.modules.geotransformer.GeometricTransformer' in done
        # what experiments/*/model.py:35-45 does, with its own import statement
        ns = {}
        exec('from geotransformer.modules.geotransformer import GeometricTransformer', ns)
        m = ns['GeometricTransformer'](1024, 256, 256, 4, ['self', 'cross'], 0.2, 15, 3, reduction_a='max')
        assert isinstance(m, pkg.GeometricTransformer)
    finally:
        pkg.uninstall_from_reference()
    g = importlib.import_module('geotransformer.modules.geotransformer.geotransformer')
    assert g.GeometricTransformer is GT and g.GeometricStructureEmbedding is GSE


def test_embedding_kernel_switch(pkg, monkeypatch):
    """GEOTX_EMBED selects the embedding kernel at construction; the tabulated kernel only takes the shapes it was
    built for (every other request goes to the tcgen05 / fp32 kernels, never to a CPU path)."""
    import torch
    m = pkg.GeometricStructureEmbedding(256, 0.2, 15, 3)
    assert m.embed_kernel == 'cheb'  # K2c, the low-rank contraction, is the default
    m.embed_kernel = 'table'
    assert m._use_table(torch.float16, None)
    assert not m._use_table(torch.float32, None) and not m._use_table(torch.float16, object())
    monkeypatch.setenv('GEOTX_EMBED', 'table')
    assert not pkg.GeometricStructureEmbedding(256, 0.2, 15, 2)._use_table(torch.float16, None)
    assert not pkg.GeometricStructureEmbedding(256, 0.2, 15, 3, reduction_a='mean')._use_table(torch.float16, None)
    assert not pkg.GeometricStructureEmbedding(64, 0.2, 15, 3)._use_table(torch.float16, None)
    assert pkg.GeometricStructureEmbedding(128, 4.8, 15, 3)._use_table(torch.float16, None)
    monkeypatch.setenv('GEOTX_EMBED', 'tensor')
    assert pkg.GeometricStructureEmbedding(256, 0.2, 15, 3).embed_kernel == 'tensor'
    monkeypatch.setenv('GEOTX_EMBED', 'cpu')
    import pytest
    with pytest.raises(ValueError):
        pkg.GeometricStructureEmbedding(256, 0.2, 15, 3)


def test_joint_view_only_inside_one_storage(pkg):
    """The fused schedule views two cloud slices as one token buffer only when they are adjacent slices of ONE storage;
    two separate allocations that merely sit back to back in memory must be concatenated instead."""
    import torch
    jv = pkg.modules._joint_view
    buf = torch.arange(40.).view(1, 10, 4)
    v = jv(buf[:, :6], buf[:, 6:])
    assert v is not None and tuple(v.shape) == (1, 10, 4) and v.data_ptr() == buf.data_ptr() and torch.equal(v, buf)
    assert jv(buf[:, :5], buf[:, 6:]) is None            # a gap
    assert jv(buf[:, :6].clone(), buf[:, 6:].clone()) is None  # separate storages, wherever the allocator put them
    assert jv(torch.zeros(2, 3, 4)[:, :2], torch.zeros(2, 3, 4)[:, 2:]) is None  # B > 1 is not one token buffer
    big = torch.zeros(1 << 18)
    a, b = big[:1024].view(1, 4, 256), big[1024:2048].view(1, 4, 256)
    assert jv(a, b) is not None
    # adjacent in address space but different storages (emulated with two views created through from-buffer copies)
    import numpy as np
    raw = np.zeros(2048, dtype=np.float32)
    a = torch.from_numpy(raw[:1024]).view(1, 4, 256)
    b = torch.from_numpy(raw[1024:]).view(1, 4, 256)
    assert a.data_ptr() + a.numel() * 4 == b.data_ptr() and jv(a, b) is None


def test_attention_operand_switch_and_projection_fallback(pkg, monkeypatch):
    """GEOTX_ATT selects the attention core's operand form (half q | k | v written by the projection kernel, or fp32);
    `_linear(..., half_cols=)` on a shape / device K13 does not take falls back to F.linear and rounds the leading
    columns exactly as the kernel's epilogue does (round to nearest, SATURATING: a stray large activation stays finite)."""
    import torch
    m = pkg.modules
    assert m._attention_f16()
    monkeypatch.setenv('GEOTX_ATT', 'f32')
    assert not m._attention_f16()
    monkeypatch.setenv('GEOTX_ATT', 'int8')
    with pytest.raises(ValueError):
        m._attention_f16()
    g = torch.Generator().manual_seed(3)
    x, w, b = torch.randn(5, 64, generator=g), torch.randn(192, 64, generator=g), torch.randn(192, generator=g)
    b[0], b[1] = 1e6, -1e6
    full = torch.nn.functional.linear(x, w, b)
    yh, rest = m._linear(x, torch.nn.Module(), 'w', w, b, half_cols=128)  # CPU tensors: the F.linear branch
    assert yh.dtype == torch.float16 and tuple(yh.shape) == (5, 128) and torch.isfinite(yh).all()
    assert float(yh[0, 0]) == 65504.0 and float(yh[0, 1]) == -65504.0
    assert torch.equal(yh[:, 2:], full[:, 2:128].to(torch.float16)) and torch.equal(rest, full[:, 128:])
    yh, rest = m._linear(x, torch.nn.Module(), 'w', w, b, half_cols=192)
    assert rest is None and tuple(yh.shape) == (5, 192)
    sched = pkg.RPEConditionalTransformer(['self', 'cross'], 64, 4)
    assert sched.throughput is False  # one forward at a time unless a graph in flight asks for the other schedule
