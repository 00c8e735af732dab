"""-m gpu: K2, the fused GeometricStructureEmbedding kernels (tcgen05 bf16 path and the fp32 CUDA-core path)
through the C ABI, against golden E written by the reference and against the oracle."""
import pytest
import torch

from gpu_common import DEV, TOL_BF16, TOL_FP32, rel, sd_from_npz

pytestmark = pytest.mark.gpu


def _gse(pkg, cfg, sd, precision, kernel='tensor'):
    """`kernel`: which kernel serves a 16-bit E ('tensor' = K2 tcgen05, 'table' = K2t, 'cheb' = K2c, the default of the
    module); fp32 output always comes from K2 / the CUDA-core kernel."""
    m = pkg.GeometricStructureEmbedding(cfg['hidden_dim'], cfg['sigma_d'], cfg['sigma_a'], cfg['angle_k'],
                                        reduction_a=cfg.get('reduction_a', 'max'), precision=precision)
    m.load_state_dict({k[len('embedding.'):]: v for k, v in sd.items() if k.startswith('embedding.')}, strict=True)
    m.embed_kernel = kernel
    return m.to(DEV).eval()


@pytest.mark.parametrize('tag', ['tiny_max', 'tiny_mean'])
def test_fp32_kernel_vs_reference_golden(pkg, golden, tag):
    """hidden_dim 32 (no tensor-core path): fp32 CUDA-core kernel, both reductions, <= 1e-4."""
    z = golden(f'{tag}.npz')
    cfg = eval(str(z['cfg']))
    m = _gse(pkg, cfg, sd_from_npz(z), 'auto')
    with torch.no_grad():
        for pts, key in ((z['ref_points'], 'E_ref'), (z['src_points'], 'E_src')):
            e = m(torch.from_numpy(pts).to(DEV))
            assert e.dtype == torch.float32 and tuple(e.shape) == z[key].shape
            assert rel(e, z[key]) <= TOL_FP32


@pytest.mark.parametrize('name', ['3dmatch_n192', 'kitti_n256', 'modelnet_n128'])
@pytest.mark.parametrize('precision', ['bf16', 'fp32'])
def test_shipped_configs_vs_reference_golden(pkg, oracle, golden, name, precision):
    """The three shipped configs (C = 256 / 128 / 256): 8 golden rows of E from the reference."""
    z = golden(f'{name}.npz')
    cfg = eval(str(z['cfg']))
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    pts = pkg.synth.make_tie_free_pair(str(z['shape']), int(z['n']), int(z['m']), seed=int(z['seed']))[0]
    m = _gse(pkg, cfg, sd, precision)
    rows = torch.from_numpy(z['rows'])
    with torch.no_grad():
        e = m(pts.to(DEV))
    err = rel(e[0].cpu()[rows], z['E_rows'])
    print(f'{name} {precision}: E rel err {err:.3e}')
    assert err <= (TOL_BF16 if precision == 'bf16' else TOL_FP32)
    if precision == 'bf16':
        assert err <= 8e-3, 'bf16 operands + fp32 accumulation should sit near 4e-3'


@pytest.mark.parametrize('C,reduction', [(128, 'max'), (256, 'mean'), (256, 'max')])
def test_tensor_core_kernel_vs_oracle(pkg, oracle, C, reduction):
    """Ragged n (tile tail), B = 2, k = 2, both reductions, f16 / bf16 / fp32 outputs, row slabs."""
    cfg = dict(hidden_dim=C, sigma_d=0.3, sigma_a=10, angle_k=2, reduction_a=reduction)
    sd = oracle.init_state_dict(8, 8, C, [], seed=5)
    g = torch.Generator().manual_seed(C)
    pts = torch.rand(2, 173, 3, generator=g) * 2.5
    exp = oracle.gse(pts, sd, 'embedding.', cfg['sigma_d'], cfg['sigma_a'], cfg['angle_k'], reduction)
    m = _gse(pkg, cfg, sd, 'bf16')
    with torch.no_grad():
        geom = m.geometry(pts.to(DEV))
        e32 = m.embed(geom, torch.float32)
        e16 = m.embed(geom, torch.bfloat16)
        eh = m.embed(geom, torch.float16)
        slab = m.embed(geom, torch.float32, row_begin=40, row_count=77)
    assert m.storage_dtype() == torch.float16  # default nn.Linear init is far inside the half range
    assert rel(e32, exp) <= 8e-3
    assert rel(e16, exp) <= 1e-2
    assert rel(eh, exp) <= 8e-3
    assert torch.equal(slab, e32[:, 40:117]), 'row slabs must reproduce the whole-cloud result bit for bit'
    assert torch.equal(eh, e32.to(torch.float16)), 'f16 output is the rounded fp32 output'
    assert torch.equal(e16, e32.to(torch.bfloat16)), 'bf16 output is the rounded fp32 output'


def test_full_size_properties_n4096(pkg, oracle):
    """BASELINE full size (n = 4096, C = 256) through size-independent properties on a slab of rows:
    (1) relabelling the points permutes E exactly (bit for bit); (2) sampled rows match the oracle evaluated on those
    rows only (the reference itself cannot materialise n = 4096) within the bf16 tolerance."""
    cfg = dict(pkg.synth.CONFIGS['3dmatch'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    n = 4096
    pts = pkg.synth.make_tie_free_pair('3dmatch', n, n, seed=pkg.synth.SEED, input_dim=4)[0]
    m = _gse(pkg, cfg, sd, 'bf16')
    perm = torch.randperm(n, generator=torch.Generator().manual_seed(3))
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(n)
    rows = torch.tensor([0, 1, 777, 2048, 4095])
    with torch.no_grad():
        geom = m.geometry(pts.to(DEV))
        geom_p = m.geometry(pts[:, perm].contiguous().to(DEV))
        for r in rows.tolist():
            e = m.embed(geom, torch.bfloat16, row_begin=r, row_count=1)[0, 0]             # (n, C): E[r, :, :]
            ep = m.embed(geom_p, torch.bfloat16, row_begin=int(inv[r]), row_count=1)[0, 0]  # same point, relabelled
            assert torch.equal(ep.cpu(), e.cpu()[perm]), 'E is not equivariant under point relabelling'
        # oracle on the sampled rows only (the reference cannot materialise n = 4096)
        d_idx, a_idx, _ = oracle.embedding_indices(pts, cfg['sigma_d'], cfg['sigma_a'], cfg['angle_k'])
        exp = oracle.gse_from_indices(d_idx[:, rows], a_idx[:, rows], sd, 'embedding.')
        got = torch.stack([m.embed(geom, torch.float32, row_begin=r, row_count=1)[0, 0] for r in rows.tolist()])
    assert rel(got, exp[0]) <= 8e-3


def test_embed_rejects_bad_arguments(pkg, oracle):
    sd = oracle.init_state_dict(8, 8, 96, [], seed=5)
    with pytest.raises(RuntimeError, match='128 or 256'), torch.no_grad():
        _gse(pkg, dict(hidden_dim=96, sigma_d=0.2, sigma_a=15, angle_k=3), sd, 'bf16')(
            torch.rand(1, 16, 3).to(DEV))
    m = _gse(pkg, dict(hidden_dim=96, sigma_d=0.2, sigma_a=15, angle_k=3), sd, 'auto')
    pts = torch.rand(1, 16, 3)
    with torch.no_grad():
        e = m(pts.to(DEV))  # falls to the fp32 CUDA-core kernel, any even C
    exp = oracle.gse(pts, sd, 'embedding.', 0.2, 15, 3)
    assert rel(e, exp) <= TOL_FP32
    with pytest.raises(RuntimeError, match='forward-only'):
        m(pts.to(DEV))  # grad enabled + parameters require grad


def test_storage_format_falls_back_to_bf16_when_half_could_overflow(pkg, oracle):
    """|E_c| <= ||W_d[c]||_1 + ||W_a[c]||_1 + |b|: with weights scaled far beyond any checkpoint the module must pick
    bf16 storage (same exponent range as fp32) and still match the oracle."""
    cfg = dict(hidden_dim=128, sigma_d=0.3, sigma_a=10, angle_k=3, reduction_a='max')
    sd = oracle.init_state_dict(8, 8, 128, [], seed=11)
    sd = {k: (v * 40000.0 if k.endswith('proj_d.weight') else v) for k, v in sd.items()}
    g = torch.Generator().manual_seed(3)
    pts = torch.rand(1, 150, 3, generator=g) * 2.0
    exp = oracle.gse(pts, sd, 'embedding.', cfg['sigma_d'], cfg['sigma_a'], cfg['angle_k'], 'max')
    m = _gse(pkg, cfg, sd, 'bf16')
    assert m.storage_dtype() == torch.bfloat16
    with torch.no_grad():
        e = m.embed(m.geometry(pts.to(DEV)), m.storage_dtype())
    assert torch.isfinite(e).all()
    assert rel(e, exp) <= 1e-2


@pytest.mark.parametrize('C,H,n,reduction', [(256, 4, 173, 'max'), (128, 4, 256, 'max'), (256, 2, 128, 'mean'), (256, 4, 384, 'max')])
def test_fused_first_block_scores(pkg, oracle, C, H, n, reduction):
    """geotx_gse_embed_scores: scores_p[b,h,i,j] = qt[b,h,i,:] . E[b,i,j,:] out of the embedding kernel (ragged n so
    that tiles straddle query rows, B = 2, row slab), against the fp32 output of the same launch."""
    cfg = dict(hidden_dim=C, sigma_d=0.3, sigma_a=10, angle_k=3, reduction_a=reduction)
    sd = oracle.init_state_dict(8, 8, C, [], seed=5)
    g = torch.Generator().manual_seed(n)
    pts = torch.rand(2, n, 3, generator=g) * 2.5
    qt = torch.randn(2, H, n, C, generator=g).to(DEV)
    m = _gse(pkg, cfg, sd, 'bf16')
    with torch.no_grad():
        geom = m.geometry(pts.to(DEV))
        e32, sc = m.embed(geom, torch.float32, qt=qt)
        plain = m.embed(geom, torch.float32)
        assert torch.equal(e32, plain), 'E must not depend on whether the scores are fused'
        want = torch.einsum('bhic,bijc->bhij', qt.double(), e32.double()).float()
        assert rel(sc, want) <= 1e-5
        # slab of rows [30, 100): only those rows of scores are written
        part = torch.full_like(sc, -7.0)
        pkg.ops.gse_embed(geom, m.embedding.div_term, m._packed_weights(), C, m.sigma_d, m.factor_a,
                          reduction == 'mean', torch.float16, 30, 70, qt=qt, scores_out=part)
        if n % 128 == 0:  # fused in the kernel: independent of the storage format of E
            assert torch.equal(part[:, :, 30:100], sc[:, :, 30:100])
        else:             # two-kernel fallback: the scores kernel reads the f16-rounded E
            assert rel(part[:, :, 30:100], want[:, :, 30:100]) <= 2e-3
        assert float((part[:, :, :30] + 7.0).abs().max()) == 0.0 and float((part[:, :, 100:] + 7.0).abs().max()) == 0.0
    assert pkg._abi.watchdog() is None


# ---- K2t: the tabulated embedding (GEOTX_EMBED=table; superseded by K2c) -----------------------------------------------
TOL_TABLE = 2e-3  # Frobenius-relative: half tables + half arithmetic; measured ~5e-4 (the bf16 contraction: ~4e-3)


@pytest.mark.parametrize('name', ['3dmatch_n192', 'kitti_n256', 'modelnet_n128'])
def test_table_kernel_vs_reference_golden(pkg, oracle, golden, name):
    """The three shipped configs (C = 256 / 128 / 256): golden rows of E written by the reference, through the
    tabulated kernel."""
    z = golden(f'{name}.npz')
    cfg = eval(str(z['cfg']))
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    pts = pkg.synth.make_tie_free_pair(str(z['shape']), int(z['n']), int(z['m']), seed=int(z['seed']))[0]
    m = _gse(pkg, cfg, sd, 'auto')
    m.embed_kernel = 'table'
    rows = torch.from_numpy(z['rows'])
    with torch.no_grad():
        e = m(pts.to(DEV), out_dtype=torch.float16)
    assert e.dtype == torch.float16 and int(m._tables().overflow.item()) == 0
    err = rel(e[0].float().cpu()[rows], z['E_rows'])
    print(f'{name} table: E rel err {err:.3e}')
    assert err <= TOL_TABLE


def test_table_kernel_vs_oracle_and_beyond_the_table(pkg, oracle):
    """Ragged n, B = 2, row slabs; then the same clouds with a table that ends at a tenth of their extent, so that most
    distances take the exact path: both must agree with the oracle."""
    cfg = dict(hidden_dim=256, sigma_d=0.2, sigma_a=15, angle_k=3, reduction_a='max')
    sd = oracle.init_state_dict(8, 8, 256, [], seed=6)
    pts = torch.rand(2, 173, 3, generator=torch.Generator().manual_seed(3)) * 2.5
    exp = oracle.gse(pts, sd, 'embedding.', cfg['sigma_d'], cfg['sigma_a'], cfg['angle_k'], 'max')
    m = _gse(pkg, cfg, sd, 'auto')
    m.embed_kernel = 'table'
    with torch.no_grad():
        geom = m.geometry(pts.to(DEV))
        e = m.embed(geom, torch.float16)
        slab = m.embed(geom, torch.float16, row_begin=40, row_count=77)
        assert m._tables().rows_d > 150
        m.table_d_index_max = 2.0  # 0.4 m of a 2.5 m cloud
        short = m.embed(geom, torch.float16)
        assert m._tables().rows_d < 20
        m.embed_kernel = 'tensor'
        tensor = m.embed(geom, torch.float16)
    assert rel(e.float(), exp) <= TOL_TABLE and rel(short.float(), exp) <= TOL_TABLE
    assert torch.equal(slab, e[:, 40:117])
    assert rel(e.float(), exp) < rel(tensor.float(), exp), 'the tabulated kernel should sit closer to fp32 than bf16 MMA'
    assert int(m._tables().overflow.item()) == 0


def test_forward_with_table_embedding(pkg, oracle, golden):
    """GeometricTransformer.forward with the tabulated embedding against the reference's golden outputs (3DMatch 1024,
    BASELINE.json configs[1]); it must not be further from the reference than the tensor-core embedding."""
    z = golden('3dmatch_n1024_out.npz')
    cfg = eval(str(z['cfg']))
    n = int(z['n'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    rp, sp, rf, sf, _ = pkg.synth.make_tie_free_pair(str(z['shape']), n, n, seed=int(z['seed']))
    model = pkg.GeometricTransformer(**cfg).to(DEV).eval()
    model.load_state_dict(sd, strict=True)
    errs = {}
    for kernel in ('tensor', 'table'):
        model.embedding.embed_kernel = kernel
        with torch.no_grad():
            o0, o1 = model(rp.to(DEV), sp.to(DEV), rf.to(DEV), sf.to(DEV))
        errs[kernel] = max(rel(o0, z['out_ref']), rel(o1, z['out_src']))
    print(f'forward rel err: {errs}')
    assert errs['table'] <= TOL_BF16 and errs['table'] <= errs['tensor'] * 1.5


# ---- K2c: the low-rank (Chebyshev) contraction, the default embedding kernel ------------------------------------------
TOL_CHEB = 1e-3  # Frobenius-relative: half operands, fp32 accumulation, half storage of E; measured ~3e-4


@pytest.mark.parametrize('name', ['3dmatch_n192', 'kitti_n256', 'modelnet_n128'])
def test_cheb_kernel_vs_reference_golden(pkg, oracle, golden, name):
    """The three shipped configs (C = 256 / 128 / 256): golden rows of E written by the reference, through K2c, which
    is what GeometricTransformer runs by default."""
    z = golden(f'{name}.npz')
    cfg = eval(str(z['cfg']))
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    pts = pkg.synth.make_tie_free_pair(str(z['shape']), int(z['n']), int(z['m']), seed=int(z['seed']))[0]
    m = _gse(pkg, cfg, sd, 'auto', kernel='cheb')
    assert pkg.GeometricStructureEmbedding(256, 0.2, 15, 3).embed_kernel == 'cheb'
    rows = torch.from_numpy(z['rows'])
    with torch.no_grad():
        assert m._use_cheb(torch.float16, None)
        e = m(pts.to(DEV), out_dtype=torch.float16)
    m.check_overflow()
    rep = m.cheb_report()
    print(f'{name} cheb report: {[(r["range"], r["coefficients"], "%.1e" % r["max_error"]) for r in rep]}')
    assert all(r['max_error'] <= 2e-3 * r['max_abs_f'] for r in rep)
    err = rel(e[0].float().cpu()[rows], z['E_rows'])
    print(f'{name} cheb: E rel err {err:.3e}')
    assert e.dtype == torch.float16 and err <= TOL_CHEB


@pytest.mark.parametrize('C,reduction,k', [(256, 'max', 3), (128, 'max', 3), (256, 'mean', 3), (256, 'max', 1),
                                           (128, 'mean', 2), (256, 'max', 5), (128, 'max', 7)])
def test_cheb_kernel_vs_oracle(pkg, oracle, C, reduction, k):
    """Ragged n (tile tail), B = 2, every reduction and several k, f16 / bf16 outputs, row slabs, the oracle as checker;
    K2c must sit closer to fp32 than the bf16 tcgen05 contraction."""
    cfg = dict(hidden_dim=C, sigma_d=0.3, sigma_a=10, angle_k=k, reduction_a=reduction)
    sd = oracle.init_state_dict(8, 8, C, [], seed=5)
    pts = torch.rand(2, 173, 3, generator=torch.Generator().manual_seed(C + k)) * 2.5
    exp = oracle.gse(pts, sd, 'embedding.', cfg['sigma_d'], cfg['sigma_a'], k, reduction)
    m = _gse(pkg, cfg, sd, 'auto', kernel='cheb')
    with torch.no_grad():
        geom = m.geometry(pts.to(DEV))
        assert m._use_cheb(torch.float16, None)
        eh = m.embed(geom, torch.float16)
        eb = m.embed(geom, torch.bfloat16)
        slab = m.embed(geom, torch.float16, row_begin=40, row_count=77)
        m.embed_kernel = 'tensor'
        tensor = m.embed(geom, torch.float16)
    m.check_overflow()
    err = rel(eh.float(), exp)
    print(f'C={C} {reduction} k={k}: cheb {err:.3e}, tcgen05 bf16 {rel(tensor.float(), exp):.3e}')
    assert err <= TOL_CHEB and rel(eb.float(), exp) <= 4e-3
    assert torch.equal(slab, eh[:, 40:117]), 'row slabs must reproduce the whole-cloud result bit for bit'
    assert err < rel(tensor.float(), exp)


@pytest.mark.parametrize('extent,want_set', [(1.5, 0), (6.0, 1), (11.0, 2), (17.0, 3), (28.0, 4), (40.0, 5)])
def test_cheb_distance_sets(pkg, oracle, extent, want_set):
    """The kernel picks the distance expansion from the cloud's largest distance (left behind by geotx_gse_knn): clouds
    of growing extent walk through all six prebuilt sets (ranges 13 .. 210 index units at sigma_d = 0.2)."""
    cfg = dict(hidden_dim=128, sigma_d=0.2, sigma_a=15, angle_k=3, reduction_a='max')
    sd = oracle.init_state_dict(8, 8, 128, [], seed=9)
    pts = torch.rand(1, 200, 3, generator=torch.Generator().manual_seed(int(extent * 10)))
    pts[0, 0] = 0.0
    pts[0, 1] = torch.tensor([1.0, 0.0, 0.0])  # the largest distance is >= 1 * extent and <= sqrt(3) * extent
    pts = pts * extent / 3 ** 0.5
    d_max = float(torch.cdist(pts[0].double(), pts[0].double()).max()) / cfg['sigma_d']
    lib = pkg._abi.load()
    ranges = [float(lib.geotx_gse_cheb_d_range(s)) for s in range(6)]
    assert min(s for s in range(6) if d_max <= ranges[s]) == want_set, (d_max, ranges)
    exp = oracle.gse(pts, sd, 'embedding.', cfg['sigma_d'], cfg['sigma_a'], 3, 'max')
    m = _gse(pkg, cfg, sd, 'auto', kernel='cheb')
    with torch.no_grad():
        e = m(pts.to(DEV), out_dtype=torch.float16)
    m.check_overflow()
    err = rel(e.float(), exp)
    print(f'extent {extent}: d_max {d_max:.1f} index units -> set {want_set}, err {err:.3e}')
    assert err <= TOL_CHEB


def test_cheb_overflow_and_guard(pkg, oracle):
    """(1) A cloud larger than cheb_d_index_max allows: the kernel raises its overflow flag; the first eager call of the
    shape reads it, widens the range to the kernels' maximum (210 index units) with a warning and repeats the launch --
    the caller never sees clamped values; a cloud beyond even that raises.  (2) Weights scaled x50 and x2000: the
    measured error of the expansions stays RELATIVE to the weights (a sampled table's would not), the module keeps K2c
    and still matches the oracle."""
    cfg = dict(hidden_dim=128, sigma_d=0.2, sigma_a=15, angle_k=3, reduction_a='max')
    sd = oracle.init_state_dict(8, 8, 128, [], seed=9)
    pts = torch.rand(1, 100, 3, generator=torch.Generator().manual_seed(2)) * 6.0
    exp = oracle.gse(pts, sd, 'embedding.', cfg['sigma_d'], cfg['sigma_a'], 3, 'max')
    m = _gse(pkg, cfg, sd, 'auto', kernel='cheb')
    assert m.cheb_d_index_max == 64.0
    m.cheb_d_index_max = 13.0  # set 0 only: 2.6 m of a cloud that spans up to 10 m
    with pytest.warns(UserWarning, match='cheb_d_index_max'), torch.no_grad():
        e = m(pts.to(DEV), out_dtype=torch.float16)
    assert m.cheb_d_index_max == 210.0 and rel(e.float(), exp) <= TOL_CHEB
    m.check_overflow()  # the flag was cleared before the repeated launch
    far = torch.rand(1, 90, 3, generator=torch.Generator().manual_seed(3)) * 60.0  # a new shape: its first call is checked
    with pytest.raises(RuntimeError, match='cheb_d_index_max'), torch.no_grad():
        m(far.to(DEV), out_dtype=torch.float16)  # up to 520 index units: beyond every prebuilt set
    for scale in (50.0, 2000.0):
        big = {k: (v * scale if k.endswith(('proj_a.weight', 'proj_d.weight')) else v) for k, v in sd.items()}
        exp = oracle.gse(pts, big, 'embedding.', cfg['sigma_d'], cfg['sigma_a'], 3, 'max')
        m = _gse(pkg, cfg, big, 'auto', kernel='cheb')
        with torch.no_grad():
            assert m._use_cheb(m.storage_dtype(), None)
            e = m(pts.to(DEV), out_dtype=m.storage_dtype())
        err = rel(e.float(), exp)
        print(f'weights x{scale:g}: storage {m.storage_dtype()}, cheb err {err:.3e}')
        assert err <= (TOL_CHEB if m.storage_dtype() == torch.float16 else 4e-3)


def test_cheb_full_size_properties_n4096(pkg, oracle):
    """BASELINE full size (n = 4096, C = 256) through K2c, the kernel behind every headline number: (1) relabelling the
    points permutes E exactly; (2) a row slab equals the same rows of a larger slab; (3) sampled rows match the oracle
    evaluated on those rows (the reference itself cannot materialise n = 4096)."""
    cfg = dict(pkg.synth.CONFIGS['3dmatch'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    n = 4096
    pts = pkg.synth.make_tie_free_pair('3dmatch', n, n, seed=pkg.synth.SEED, input_dim=4)[0]
    m = _gse(pkg, cfg, sd, 'auto', kernel='cheb')
    perm = torch.randperm(n, generator=torch.Generator().manual_seed(3))
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(n)
    rows = torch.tensor([0, 1, 777, 2048, 4095])
    with torch.no_grad():
        geom = m.geometry(pts.to(DEV))
        geom_p = m.geometry(pts[:, perm].contiguous().to(DEV))
        got = []
        for r in rows.tolist():
            e = m.embed(geom, torch.float16, row_begin=r, row_count=1)[0, 0]             # (n, C): E[r, :, :]
            ep = m.embed(geom_p, torch.float16, row_begin=int(inv[r]), row_count=1)[0, 0]  # same point, relabelled
            assert torch.equal(ep.cpu(), e.cpu()[perm]), 'E is not equivariant under point relabelling'
            got.append(e.float().cpu())
        big = m.embed(geom, torch.float16, row_begin=700, row_count=100)
        assert torch.equal(big[0, 77].cpu().float(), got[2])
        d_idx, a_idx, _ = oracle.embedding_indices(pts, cfg['sigma_d'], cfg['sigma_a'], cfg['angle_k'])
        exp = oracle.gse_from_indices(d_idx[:, rows], a_idx[:, rows], sd, 'embedding.')
    m.check_overflow()
    err = rel(torch.stack(got), exp[0])
    print(f'n = 4096 cheb rows: {err:.3e}')
    assert err <= TOL_CHEB


@pytest.mark.parametrize('C,H,n,reduction,k', [(256, 4, 173, 'max', 3), (128, 4, 256, 'max', 3), (256, 8, 130, 'mean', 2),
                                               (256, 2, 384, 'max', 1), (128, 1, 64, 'max', 5)])
def test_cheb_scores_without_e(pkg, oracle, C, H, n, reduction, k):
    """K3r (geotx_rpe_scores_cheb): scores_p = qt . E straight from the geometry must equal the two-kernel path
    (K2c writes the half E, K3a reads it back) up to the summation order, and the oracle's E within K2c's tolerance;
    ragged n (key tiles with a tail), B = 2, 1..8 heads, row slabs, strided qt."""
    cfg = dict(hidden_dim=C, sigma_d=0.3, sigma_a=10, angle_k=k, reduction_a=reduction)
    sd = oracle.init_state_dict(8, 8, C, [], seed=5)
    g = torch.Generator().manual_seed(n + H)
    pts = torch.rand(2, n, 3, generator=g) * 2.5
    wide = torch.randn(2, n, H * C + 32, generator=g).to(DEV)
    qt = wide[:, :, 32:].unflatten(2, (H, C)).permute(0, 2, 1, 3)  # (B,H,n,C) view of a wider projection output
    m = _gse(pkg, cfg, sd, 'auto', kernel='cheb')
    with torch.no_grad():
        geom = m.geometry(pts.to(DEV))
        E = m.embed(geom, torch.float16)
        want = pkg.ops.rpe_scores(qt, E)
        got = m.scores(geom, qt)
        part = torch.full_like(got, -7.0)
        m.scores(geom, qt, row_begin=30, row_count=20, out=part)
        # the distance split: u = G_d qt in coefficient space replaces the distance term's share of the contraction
        # (fast instances: k = 3, 'max'; every other shape ignores u and must reproduce `got` bit for bit)
        gd = m.distance_split_rows()
        assert tuple(gd.shape) == (80, C)
        u_wide = torch.zeros(2, n, H * 80 + 8, device=DEV)
        u_wide[:, :, 8:] = torch.einsum('bhnc,mc->bnhm', qt.double(), gd.double()).float().flatten(2)
        u = u_wide[:, :, 8:].unflatten(2, (H, 80)).permute(0, 2, 1, 3)  # (B,H,n,80) view of a wider buffer
        split = m.scores(geom, qt, u=u)
    m.check_overflow()
    if k == 3 and reduction == 'max':
        print(f'    distance split vs unsplit {rel(split, got):.3e}, vs oracle (below)')
        assert rel(split, got) <= 1e-3
    else:
        assert torch.equal(split, got)
    exp = torch.einsum('bhic,bijc->bhij', qt.double().cpu(),
                       oracle.gse(pts, sd, 'embedding.', 0.3, 10, k, reduction).double()).float()
    print(f'C={C} H={H} n={n}: K3r vs K2c+K3a {rel(got, want):.3e}, vs oracle {rel(got, exp):.3e}')
    assert rel(got, want) <= 1e-5
    assert rel(got, exp) <= TOL_CHEB and rel(split, exp) <= TOL_CHEB
    assert torch.equal(part[:, :, 30:50], got[:, :, 30:50])
    assert float((part[:, :, :30] + 7.0).abs().max()) == 0.0 and float((part[:, :, 50:] + 7.0).abs().max()) == 0.0


@pytest.mark.parametrize('C,sigma_d,want_set', [(256, 0.05, 3), (128, 0.05, 3), (256, 0.08, 2)])
def test_cheb_split_scores_wide_clouds(pkg, oracle, C, sigma_d, want_set):
    """The split instances of K3r on clouds wider than 64 index units: distance set 3 (64 coefficients, u of 144 columns:
    sets 1, 2, 3) against the unsplit kernel and the oracle; set 2 through the same wide preparation."""
    H, n, k = 4, 200, 3
    cfg = dict(hidden_dim=C, sigma_d=sigma_d, sigma_a=10, angle_k=k, reduction_a='max')
    sd = oracle.init_state_dict(8, 8, C, [], seed=5)
    g = torch.Generator().manual_seed(C + want_set)
    pts = torch.rand(1, n, 3, generator=g) * 2.5
    qt = torch.randn(1, H, n, C, generator=g).to(DEV)
    m = _gse(pkg, cfg, sd, 'auto', kernel='cheb')
    m.cheb_d_index_max = 93.0
    with torch.no_grad():
        geom = m.geometry(pts.to(DEV))
        got = m.scores(geom, qt)
        gd = m.distance_split_rows()
        assert tuple(gd.shape) == (144, C)
        u = torch.einsum('bhnc,mc->bhnm', qt.double(), gd.double()).float().contiguous()
        split = m.scores(geom, qt, u=u)
    m.check_overflow()
    d_max = float(torch.cdist(pts[0].double(), pts[0].double()).max()) / sigma_d
    assert pkg.ops.gse_cheb_max_d_set(d_max) == want_set, d_max
    exp = torch.einsum('bhic,bijc->bhij', qt.double().cpu(), oracle.gse(pts, sd, 'embedding.', sigma_d, 10, k, 'max').double()).float()
    print(f'C={C} set {want_set}: split vs unsplit {rel(split, got):.3e}, split vs oracle {rel(split, exp):.3e}, unsplit vs oracle {rel(got, exp):.3e}')
    assert rel(split, got) <= 1e-3 and rel(split, exp) <= TOL_CHEB and rel(got, exp) <= TOL_CHEB
