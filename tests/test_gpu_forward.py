"""-m gpu: GeometricTransformer.forward end to end through the drop-in module (C ABI underneath) against golden
outputs written by the reference itself (tests/golden/make_golden.py) and against the oracle."""
import pytest
import torch

from gpu_common import DEV, TOL_BF16, TOL_FP32, build_model, rel, sd_from_npz

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('tag', ['tiny_max', 'tiny_mean'])
def test_tiny_fp32_path_vs_reference_golden(pkg, golden, tag):
    """hidden 32, 2 heads (c=16), n=40 / m=33: fp32 embedding kernel + TF32 attention."""
    z = golden(f'{tag}.npz')
    cfg = eval(str(z['cfg']))
    model = build_model(pkg, cfg, sd_from_npz(z))
    with torch.no_grad():
        o0, o1 = model(*[torch.from_numpy(z[k]).to(DEV) for k in ('ref_points', 'src_points', 'ref_feats', 'src_feats')])
    e0, e1 = rel(o0, z['out_ref']), rel(o1, z['out_src'])
    print(f'{tag}: {e0:.3e} {e1:.3e}')
    assert e0 <= 2e-3 and e1 <= 2e-3  # TF32 attention contractions bound this path


@pytest.mark.parametrize('name', ['3dmatch_n192', 'kitti_n256', 'modelnet_n128'])
def test_shipped_configs_vs_reference_golden(pkg, oracle, golden, name):
    z = golden(f'{name}.npz')
    cfg = eval(str(z['cfg']))
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    rp, sp, rf, sf, _ = pkg.synth.make_tie_free_pair(str(z['shape']), int(z['n']), int(z['m']), seed=int(z['seed']))
    model = build_model(pkg, cfg, sd)
    with torch.no_grad():
        o0, o1 = model(rp.to(DEV), sp.to(DEV), rf.to(DEV), sf.to(DEV))
    e0, e1 = rel(o0, z['out_ref']), rel(o1, z['out_src'])
    print(f'{name}: {e0:.3e} {e1:.3e}')
    assert e0 <= TOL_BF16 and e1 <= TOL_BF16


@pytest.mark.parametrize('name', ['modelnet_n717_out', '3dmatch_n1024_out'])
def test_baseline_sizes_vs_reference_golden(pkg, oracle, golden, name):
    """BASELINE.json configs[0] (ModelNet 717) and configs[1] (3DMatch 1024): full outputs of the reference."""
    z = golden(f'{name}.npz')
    cfg = eval(str(z['cfg']))
    n = int(z['n'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    rp, sp, rf, sf, _ = pkg.synth.make_tie_free_pair(str(z['shape']), n, n, seed=int(z['seed']))
    model = build_model(pkg, cfg, sd)
    with torch.no_grad():
        o0, o1 = model(rp.to(DEV), sp.to(DEV), rf.to(DEV), sf.to(DEV))
    e0, e1 = rel(o0, z['out_ref']), rel(o1, z['out_src'])
    print(f'{name}: {e0:.3e} {e1:.3e}')
    assert e0 <= TOL_BF16 and e1 <= TOL_BF16


def test_masks_batch_and_fp32_precision_vs_oracle(pkg, oracle):
    """B = 2, n != m, key masks on both clouds (rpe_transformer.py:63-64), precision='fp32' override."""
    cfg = dict(pkg.synth.CONFIGS['kitti'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    g = torch.Generator().manual_seed(9)
    rp, sp = torch.rand(2, 90, 3, generator=g) * 40, torch.rand(2, 70, 3, generator=g) * 40
    rf, sf = torch.randn(2, 90, cfg['input_dim'], generator=g), torch.randn(2, 70, cfg['input_dim'], generator=g)
    rm, sm = torch.rand(2, 90, generator=g) < 0.2, torch.rand(2, 70, generator=g) < 0.2
    e0, e1 = oracle.geometric_transformer(sd, cfg, rp, sp, rf, sf, rm, sm)
    for precision, tol in (('auto', TOL_BF16), ('fp32', 3e-3)):
        model = build_model(pkg, cfg, sd, precision)
        with torch.no_grad():
            o0, o1 = model(rp.to(DEV), sp.to(DEV), rf.to(DEV), sf.to(DEV), rm.to(DEV), sm.to(DEV))
        r0, r1 = rel(o0, e0), rel(o1, e1)
        print(f'masks {precision}: {r0:.3e} {r1:.3e}')
        assert r0 <= tol and r1 <= tol


def test_forward_only_guard_and_cpu_rejection(pkg, oracle):
    cfg = dict(pkg.synth.CONFIGS['modelnet'])
    model = pkg.GeometricTransformer(**cfg).to(DEV)
    rp, sp, rf, sf = pkg.synth.make_pair('modelnet', 32, 32)
    with pytest.raises(RuntimeError, match='forward-only'):
        model(rp.to(DEV), sp.to(DEV), rf.to(DEV), sf.to(DEV))
    with torch.no_grad(), pytest.raises(RuntimeError, match='CUDA tensor'):
        pkg.GeometricTransformer(**cfg)(rp, sp, rf, sf)


@pytest.mark.parametrize('slab_rows', [1, 7, 64])
def test_l2_slab_interleave_is_bit_identical(pkg, oracle, slab_rows):
    """The first 'self' block's scores computed slab by slab right behind the embedding kernel (StreamedEmbedding)
    must equal the one-launch schedule bit for bit: same kernels, same per-row arithmetic, only the launch split."""
    cfg = dict(pkg.synth.CONFIGS['3dmatch'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    n, m = 200, 131
    rp, sp, rf, sf = pkg.synth.make_pair('3dmatch', n, m, seed=5)
    model = build_model(pkg, cfg, sd)
    args = (rp.to(DEV), sp.to(DEV), rf.to(DEV), sf.to(DEV))
    model.fuse_scores = False  # compare the two schedules that run the same two kernels
    with torch.no_grad():
        model.l2_slab_bytes = 0
        a0, a1 = [t.clone() for t in model(*args)]
        model.l2_slab_bytes = slab_rows * n * cfg['hidden_dim'] * 2  # slab_rows rows of the larger cloud
        b0, b1 = model(*args)
    assert torch.equal(a0, b0) and torch.equal(a1, b1)
    assert pkg._abi.watchdog() is None


def test_pipelined_host_to_host_serving_matches_direct_forward(pkg, oracle):
    """PipelinedGeometricTransformer: 5 distinct host pairs through 2 graph instances with overlapped copies come back
    in order and bit-identical to model(...) on the same inputs."""
    cfg = dict(pkg.synth.CONFIGS['3dmatch'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    model = build_model(pkg, cfg, sd)
    n, m = 160, 96
    host = [tuple(t.pin_memory() for t in pkg.synth.make_pair('3dmatch', n, m, seed=20 + p)) for p in range(5)]
    with torch.no_grad():
        want = [tuple(o.cpu() for o in model(*[t.to(DEV) for t in pr])) for pr in host]
    pipe = pkg.PipelinedGeometricTransformer(model, *[t.to(DEV) for t in host[0]], depth=2)
    got = [tuple(o.clone() for o in out) for out in pipe.map(host)]
    assert len(got) == 5
    for w, g in zip(want, got):
        assert torch.equal(w[0], g[0]) and torch.equal(w[1], g[1])
    # a second pass reuses the slots
    again = [tuple(o.clone() for o in out) for out in pipe.map(host[::-1])]
    for w, g in zip(want[::-1], again):
        assert torch.equal(w[0], g[0]) and torch.equal(w[1], g[1])


def test_first_block_scores_fused_into_the_embedding_kernel(pkg, oracle):
    """geotx_gse_embed_scores: the first 'self' block's scores_p from the embedding kernel's accumulators must agree with
    the separate scores kernel (which reads the stored, 16-bit-rounded E) and both with the oracle."""
    for shape, n, m in (('3dmatch', 200, 131), ('kitti', 256, 256)):
        cfg = dict(pkg.synth.CONFIGS[shape])
        sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
        rp, sp, rf, sf = pkg.synth.make_pair(shape, n, m, seed=7)
        model = build_model(pkg, cfg, sd)
        args = (rp.to(DEV), sp.to(DEV), rf.to(DEV), sf.to(DEV))
        exp0, exp1 = oracle.geometric_transformer(sd, cfg, rp, sp, rf, sf)[:2]
        with torch.no_grad():
            model.fuse_scores = True
            f0, f1 = [t.clone() for t in model(*args)]
            model.fuse_scores = False
            u0, u1 = model(*args)
        assert rel(f0, u0) <= 2e-3 and rel(f1, u1) <= 2e-3
        assert rel(f0, exp0) <= TOL_BF16 and rel(f1, exp1) <= TOL_BF16
        print(f'{shape}: fused vs separate {rel(f0, u0):.2e}, fused vs oracle {rel(f0, exp0):.2e}')
    assert pkg._abi.watchdog() is None


def test_regenerated_embedding_equals_stored(pkg, oracle):
    """E-storage modes (clouds whose n^2 C bytes do not fit, the N = 16384 sweep point): 'regenerate'
    rebuilds E in row slabs for every 'self' block and never stores it.  Same kernels on the same values, so the
    outputs must be bit-identical to the stored mode -- eagerly and through the CUDA-graph wrapper."""
    cfg = dict(pkg.synth.CONFIGS['3dmatch'])
    rp, sp, rf, sf, _ = pkg.synth.make_tie_free_pair('3dmatch', 640, 512, seed=pkg.synth.SEED)
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    model = build_model(pkg, cfg, sd)
    args = [t.to(DEV) for t in (rp, sp, rf, sf)]
    with torch.no_grad():
        model.e_mode = 'store'
        want = [t.clone() for t in model(*args)]
        assert model.last_e_mode == 'store'
        model.e_mode = 'regenerate'
        model.regenerate_slab_bytes = 100 * 640 * 256 * 2  # 100 query rows per slab: 7 / 6 slabs, ragged tails
        got = [t.clone() for t in model(*args)]
        assert model.last_e_mode == 'regenerate'
        graphed = pkg.GraphedGeometricTransformer(model, *args)
        replay = [t.clone() for t in graphed(*args)]
        model.e_mode = 'auto'
        model(*args)
        assert model.last_e_mode == 'store'  # 0.3 GB fits
    for a, b, c in zip(want, got, replay):
        assert torch.equal(a, b) and torch.equal(a, c)
