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


@pytest.mark.parametrize('n,m', [(4, 7), (5, 4), (129, 4)])
def test_smallest_clouds_default_path_vs_oracle(pkg, oracle, n, m):
    """n = k + 1 (every other point is a neighbour: geotransformer.py:42 takes the k+1 nearest and drops the first) and
    clouds of a handful of points next to one that crosses a 128-key tile boundary, B = 1, DEFAULT path (K3r scores,
    own dense kernels): almost every lane of every tile is padding."""
    cfg = dict(pkg.synth.CONFIGS['3dmatch'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    g = torch.Generator().manual_seed(n * 31 + m)
    rp, sp = torch.rand(1, n, 3, generator=g) * 2, torch.rand(1, m, 3, generator=g) * 2
    rf, sf = torch.randn(1, n, cfg['input_dim'], generator=g), torch.randn(1, m, cfg['input_dim'], generator=g)
    e0, e1 = oracle.geometric_transformer(sd, cfg, rp, sp, rf, sf)
    model = build_model(pkg, cfg, sd)
    with torch.no_grad():
        o0, o1 = model(rp.to(DEV), sp.to(DEV), rf.to(DEV), sf.to(DEV))
    assert model.last_e_mode == 'scores'
    assert tuple(o0.shape) == (1, n, cfg['output_dim']) and tuple(o1.shape) == (1, m, cfg['output_dim'])
    r0, r1 = rel(o0, e0), rel(o1, e1)
    print(f'n = {n}, m = {m}: {r0:.3e} {r1:.3e}')
    assert r0 <= 1e-3 and r1 <= 1e-3


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
    model.e_mode = 'store'     # (the default, E-free 'scores' mode has no E to interleave)
    with torch.no_grad():
        model.l2_slab_bytes = 0
        a0, a1 = [t.clone() for t in model(*args)]
        model.l2_slab_bytes = slab_rows * n * cfg['hidden_dim'] * 2  # slab_rows rows of the larger cloud
        b0, b1 = model(*args)
    assert torch.equal(a0, b0) and torch.equal(a1, b1)
    assert pkg._abi.watchdog() is None


def test_pipelined_host_to_host_serving_matches_direct_forward(pkg, oracle):
    """PipelinedGeometricTransformer: 5 distinct host pairs through 2 graph instances with overlapped copies come back
    in order and bit-identical to model(...) on the same inputs under the same schedule (the pipeline captures its graphs
    with transformer.throughput = True: attention keys are not split over clusters, which changes the summation order)."""
    cfg = dict(pkg.synth.CONFIGS['3dmatch'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    model = build_model(pkg, cfg, sd)
    n, m = 160, 96
    host = [tuple(t.pin_memory() for t in pkg.synth.make_pair('3dmatch', n, m, seed=20 + p)) for p in range(5)]
    with torch.no_grad():
        latency = [tuple(o.cpu() for o in model(*[t.to(DEV) for t in pr])) for pr in host]
        model.transformer.throughput = True
        want = [tuple(o.cpu() for o in model(*[t.to(DEV) for t in pr])) for pr in host]
        model.transformer.throughput = False
    pipe = pkg.PipelinedGeometricTransformer(model, *[t.to(DEV) for t in host[0]], depth=2)
    assert model.transformer.throughput is False  # the capture restores the caller's schedule
    got = [tuple(o.clone() for o in out) for out in pipe.map(host)]
    assert len(got) == 5
    # the two schedules round the softmax weights to half against different running maxima: close, not equal
    assert all(rel(a[0], b[0]) < 1e-3 and rel(a[1], b[1]) < 1e-3 for a, b in zip(latency, want))
    for w, g in zip(want, got):
        assert torch.equal(w[0], g[0]) and torch.equal(w[1], g[1])
    # a second pass reuses the slots
    again = [tuple(o.clone() for o in out) for out in pipe.map(host[::-1])]
    for w, g in zip(want[::-1], again):
        assert torch.equal(w[0], g[0]) and torch.equal(w[1], g[1])


def test_graphed_forward_follows_weight_updates(pkg, oracle):
    """A captured graph bakes in derived weight copies (fused q|k|v|folded-query projections, embedding coefficients).
    After load_state_dict the next call must serve the NEW weights everywhere (it recaptures), for the single graph and
    for every slot of the host-to-host pipeline -- never a mix of old and new."""
    cfg = dict(pkg.synth.CONFIGS['3dmatch'])
    sd_a = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    sd_b = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'], seed=99)
    n, m = 160, 96
    pair = [t.to(DEV) for t in pkg.synth.make_pair('3dmatch', n, m, seed=31)]
    model = build_model(pkg, cfg, sd_a)
    fast = pkg.GraphedGeometricTransformer(model, *pair)
    host = [tuple(t.pin_memory() for t in pkg.synth.make_pair('3dmatch', n, m, seed=40 + p)) for p in range(4)]
    pipe = pkg.PipelinedGeometricTransformer(model, *pair, depth=2)
    with torch.no_grad():
        want_a = [o.clone() for o in model(*pair)]
        got_a = [o.clone() for o in fast(*pair)]
        assert torch.equal(want_a[0], got_a[0]) and torch.equal(want_a[1], got_a[1]) and fast.captures == 1
        assert not fast.stale()
        model.load_state_dict(sd_b, strict=True)
        assert fast.stale()
        want_b = [o.clone() for o in model(*pair)]
        got_b = [o.clone() for o in fast(*pair)]
        assert fast.captures == 2 and not fast.stale()
        assert torch.equal(want_b[0], got_b[0]) and torch.equal(want_b[1], got_b[1])
        assert rel(got_b[0], want_a[0]) > 0.1  # the two weight sets really differ
        # only one set of weights changed in place (a LayerNorm gain, read directly by the graph): still detected
        model.transformer.layers[0].attention.norm.weight.mul_(1.5)
        want_c = [o.clone() for o in model(*pair)]
        got_c = fast(*pair)
        assert fast.captures == 3 and torch.equal(want_c[0], got_c[0])
        model.transformer.throughput = True  # the pipeline's schedule
        want = [tuple(o.cpu() for o in model(*[t.to(DEV) for t in pr])) for pr in host]
        model.transformer.throughput = False
    got = [tuple(o.clone() for o in out) for out in pipe.map(host)]
    assert all(g.captures == 2 for g, _, _ in pipe.slots)
    for w, g in zip(want, got):
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
        direct = [t.clone() for t in model(*args)]
        assert model.last_e_mode == 'scores'  # the default: K3r, E is never formed in memory
        model.embedding.embed_kernel = 'table'
        model(*args)
        assert model.last_e_mode == 'store'  # without the low-rank kernel: 0.3 GB of E fits
    for a, b, c in zip(want, got, replay):
        assert torch.equal(a, b) and torch.equal(a, c)
    # K3r contracts the same half-rounded E with the same hi + lo query, only in a different summation order: the first
    # block's scores agree to 3e-7, and the six blocks amplify that ~5x each (tools/debug_modes.py); both modes sit
    # equally close to the oracle (4.6e-4)
    for a, d in zip(want, direct):
        assert rel(d, a) <= 1e-3


@pytest.mark.parametrize('name', ['kitti_n2048', '3dmatch_n2048', '3dmatch_n4096', '3dmatch_n1024_m921'])
@pytest.mark.parametrize('e_mode', ['scores', 'store'])
def test_large_sizes_vs_reference_golden_rows(pkg, oracle, golden, name, e_mode):
    """BASELINE.json configs[2] (KITTI 2048), the sweep points 2048 / 4096 (north_star's target size) and M != N at
    1024 / 921, DEFAULT path (low-rank embedding; E-free scores K3r, and the stored-E variant): 128 output rows per
    cloud written by the UNMODIFIED reference driven in query-row chunks (tests/golden/make_golden_large.py)."""
    z = golden(f'{name}_rows.npz')
    cfg = eval(str(z['cfg']))
    n, m = int(z['n']), int(z['m'])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    rp, sp, rf, sf, used = pkg.synth.make_tie_free_pair(str(z['shape']), n, m, seed=int(z['seed']))
    assert used == int(z['seed_used'])
    model = build_model(pkg, cfg, sd)
    assert model.embedding.embed_kernel == 'cheb'
    model.e_mode = e_mode
    with torch.no_grad():
        o0, o1 = model(rp.to(DEV), sp.to(DEV), rf.to(DEV), sf.to(DEV))
    assert model.last_e_mode == e_mode
    model.embedding.check_overflow()
    e0 = rel(o0[0].cpu()[torch.from_numpy(z['out_rows_ref'])], z['out_ref'])
    e1 = rel(o1[0].cpu()[torch.from_numpy(z['out_rows_src'])], z['out_src'])
    print(f'{name} [{e_mode}]: {e0:.3e} {e1:.3e}')
    assert e0 <= 1e-3 and e1 <= 1e-3  # north_star: <= 2e-2 for 16-bit operands; measured ~1e-4
    assert abs(float(o0.norm()) / float(z['out_ref_norm']) - 1) <= 1e-3


def test_large_size_embedding_rows_vs_reference_golden(pkg, oracle, golden):
    """Rows of E at n = 4096 and KITTI 2048 written by the reference, through the DEFAULT embedding kernel (K2c) and
    through the tabulated one (the round-1 4096 test only reached the tcgen05 kernel)."""
    for name in ('3dmatch_n4096', 'kitti_n2048'):
        z = golden(f'{name}_rows.npz')
        cfg = eval(str(z['cfg']))
        n = int(z['n'])
        sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
        rp = pkg.synth.make_tie_free_pair(str(z['shape']), n, int(z['m']), seed=int(z['seed']))[0]
        model = build_model(pkg, cfg, sd)
        gse = model.embedding
        cols = torch.from_numpy(z['E_cols'])
        with torch.no_grad():
            geom = gse.geometry(rp.to(DEV))
            for kernel, tol in (('cheb', 1e-3), ('table', 2e-3)):
                gse.embed_kernel = kernel
                got = torch.stack([gse.embed(geom, torch.float16, int(r), 1)[0, 0].float().cpu()[cols]
                                   for r in z['E_rows'].tolist()])
                err = rel(got, z['E'])
                print(f'{name} E rows through {kernel}: {err:.3e}')
                assert err <= tol
        gse.check_overflow()
