"""The three embedding kernels side by side (K2c 'cheb' = default, K2t 'table', K2 'tensor'): time for the whole cloud
(CUDA events, back-to-back launches over an output larger than L2), GB/s of E written against MEASURED_PEAKS.json, and
the error of a row slab against the fp32 CUDA-core kernel.  usage: k2_embed_bench.py [once <kernel> <n> <C>]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
dev = 'cuda:0'
PEAK = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']


def setup(n, C, kernel):
    torch.manual_seed(1)
    sigma_d = 0.2 if C == 256 else 4.8
    m = pkg.GeometricStructureEmbedding(C, sigma_d, 15, 3).to(dev).eval()
    m.embed_kernel = kernel
    pts = pkg.synth.make_pair('3dmatch' if C == 256 else 'kitti', n)[0].to(dev)
    with torch.no_grad():
        geom = m.geometry(pts)
    return m, geom


def geom_pts(geom):
    """(n, 3) points back out of the packed geometry (float4 x, y, z, |p|^2)."""
    return geom.buf[:geom.n * 16].view(torch.float32).view(geom.n, 4)[:, :3].double()


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


if len(sys.argv) > 1 and sys.argv[1] == 'once':  # for ncu: three launches of one kernel
    kernel, n, C = sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
    m, geom = setup(n, C, kernel)
    if os.environ.get('K2_SCORES') == '1':  # K3r: scores straight from the geometry
        qt = torch.randn(1, 4, n, C, device=dev)
        sc = torch.empty(1, 4, n, n, device=dev)
        uu = torch.randn(1, 4, n, 80, device=dev) if os.environ.get('K2_SPLIT', '1') == '1' else None
        with torch.no_grad():
            for _ in range(3):
                m.scores(geom, qt, out=sc, u=uu)
        torch.cuda.synchronize()
        print('ok', float(sc[0, 0, :8].abs().max()))
        sys.exit(0)
    out = torch.empty(1, n, n, C, dtype=torch.float16, device=dev)
    with torch.no_grad():
        for _ in range(3):
            m.embed(geom, torch.float16, out=out)
    torch.cuda.synchronize()
    print('ok', float(out[0, :8].float().abs().max()))
    sys.exit(0)

res = {}
QUICK = os.environ.get('K2_QUICK') == '1'  # variant libraries (GEOTX_LIB): the low-rank kernel only, three sizes
SIZES = ((1024, 256), (4096, 256), (2048, 128)) if QUICK else ((512, 256), (1024, 256), (2048, 256), (4096, 256), (717, 256), (2048, 128))
for n, C in SIZES:
    row = {}
    out = torch.empty(1, n, n, C, dtype=torch.float16, device=dev)
    exact = None
    for kernel in (('cheb',) if QUICK else ('cheb', 'table', 'tensor')):
        m, geom = setup(n, C, kernel)
        with torch.no_grad():
            if exact is None:
                w = [t.detach() for t in (m.proj_d.weight, m.proj_d.bias, m.proj_a.weight, m.proj_a.bias)]
                exact = pkg.ops.gse_embed_fp32(geom, m.embedding.div_term, *w, m.sigma_d, m.factor_a, False,
                                               torch.float32, 0, 64)
            got = m.embed(geom, torch.float16, 0, 64).float()
            ms = timed(lambda: m.embed(geom, torch.float16, out=out))
        gbs = n * n * C * 2 / ms / 1e6
        row[kernel] = {'ms': ms, 'e_write_gbs': gbs, 'frac_of_measured_hbm_peak': gbs / PEAK,
                       'rel_err_vs_fp32_kernel': float((got - exact).norm() / exact.norm()),
                       'max_abs_err': float((got - exact).abs().max())}
        if kernel == 'cheb':
            # K3r: the same expansion + the folded-query scores, E never stored (what the forward runs by default)
            H = 4
            qt = torch.randn(1, H, n, C, device=dev)
            sc = torch.empty(1, H, n, n, device=dev)
            with torch.no_grad():
                ms_s = timed(lambda: m.scores(geom, qt, out=sc))
            kd = 16 * (1, 2, 3, 4, 6, 8)[pkg.ops.gse_cheb_max_d_set(float(torch.cdist(geom_pts(geom), geom_pts(geom)).max()) / m.sigma_d)] \
                if n <= 4096 else 48
            flops = n * n * (2.0 * C * (3 * 16 + kd) + 2.0 * C * 8 * 2)
            uu = torch.randn(1, H, n, 80, device=dev)
            with torch.no_grad():
                ms_u = timed(lambda: m.scores(geom, qt, out=sc, u=uu))
            fl_u = n * n * (2.0 * C * (3 * 16) + 2.0 * C * 8 * 2 + 2.0 * kd * 8 * 2)
            row['cheb_scores_K3r_split'] = {'ms': ms_u, 'executed_tflops': fl_u / ms_u / 1e9,
                                            'clk_per_pair_per_sm_at_1965MHz': ms_u * 1e-3 * 1.965e9 * 148 / (n * n)}
            row['cheb_scores_K3r'] = {'ms': ms_s, 'executed_tflops': flops / ms_s / 1e9, 'kd': kd,
                                      'hmma_per_pair': ((3 + kd // 16) * 4 + 4) * (C // 32) / 16.0,
                                      'clk_per_pair_per_sm_at_1965MHz': ms_s * 1e-3 * 1.965e9 * 148 / (n * n)}
            m.check_overflow()
            row['cheb_sets'] = [(r['range'], r['coefficients'], r['max_error'], r['max_abs_f']) for r in m.cheb_report()]
    res[f'n{n}_C{C}'] = row
    print(json.dumps({f'n{n}_C{C}': row}), flush=True)
os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, 'gpurun_out', os.environ.get('K2_OUT', 'k2_embed_bench.json')), 'w'), indent=1)
