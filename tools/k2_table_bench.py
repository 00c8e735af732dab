"""Tabulated embedding (geotx_gse_embed_table, experimental) against the tcgen05 kernel and the fp32 CUDA-core kernel:
accuracy on a row slab and time for the whole cloud."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
dev = 'cuda:0'
res = {}
for n in (1024, 2048):
    torch.manual_seed(1)
    m = pkg.GeometricStructureEmbedding(256, 0.2, 15, 3).to(dev).eval()
    ref, _, _, _ = pkg.synth.make_pair('3dmatch', n)
    pts = ref.to(dev)
    with torch.no_grad():
        geom = m.geometry(pts)
        d_max = float(torch.cdist(pts[0], pts[0]).max()) / m.sigma_d
        for h, hd, dmax in ((1.0 / 8, 0.25, None), (1.0 / 8, 0.25, 48.0), (1.0 / 8, 0.5, 48.0), (1.0 / 8, 0.5, 96.0)):
            tb = pkg.ops.gse_build_tables(m.embedding.div_term, m.proj_d.weight.detach(), m.proj_d.bias.detach(),
                                          m.proj_a.weight.detach(), m.proj_a.bias.detach(), m.factor_a,
                                          d_max * 1.01 if dmax is None else dmax, h_a=h, h_d=hd)
            rows = 64
            exact = pkg.ops.gse_embed_fp32(geom, m.embedding.div_term, m.proj_d.weight.detach(), m.proj_d.bias.detach(),
                                           m.proj_a.weight.detach(), m.proj_a.bias.detach(), m.sigma_d, m.factor_a, False,
                                           torch.float32, 0, rows)
            tc = m.embed(geom, torch.float16, 0, rows).float()
            tab = pkg.ops.gse_embed_table(geom, tb, m.sigma_d, m.factor_a, 0, rows).float()
            scale = float(exact.abs().max())
            out_tc = torch.empty(1, n, n, 256, dtype=torch.float16, device=dev)
            out_tb = torch.empty(1, n, n, 256, dtype=torch.float16, device=dev)

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
            res[f'n{n}_ha{h}_hd{hd}_dmax{dmax}'] = {
                'table_rows': [tb.rows_a, tb.rows_d], 'table_bytes': int(tb.buf.numel()),
                'max_abs_E': scale,
                'table_vs_fp32_max_abs': float((tab - exact).abs().max()),
                'tcgen05_vs_fp32_max_abs': float((tc - exact).abs().max()),
                'table_vs_fp32_rms': float((tab - exact).pow(2).mean().sqrt()),
                'tcgen05_vs_fp32_rms': float((tc - exact).pow(2).mean().sqrt()),
                'table_ms': timed(lambda: pkg.ops.gse_embed_table(geom, tb, m.sigma_d, m.factor_a, out=out_tb)),
                'tcgen05_ms': timed(lambda: m.embed(geom, torch.float16, out=out_tc)),
                'overflow': int(tb.overflow.item()),
                'full_max_abs_table_vs_tcgen05': float((out_tb.float() - out_tc.float()).abs().max()),
            }
            print(json.dumps({f'n{n}_ha{h}_hd{hd}_dmax{dmax}': res[f'n{n}_ha{h}_hd{hd}_dmax{dmax}']}), flush=True)
json.dump(res, open(os.path.join(ROOT, 'gpurun_out', 'k2_table.json'), 'w'), indent=1)
