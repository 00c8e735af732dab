"""What each kernel family costs INSIDE the pipelined step: the bench's timed region (graph instances in flight, rotated
resident pairs) re-captured with one family's launches left out (outputs are then garbage: timing only).  The difference
to the full step is that family's marginal cost under the real schedule, which the serialised launch list cannot show.

    python tools/ablate_step.py [--workload 3dmatch_1024] [--steps 60]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


FAMILIES = [
    ('full step', ()),
    ('- attention (K3b)', ('geotx_attention',)),
    ('- post_attention (K4f)', ('geotx_post_attention',)),
    ('- linear (K13)', ('geotx_linear_f16',)),
    ('- K4f and K13', ('geotx_post_attention', 'geotx_linear_f16')),
    ('- K3b, K4f, K13', ('geotx_attention', 'geotx_post_attention', 'geotx_linear_f16')),
    ('- embedding scores (K3r)', ('geotx_rpe_scores_cheb',)),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', default='3dmatch_1024')
    ap.add_argument('--steps', type=int, default=60)
    ap.add_argument('--inflight', type=int, default=0)
    ap.add_argument('--full-only', action='store_true', help='time the full step only (A/B runs of environment knobs)')
    args = ap.parse_args()
    import torch
    import pkgpath
    pkg = pkgpath.load_package()
    dev = torch.device('cuda', 0)
    shape, n, m = bench.WORKLOADS[args.workload]
    cfg = dict(pkg.synth.CONFIGS[shape])
    torch.manual_seed(pkg.synth.SEED)
    model = pkg.GeometricTransformer(**cfg).to(dev).eval()
    pair_bytes = (n + m) * (cfg['input_dim'] + 3) * 4
    n_pairs = max(4, -(-140_000_000 // pair_bytes))
    host = [pkg.synth.make_pair(shape, n, m, seed=pkg.synth.SEED + p) for p in range(n_pairs)]
    resident = [tuple(t.to(dev) for t in pr) for pr in host]
    widest = max(bench.host_d_set(pkg, pr[c], cfg['sigma_d'])[1] for pr in host for c in (0, 1))
    if widest > model.embedding.cheb_d_index_max:
        lib = pkg._abi.load()
        model.embedding.cheb_d_index_max = next(float(lib.geotx_gse_cheb_d_range(s)) for s in range(6)
                                                if widest <= float(lib.geotx_gse_cheb_d_range(s)) or s == 5)
    inflight = args.inflight or bench.default_inflight(max(n, m))
    with torch.no_grad():
        model(*resident[0])  # packs the weights, builds the coefficient tables
    torch.cuda.synchronize()
    real_call = pkg._abi.call
    skip = set()

    def call(name, *a):
        if name in skip:
            return
        real_call(name, *a)
    pkg._abi.call = call
    base = None
    for tag, names in (FAMILIES[:1] if args.full_only else FAMILIES):
        skip.clear()
        skip.update(names)
        with torch.no_grad():
            fast = bench.InFlight(pkg, model, resident[0], inflight)
            best = None
            for rep in range(3):
                ms, _ = bench.timed_value(torch, fast, resident, n_pairs, args.steps, 2 * inflight)
                best = ms if best is None else min(best, ms)
            del fast
        us = best / args.steps * 1e3
        base = us if base is None else base
        print(f'{tag:32s} {us:8.1f} us per pair   {args.steps / (best * 1e-3):8.1f} pairs/s   ({us - base:+7.1f} us)', flush=True)


if __name__ == '__main__':
    main()
