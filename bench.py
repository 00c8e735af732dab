#!/usr/bin/env python
"""bench.py -- registration pairs/sec through GeometricTransformer.forward (GSE embedding + RPE attention).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload 3dmatch_1024]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one forward of the hot path over one synthetic registration pair of the workload (BASELINE.json
configs[1]: 3DMatch stage-4 config, N = M = 1024 superpoints, random-init weights, seed 7351).  Rank r works on its
own pairs (pair sharding, no collective on the path; one NCCL all_gather of the outputs at the end).
Prints ONE JSON line on rank 0.

  value     pairs/s, inputs resident in HBM, CUDA-event timed, max over ranks; up to N = 2048 three CUDA-graph instances
            of the forward are in flight on three streams (independent pairs overlap; GEOTX_INFLIGHT)
  e2e       pairs/s through PipelinedGeometricTransformer.map with HOST (pinned) inputs and a host read of the outputs
  roofline  the dominant kernel, timed by CUDA events captured inside the forward's graph: with the tabulated embedding
            (default) the HBM-bound RPE score kernel (algorithmic bytes / time against the measured copy bandwidth), with
            GEOTX_EMBED=tensor the tcgen05 embedding kernel (algorithmic flops against the measured bf16 rate)
  kernels   the other of the two; variants.other_embed_kernel: the same run with the other embedding kernel
  cpu_baseline  the oracle (CPU restatement of the reference's formulation) on this box's host cores

--impl reference times that CPU oracle alone (rank 0 only)."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (synth shape / config, n, m)
    '3dmatch_1024': ('3dmatch', 1024, 1024),   # BASELINE.json configs[1]  (the metric's configuration)
    'kitti_2048': ('kitti', 2048, 2048),       # configs[2]
    'modelnet_717': ('modelnet', 717, 717),    # configs[0]
    '3dmatch_512': ('3dmatch', 512, 512),
    '3dmatch_2048': ('3dmatch', 2048, 2048),
    '3dmatch_4096': ('3dmatch', 4096, 4096),   # configs[3] sweep point named by north_star
    '3dmatch_8192': ('3dmatch', 8192, 8192),   # configs[3]: E = 34 GB per cloud, stored
    '3dmatch_16384': ('3dmatch', 16384, 16384),  # configs[3]: E = 137 GB per cloud: never stored, regenerated per 'self' block
}
# dram__bytes_read.sum + dram__bytes_write.sum of one rpe_scores_tc_kernel launch (ncu --set full, profiles/r01_ncu_k3a_r01.txt)
K3A_NCU_TRAFFIC_BYTES = {'3dmatch_1024': 560.4e6}
# dram__bytes_read.sum + dram__bytes_write.sum of one gse_embed_tc_kernel launch (ncu --set full, profiles/r01_ncu_k2_*)
K2_NCU_TRAFFIC_BYTES = {"3dmatch_1024": 483.98e6}  # profiles/r01_ncu_k2_v4.txt
METRIC = 'registration pairs/sec (embed+RPE-attn) at N superpoints'
UNIT = 'pairs/s'


def measured_peaks():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            p = json.load(f)
        return dict(tensor_sustained=float(p['bf16_tflops_sustained']), tensor_burst=float(p['bf16_tflops']),
                    hbm=float(p['hbm_gbs']), source='measured')
    except Exception:
        return dict(tensor_sustained=1400.0, tensor_burst=1590.0, hbm=6650.0, source='fallback')


def flops_gse(n, C, k):
    """algorithmic flops of one GSE call (one cloud): 2 n^2 C^2 (1+k)."""
    return 2.0 * n * n * C * C * (1 + k)


def flops_pair(n, m, cfg, regenerate=False):
    """SURVEY.md 8d F_min (proj_p folded, E stored); regenerate=True: E is rebuilt for every 'self' block, so the
    embedding flops are executed Ls times ("add (Ls-1) gse and say so")."""
    C, H, k = cfg['hidden_dim'], cfg['num_heads'], cfg['angle_k']
    Ls = sum(b == 'self' for b in cfg['blocks'])
    Lc = len(cfg['blocks']) - Ls
    f = 0.0
    for x in (n, m):
        f += (Ls if regenerate else 1) * flops_gse(x, C, k) + Ls * (16 * x * C * C + 2 * x * C * C + 2 * x * x * C * H + 4 * x * x * C)
        f += 2 * x * cfg['input_dim'] * C + 2 * x * C * cfg['output_dim']
    f += Lc * ((12 * n * C * C + 4 * m * C * C + 4 * n * m * C) + (12 * m * C * C + 4 * n * C * C + 4 * n * m * C))
    return f


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), f'--query-gpu={self.Q}',
                                          '--format=csv,noheader,nounits', '-lms', '200'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for ln in self.lines:
            f = [x.strip() for x in ln.split(',')]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith('active'):
                    reasons.add(nm)
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'samples': len(sm), 'reasons': sorted(reasons)}


# ------------------------------------------------------------------------------------------------------------
# CPU legs (oracle): test infrastructure used only as the timed baseline
# ------------------------------------------------------------------------------------------------------------
def cpu_setup(workload):
    import torch
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import geotx_oracle as oracle
    import pkgpath
    import importlib.util
    spec = importlib.util.spec_from_file_location('geotx_synth', os.path.join(pkgpath.PKG_DIR, 'synth.py'))
    synth = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(synth)
    shape, n, m = WORKLOADS[workload]
    cfg = dict(synth.CONFIGS[shape])
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    inputs = synth.make_pair(shape, n, m, seed=synth.SEED)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    return oracle, cfg, sd, inputs, n, m, cores


def cpu_sample_seconds(oracle, cfg, sd, inputs, rows):
    t0 = time.perf_counter()
    oracle.sampled_forward(sd, cfg, *inputs, rows=rows)
    return time.perf_counter() - t0


def cpu_baseline(workload, budget_s=25.0):
    """Bounded CPU baseline beside the GPU number (rank 0, N=1 only)."""
    oracle, cfg, sd, inputs, n, m, cores = cpu_setup(workload)
    rows = min(64, n)
    cpu_sample_seconds(oracle, cfg, sd, inputs, min(16, n))  # warm the thread pool / allocator
    t = cpu_sample_seconds(oracle, cfg, sd, inputs, rows)
    est_full = t * n / rows
    if est_full <= budget_s:
        t0 = time.perf_counter()
        oracle.geometric_transformer_chunked(sd, cfg, *inputs, rows=128)
        t_full = time.perf_counter() - t0
        return {'value': 1.0 / t_full, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                'sample': f'1 full pair ({workload}), row-chunked oracle (128 rows), {t_full:.2f} s'}
    rows2 = max(rows, min(n, int(rows * budget_s / max(t, 1e-3) / 2)))
    t2 = cpu_sample_seconds(oracle, cfg, sd, inputs, rows2)
    return {'value': (rows2 / n) / t2, 'unit': UNIT, 'cores': cores, 'kind': 'port',
            'sample': f'{rows2} of {n} query rows of every block of 1 pair ({workload}), {t2:.2f} s, scaled by n/rows'}


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    oracle, cfg, sd, inputs, n, m, cores = cpu_setup(args.workload)
    rows = min(128, n)
    for _ in range(args.warmup):
        cpu_sample_seconds(oracle, cfg, sd, inputs, rows)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_sample_seconds(oracle, cfg, sd, inputs, rows)
    dt = (time.perf_counter() - t0) / args.steps
    value = (rows / n) / dt
    sample = f'each step = {rows} of {n} query rows of every block of 1 pair, scaled by n/rows'
    print(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': dt * 1e3, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': args.workload, 'n': n, 'm': m, 'hidden_dim': cfg['hidden_dim'],
                   'note': 'CPU oracle port of the reference formulation (reference is pure PyTorch; /root/reference '
                           'does not travel to the GPU box), all host threads'},
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }))


# ------------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------------
def time_matching(pkg, outputs, num_corr=256, calls=10, replays=5):
    """geotx_superpoint_matching (10 kernels: unit rows, fp32 score tiles, sums, radix select, sort) on one pair's
    output features, replayed from a CUDA graph; returns microseconds per call."""
    import torch
    try:
        ref, src = outputs[0][0].contiguous(), outputs[1][0].contiguous()

        def once():
            return pkg.ops.superpoint_matching(ref, src, num_corr, True, normalize=True)
        once()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(calls):
                keep = once()
        g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(replays):
            g.replay()
        b.record()
        torch.cuda.synchronize()
        us = a.elapsed_time(b) * 1e3 / (calls * replays)
        n, m, d = ref.shape[0], src.shape[0], ref.shape[1]
        del keep
        return {'us_per_pair': us, 'n': n, 'm': m, 'd': d, 'num_correspondences': num_corr, 'kernels_per_call': 10,
                'fp32_tflops_scores': 2.0 * n * m * d / (us * 1e-6) / 1e12, 'dtype': 'f32',
                'timing': f'{calls} calls captured in one CUDA graph, {replays} replays, CUDA events'}
    except Exception as exc:  # keeps the headline line alive
        return {'error': f'{type(exc).__name__}: {exc}'}


def _graph_time_us(once, calls, replays):
    """`calls` invocations captured in one CUDA graph, `replays` replays between two CUDA events; us per invocation."""
    import torch
    once()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(calls):
            keep = once()
    g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(replays):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    del keep
    return a.elapsed_time(b) * 1e3 / (calls * replays)


def time_partition(pkg, shape, num_nodes, calls=10, replays=5):
    """geotx_point_to_node_partition (4 kernels) for BOTH clouds of a pair (model.py:98-103) on
    synthetic fine clouds; microseconds per pair."""
    import torch
    try:
        K, ppn = (128, 24) if shape == 'kitti' else (64, 16)
        clouds = [tuple(t.cuda() for t in pkg.synth.make_partition_inputs(shape, num_nodes, ppn, seed=pkg.synth.SEED + s))
                  for s in (0, 1)]

        def once():
            return [pkg.ops.point_to_node_partition(p, q, K, return_points=True) for p, q in clouds]
        us = _graph_time_us(once, calls, replays)
        N = clouds[0][0].shape[0]
        return {'us_per_pair': us, 'points_per_cloud': N, 'nodes_per_cloud': num_nodes, 'point_limit': K,
                'kernels_per_call': 4, 'calls_per_pair': 2, 'dtype': 'f32 distances, u64 keys',
                'distance_evals_per_s': 2.0 * N * num_nodes / (us * 1e-6),
                'timing': f'{calls} pairs captured in one CUDA graph, {replays} replays, CUDA events'}
    except Exception as exc:
        return {'error': f'{type(exc).__name__}: {exc}'}


def time_transport(pkg, shape, calls=10, replays=5):
    """geotx_log_optimal_transport (one kernel, 100 iterations inside) on the 256 patch score
    matrices of a pair (model.py:184-189); microseconds per pair."""
    import torch
    try:
        K = 128 if shape == 'kitti' else 64
        scores, rm, cm = (t.cuda() for t in pkg.synth.make_patch_scores(256, K, K, 256, seed=pkg.synth.SEED))
        alpha = torch.ones((), device='cuda')

        def once():
            return pkg.ops.log_optimal_transport(scores, alpha, 100, rm, cm)
        us = _graph_time_us(once, calls, replays)
        return {'us_per_pair': us, 'problems': 256, 'rows': K + 1, 'cols': K + 1, 'iterations': 100, 'kernels_per_call': 1,
                'dtype': 'f32', 'ns_per_half_iteration': us * 1e3 / 200,
                'bound': 'latency: one barrier-to-barrier dependency chain per half-iteration (DESIGN.md K7)',
                'timing': f'{calls} calls captured in one CUDA graph, {replays} replays, CUDA events'}
    except Exception as exc:
        return {'error': f'{type(exc).__name__}: {exc}'}


def time_procrustes(pkg, shape, calls=10, replays=5):
    """geotx_weighted_procrustes (one kernel, 3x3 SVD on the device) as the fine matching calls it
    (local_global_registration.py:169,181-191): one batch of 256 patch problems + 5 global problems; us per pair."""
    import torch
    try:
        K = 128 if shape == 'kitti' else 64
        local = [t.cuda() for t in pkg.synth.make_procrustes_inputs(256, K, seed=pkg.synth.SEED)[:3]]
        glob = [t.cuda() for t in pkg.synth.make_procrustes_inputs(1, 5000, seed=pkg.synth.SEED + 1)[:3]]

        def once():
            keep = [pkg.ops.weighted_procrustes(*local)]
            for _ in range(5):
                keep.append(pkg.ops.weighted_procrustes(*glob))
            return keep
        us = _graph_time_us(once, calls, replays)
        return {'us_per_pair': us, 'local_problems': 256, 'points_per_problem': K, 'global_problems': 5,
                'points_per_global_problem': 5000, 'kernels_per_call': 1, 'calls_per_pair': 6, 'dtype': 'f32 (SVD in f64)',
                'host_synchronisations': 0,
                'timing': f'{calls} pairs captured in one CUDA graph, {replays} replays, CUDA events'}
    except Exception as exc:
        return {'error': f'{type(exc).__name__}: {exc}'}


class InFlight:
    """`count` CUDA-graph instances of the forward, each on its own stream: step s replays instance s % count, so that
    the small kernels of one pair (the cross blocks keep few SMs busy) run under the GPU-filling kernels of the next.
    Independent pairs, identical kernels per pair; count = 1 is the plain replay loop."""

    def __init__(self, pkg, model, example, count):
        import torch
        self.torch = torch
        self.graphs = [pkg.GraphedGeometricTransformer(model, *example) for _ in range(count)]
        self.streams = [torch.cuda.Stream() for _ in range(count)] if count > 1 else [None]

    def run(self, resident, n_pairs, steps):
        torch = self.torch
        main = torch.cuda.current_stream()
        out = None
        if self.streams[0] is None:
            for s in range(steps):
                out = self.graphs[0](*resident[s % n_pairs])
            return out
        for st in self.streams:
            st.wait_stream(main)
        for s in range(steps):
            with torch.cuda.stream(self.streams[s % len(self.graphs)]):
                out = self.graphs[s % len(self.graphs)](*resident[s % n_pairs])
        for st in self.streams:
            main.wait_stream(st)
        return out


def time_embed_variant(pkg, model, resident, n_pairs, steps, default_out, inflight):
    """The same forward with the OTHER embedding kernel (tabulated K2t <-> tcgen05 contraction K2): graph replay, CUDA
    events, same pairs; reported beside `value`, not as it."""
    import torch
    default = model.embedding.embed_kernel
    other = 'tensor' if default == 'table' else 'table'
    try:
        model.embedding.embed_kernel = other
        fast = InFlight(pkg, model, resident[0], inflight)
        out = fast.run(resident, n_pairs, 2 * inflight + 1)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = fast.run(resident, n_pairs, steps)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        diff = max(float((x - y).norm() / y.norm()) for x, y in zip(out, default_out))
        del fast
        return {'embed_kernel': other, 'value': steps / (ms * 1e-3), 'unit': 'pairs/s', 'ms_per_step': ms / steps,
                'rel_diff_of_outputs_vs_default': diff}
    except Exception as exc:
        return {'error': f'{type(exc).__name__}: {exc}'}
    finally:
        model.embedding.embed_kernel = default


def time_patch_scores(pkg, shape, calls=10, replays=5):
    """geotx_patch_scores (K12: the glue of model.py:169-170,176-179,187-188 between coarse matching and optimal
    transport) for one pair's 256 node correspondences; us per pair, next to the same lines written with stock torch
    ops on the same GPU (both replayed from a CUDA graph)."""
    import torch
    try:
        K, n, M = (128, 49152, 2048) if shape == 'kitti' else (64, 16384, 1024)
        g = torch.Generator().manual_seed(pkg.synth.SEED)
        feats = [torch.randn(n, 256, generator=g).cuda() for _ in range(2)]
        knn = [torch.randint(0, n + 1, (M, K), generator=g).cuda() for _ in range(2)]
        corr = [torch.randint(0, M, (256,), generator=g).cuda() for _ in range(2)]

        def torch_glue():
            a = torch.cat([feats[0], torch.zeros_like(feats[0][:1])], dim=0)[knn[0][corr[0]]]
            b = torch.cat([feats[1], torch.zeros_like(feats[1][:1])], dim=0)[knn[1][corr[1]]]
            return torch.einsum('bnd,bmd->bnm', a, b) / 256 ** 0.5
        us = _graph_time_us(lambda: pkg.ops.patch_scores(feats[0], feats[1], knn[0], knn[1], corr[0], corr[1]),
                            calls, replays)
        ref = _graph_time_us(torch_glue, calls, replays)
        flops = 2.0 * 256 * K * K * 256
        return {'us_per_pair': us, 'stock_torch_us': ref, 'node_correspondences': 256, 'points_per_patch': K,
                'channels': 256, 'fine_points': n, 'kernels_per_call': 1, 'dtype': 'f32 FMA',
                'tflops': flops / us * 1e-6,
                'timing': f'{calls} pairs captured in one CUDA graph, {replays} replays, CUDA events'}
    except Exception as exc:
        return {'error': f'{type(exc).__name__}: {exc}'}


def time_registration(pkg, shape, calls=10, replays=5):
    """geotx_local_global_registration (4 kernels, K11) on one pair's 256 patch score matrices
    (local_global_registration.py:145-235 as model.py:198-205 calls it); us per pair on the GPU, and the oracle's torch
    restatement of the reference module on the host cores for the same input."""
    import time
    import torch
    try:
        K, k = (128, 2) if shape == 'kitti' else (64, 3)
        inp = pkg.synth.make_lgr_inputs(num_patches=256, K=K, seed=pkg.synth.SEED)
        dev = [t.cuda() for t in inp[:6]]

        def once():
            return pkg.ops.local_global_registration(*dev[:5], None, k, 0.1)
        us = _graph_time_us(once, calls, replays)
        count = int(once()[3].item())
        out = {'us_per_pair': us, 'patch_pairs': 256, 'points_per_patch': K, 'topk': k, 'correspondences': count,
               'kernels_per_call': 4, 'host_synchronisations': 0, 'dtype': 'f32 (moment algebra in f64, Jacobi SVD in f32)',
               'timing': f'{calls} pairs captured in one CUDA graph, {replays} replays, CUDA events'}
        if os.path.join(ROOT, 'oracle') not in sys.path:
            sys.path.insert(0, os.path.join(ROOT, 'oracle'))
        import geotx_oracle
        with torch.no_grad():
            geotx_oracle.local_global_registration(*inp[:6], k=k, acceptance_radius=0.1)
            t0 = time.perf_counter()
            geotx_oracle.local_global_registration(*inp[:6], k=k, acceptance_radius=0.1)
            out['oracle_cpu_us'] = (time.perf_counter() - t0) * 1e6
            out['oracle_cpu_threads'] = torch.get_num_threads()
        return out
    except Exception as exc:
        return {'error': f'{type(exc).__name__}: {exc}'}


def time_pyramid(pkg, calls=3):
    """precompute_data_stack_mode (K10 grid subsampling + K9 radius neighbours, the KPConv pyramid of
    utils/data.py:13-81) for one 3DMatch-sized pair on the GPU, and the same pipeline on the reference's own C++
    operators (oracle/_ref, compiled from the reference's sources) on one host thread; microseconds per pair."""
    import time
    import torch
    try:
        sizes, limits = (20000, 18000), [38, 36, 36, 38]
        p_cpu = torch.cat([pkg.synth.make_dense_cloud(n, 3.0, 800 + b) for b, n in enumerate(sizes)])
        p_gpu = p_cpu.cuda()

        def once():
            return pkg.precompute_data_stack_mode(p_gpu, list(sizes), 4, 0.025, 0.0625, limits)
        once()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(calls):
            once()
        b.record()
        torch.cuda.synchronize()
        out = {'us_per_pair': a.elapsed_time(b) * 1e3 / calls, 'points': list(sizes), 'stages': 4,
               'kernels_per_call': 3 * 14 + 10 * 8, 'timing': f'eager, {calls} calls, CUDA events (4 host reads: 3 subsampled lengths, then the 10 result widths together)'}
        sys.path.insert(0, os.path.join(ROOT, 'tests'))
        import ref_cpu
        if ref_cpu.available():
            t0 = time.perf_counter()
            ref_cpu.precompute_data_stack_mode(p_cpu.numpy(), list(sizes), 4, 0.025, 0.0625, limits)
            out['reference_cpu'] = {'us_per_pair': (time.perf_counter() - t0) * 1e6, 'kind': 'reference', 'cores': 1,
                                    'what': 'oracle/_ref: the reference\'s own grid_subsampling / radius_neighbors C++'}
        return out
    except Exception as exc:
        return {'error': f'{type(exc).__name__}: {exc}'}


def run_b200(args):
    import torch
    import torch.distributed as dist
    import pkgpath
    pkg = pkgpath.load_package()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py needs a CUDA device: ref-geotransformer_b200 has no CPU path')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    pkg._abi.load()

    shape, n, m = WORKLOADS[args.workload]
    cfg = dict(pkg.synth.CONFIGS[shape])
    C, k = cfg['hidden_dim'], cfg['angle_k']
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    torch.manual_seed(pkg.synth.SEED)
    model = pkg.GeometricTransformer(**cfg).to(dev).eval()  # default nn.Linear init under seed 7351 = oracle init
    big = max(n, m) >= 8192  # E of both clouds: 69 GB at 8192 (stored, one graph alive at a time), 275 GB at 16384
    if os.environ.get('GEOTX_E_MODE') is None:
        model.e_mode = 'regenerate' if 2 * max(n, m) ** 2 * C * 2 > 140e9 else 'store'

    # a small rotation of distinct pairs per rank; E (n*n*C*2 bytes per cloud) alone exceeds the 126 MB L2
    n_pairs = 4
    host = [pkg.synth.make_pair(shape, n, m, seed=pkg.synth.SEED + rank * 101 + p) for p in range(n_pairs)]
    host = [tuple(t.pin_memory() for t in pr) for pr in host]
    resident = [tuple(t.to(dev) for t in pr) for pr in host]
    h2d = sum(t.numel() * t.element_size() for t in host[0])
    out_host = [torch.empty(1, x, cfg['output_dim']).pin_memory() for x in (n, m)]
    d2h = sum(t.numel() * t.element_size() for t in out_host)

    # CUDA-event bracket around every launch of the dominant kernel (fused embedding, K2)
    k2_events = []
    orig_embed = model.embedding.embed

    external = [False]

    def timed_embed(geom, out_dtype, row_begin=0, row_count=None, out=None, qt=None):
        kw = dict(enable_timing=True, external=True) if external[0] else dict(enable_timing=True)
        e0, e1 = torch.cuda.Event(**kw), torch.cuda.Event(**kw)
        e0.record()
        res = orig_embed(geom, out_dtype, row_begin, row_count, out, qt=qt)
        e1.record()
        rows = geom.n - row_begin if row_count is None else row_count
        k2_events.append((e0, e1, flops_gse(geom.n, C, k) * rows / geom.n))  # a launch covers `rows` query rows
        return res
    model.embedding.embed = timed_embed

    # the same bracket around every launch of the RPE score kernel (K3a, HBM-bound on E): the dominant kernel when the
    # embedding runs by tabulation
    k3_events = []
    orig_scores = pkg.ops.rpe_scores

    def timed_scores(qt, E, n=None, row_begin=0, out=None):
        kw = dict(enable_timing=True, external=True) if external[0] else dict(enable_timing=True)
        e0, e1 = torch.cuda.Event(**kw), torch.cuda.Event(**kw)
        e0.record()
        res = orig_scores(qt, E, n=n, row_begin=row_begin, out=out)
        e1.record()
        heads = qt.shape[1] if qt.dim() == 4 else cfg['num_heads']
        k3_events.append((e0, e1, E.numel() * E.element_size() + E.numel() // E.shape[-1] * heads * 4))
        return res
    pkg.ops.rpe_scores = timed_scores

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    last = None
    with torch.no_grad():
        # ---- pass 1 (eager launches): per-launch CUDA-event timing of the dominant kernel + launch count ----
        for s in range(3):
            last = model(*resident[s % n_pairs])
        barrier()
        k2_events.clear()
        k3_events.clear()
        launches0 = pkg._abi.launch_count
        e_start, e_stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e_start.record()
        for s in range(args.steps):
            last = model(*resident[s % n_pairs])
        e_stop.record()
        barrier()
        eager_ms = e_start.elapsed_time(e_stop)
        launches = pkg._abi.launch_count - launches0
        k2_ms = [a.elapsed_time(b) for a, b, _ in k2_events]
        k2_launch_flops = [f for _, _, f in k2_events]
        k3_ms = [a.elapsed_time(b) for a, b, _ in k3_events]
        k3_launch_bytes = [f for _, _, f in k3_events]
        k2_timing = 'CUDA events around every launch in an eager pass over the same steps'
        # Preferred: the same event pairs captured INSIDE a CUDA graph of the forward (external event-record nodes), so
        # that the kernel is timed in exactly the context the value pass runs it in (eager launches at n ~ 1k are
        # host-bound: the GPU idles between launches and the kernel then starts from a lower power state).
        try:
            k2_events.clear()
            k3_events.clear()
            external[0] = True
            probe = pkg.GraphedGeometricTransformer(model, *resident[0])
            per_forward = max(len(k2_ms) // max(args.steps, 1), 1)
            pairs_per_replay = list(k2_events)[-per_forward:]  # those recorded during the capture itself
            per_forward3 = max(len(k3_ms) // max(args.steps, 1), 1)
            pairs3_per_replay = list(k3_events)[-per_forward3:]
            in_graph, in_graph_flops = [], []
            in_graph3, in_graph3_bytes = [], []
            for s in range(3 + args.steps):
                probe(*resident[s % n_pairs])
                torch.cuda.synchronize()
                if s >= 3:
                    in_graph += [a.elapsed_time(b) for a, b, _ in pairs_per_replay]
                    in_graph_flops += [f for _, _, f in pairs_per_replay]
                    in_graph3 += [a.elapsed_time(b) for a, b, _ in pairs3_per_replay]
                    in_graph3_bytes += [f for _, _, f in pairs3_per_replay]
            if in_graph and all(t > 0 for t in in_graph):
                k2_ms, k2_launch_flops = in_graph, in_graph_flops
                if in_graph3 and all(t > 0 for t in in_graph3):
                    k3_ms, k3_launch_bytes = in_graph3, in_graph3_bytes
                k2_timing = 'CUDA events recorded inside the CUDA graph of the forward, around every launch'
            del probe
            if big:
                torch.cuda.empty_cache()
        except Exception as exc:  # older torch: no external events in graphs
            k2_timing += f' (in-graph events unavailable: {type(exc).__name__})'
        external[0] = False
        model.embedding.embed = orig_embed
        pkg.ops.rpe_scores = orig_scores

        # ---- pass 2 (the timed region of `value`): the same forward replayed as one CUDA graph per step ----
        # graph-private buffers per instance: E of both clouds (n = 4096: 17 GB); keep the instances alive at once
        # (value, end-to-end pipeline, variant) below ~100 GB
        many = max(n, m) <= 2048
        inflight = max(1, int(os.environ.get('GEOTX_INFLIGHT', '3'))) if many else 1
        fast = InFlight(pkg, model, resident[0], inflight)
        last = fast.run(resident, n_pairs, max(args.warmup, 3, 2 * inflight))
        barrier()
        sampler = ClockSampler(local) if rank == 0 else None
        if sampler:
            sampler.start()
        t_start, t_stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_start.record()
        last = fast.run(resident, n_pairs, args.steps)
        if world > 1:  # the one collective of the job: gather the outputs of the last pair of every rank
            gathered = pkg.dist.gather_pair_outputs([torch.cat(last, dim=1)[0]], world)
        t_stop.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        ms = t_start.elapsed_time(t_stop)

        # ---- end to end: host (pinned) inputs -> device -> forward -> host outputs, every step, through the serving
        #      API (PipelinedGeometricTransformer: the upload of pair p+1 / download of pair p-1 overlap pair p) ----
        depth = 1 if big else (max(1, int(os.environ.get('GEOTX_PIPE_DEPTH', '3'))) if many else 2)
        if big or not many:  # one set of graph-private buffers at a time
            del fast
            torch.cuda.empty_cache()
        pipe = pkg.PipelinedGeometricTransformer(model, *resident[0], depth=depth)
        for o in pipe.map(host[s % n_pairs] for s in range(3)):
            pass
        barrier()
        w0 = time.perf_counter()
        checksum = 0.0
        for o in pipe.map(host[s % n_pairs] for s in range(args.steps)):
            checksum += float(o[0][0, 0, 0]) + float(o[1][0, 0, 0])  # the caller reads every result on the host
        torch.cuda.synchronize()
        barrier()
        e2e_s = time.perf_counter() - w0
        if checksum != checksum:
            raise RuntimeError('end-to-end outputs are NaN')

        # ---- next row: the path's consumer (F.normalize + SuperPointMatching), timed alone on the
        #      outputs of the last pair; not part of `value` ----
        if big:
            del pipe
            torch.cuda.empty_cache()
        matching = time_matching(pkg, last) if world == 1 else None
        shape_name = WORKLOADS[args.workload][0]
        partition = time_partition(pkg, shape_name, n) if world == 1 else None
        transport = time_transport(pkg, shape_name) if world == 1 else None
        procrustes = time_procrustes(pkg, shape_name) if world == 1 else None
        pyramid = time_pyramid(pkg) if world == 1 else None
        registration = time_registration(pkg, shape_name) if world == 1 else None
        patch_scores = time_patch_scores(pkg, shape_name) if world == 1 else None
        variant = None
        if world == 1 and not big and C in (128, 256) and k == 3:
            variant = time_embed_variant(pkg, model, resident, n_pairs, args.steps, last, inflight)

    tt = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms, e2e_ms = tt.tolist()
    if rank == 0:
        peaks = measured_peaks()
        value = world * args.steps / (ms * 1e-3)
        k2_avg_ms = sum(k2_ms) / max(len(k2_ms), 1)
        k2_flops = sum(k2_launch_flops) / max(len(k2_ms), 1)  # average over launches (row slabs of E)
        achieved = k2_flops / (k2_avg_ms * 1e-3) / 1e12 if k2_ms else 0.0
        step_ms = ms / args.steps
        k3_avg_ms = sum(k3_ms) / max(len(k3_ms), 1)
        k3_bytes = sum(k3_launch_bytes) / max(len(k3_ms), 1)
        k3_gbs = k3_bytes / (k3_avg_ms * 1e-3) / 1e9 if k3_ms else 0.0
        table = model.embedding.embed_kernel == 'table' and model.embedding._use_table(model.embedding.storage_dtype(), None)
        timing_note = k2_timing + ('; the value pass replays the identical kernels from a CUDA graph; '
                                   'share_of_step = timed launches per step / ms_per_step')
        tensor_roof = {'kernel': 'gse_embed_tc_kernel (geotx_gse_embed)', 'bound': 'tensor', 'achieved': achieved,
                       'peak': peaks['tensor_sustained'], 'unit': 'TFLOP/s',
                       'frac': achieved / peaks['tensor_sustained'],
                       'traffic': K2_NCU_TRAFFIC_BYTES.get(args.workload),
                       'peak_source': peaks['source'] + ' bf16 sustained (kernel timed inside the step)',
                       'frac_of_burst_peak': achieved / peaks['tensor_burst'],
                       'flops_per_launch': k2_flops, 'avg_launch_ms': k2_avg_ms, 'launches_timed': len(k2_ms),
                       'share_of_step': sum(k2_ms) / args.steps / step_ms if ms else None, 'timing': timing_note}
        hbm_roof = {'kernel': 'rpe_scores_tc_kernel (geotx_rpe_scores)', 'bound': 'hbm', 'achieved': k3_gbs,
                    'peak': peaks['hbm'], 'unit': 'GB/s', 'frac': k3_gbs / peaks['hbm'],
                    'traffic': K3A_NCU_TRAFFIC_BYTES.get(args.workload),
                    'peak_source': peaks['source'] + ' copy bandwidth',
                    'bytes_per_launch': k3_bytes, 'avg_launch_ms': k3_avg_ms, 'launches_timed': len(k3_ms),
                    'share_of_step': sum(k3_ms) / args.steps / step_ms if ms else None,
                    'algorithmic_bytes': 'E read once (n*n*C*2) + scores written (n*n*H*4) per launch',
                    'timing': timing_note}
        if table:  # the embedding runs by tabulation (shared-memory bound): the HBM-bound score kernel dominates
            roofline = hbm_roof
            kernel_stats = {'embedding': {'kernel': 'gse_embed_table_kernel (geotx_gse_embed_table)',
                                          'bound': 'shared-memory pipe (10 table rows = 20 C bytes of shared-memory reads per pair)',
                                          'avg_launch_ms': k2_avg_ms, 'launches_timed': len(k2_ms),
                                          'share_of_step': sum(k2_ms) / args.steps / step_ms if ms else None,
                                          'lds_bytes_per_launch': 20.0 * C * k2_flops / flops_gse(n, C, k) * n * n if n else None}}
        else:
            roofline = tensor_roof
            kernel_stats = {'rpe_scores': hbm_roof}
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': max(args.warmup, 3), 'ms_per_step': ms / args.steps, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f16' if table else 'bf16', 'data': 'synthetic',
            'config': {'workload': args.workload, 'n': n, 'm': m, 'hidden_dim': C, 'num_heads': cfg['num_heads'],
                       'blocks': len(cfg['blocks']), 'angle_k': k, 'pairs_per_step_per_gpu': 1, 'launch': 'one CUDA graph replay per step' + (f'; {inflight} graph instances in flight on {inflight} streams (independent pairs overlap)' if inflight > 1 else ''),
                       'e_storage': model.last_e_mode + (f' (E never stored: rebuilt for each of the {sum(b == "self" for b in cfg["blocks"])} self blocks in {model.regenerate_slab_bytes >> 20} MiB row slabs; pair_tflops counts those passes)' if model.last_e_mode == 'regenerate' else ' (one embedding pass per cloud)'),
                       'embed_kernel': 'table (geotx_gse_embed_table)' if table else 'tensor (geotx_gse_embed, tcgen05)',
                       'l2_slab_bytes': model.l2_slab_bytes, 'first_block_scores_fused_into_embedding': bool(model.fuse_scores),
                       'l2': f'inputs larger than L2: E is {n * n * C * 2 / 1e6:.0f} MB per cloud, '
                             f'{n_pairs} distinct pairs rotated',
                       'sharding': f'by pair over {world} GPU(s); one all_gather of outputs'},
            'e2e': {'value': world * args.steps / (e2e_ms * 1e-3), 'unit': UNIT, 'h2d_bytes_per_step': h2d,
                    'd2h_bytes_per_step': d2h,
                    'api': f'PipelinedGeometricTransformer.map over pinned host pairs, depth {depth}, wall clock'},
            'gpu_launches': launches,
            'roofline': roofline,
            'kernels': kernel_stats,
            'eager_ms_per_step': eager_ms / args.steps,
            'pair_tflops': flops_pair(n, m, cfg, model.last_e_mode == 'regenerate') / 1e12,
            'pair_tflops_min': flops_pair(n, m, cfg) / 1e12,
            'pair_tensor_frac': flops_pair(n, m, cfg, model.last_e_mode == 'regenerate') * value / world / 1e12 / peaks['tensor_sustained'],
            'clocks': clocks,
            'next_rows': {'superpoint_matching': matching, 'point_to_node_partition': partition,
                          'log_optimal_transport': transport, 'weighted_procrustes': procrustes,
                          'patch_scores': patch_scores, 'local_global_registration': registration, 'collate_pyramid': pyramid},
            'variants': {'other_embed_kernel': variant},
        }
        if world == 1 and not args.no_cpu_baseline:
            line['cpu_baseline'] = cpu_baseline(args.workload)
        else:
            line['cpu_baseline'] = None
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--workload', default='3dmatch_1024', choices=sorted(WORKLOADS))
    ap.add_argument('--no-cpu-baseline', action='store_true')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)


if __name__ == '__main__':
    main()
