#!/usr/bin/env python
"""bench.py -- registration pairs/sec through GeometricTransformer.forward (GSE embedding + RPE attention).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload 3dmatch_1024]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one forward of the hot path over one synthetic registration pair of the workload (BASELINE.json
configs[1]: 3DMatch stage-4 config, N = M = 1024 superpoints, random-init weights, seed 7351) on every rank.  Rank r
works on its own pairs (pair sharding, no collective on the path); with N > 1 the outputs of ALL timed
steps of all ranks are gathered with one NCCL all_gather inside the timed region (after a warm-up gather).
Prints ONE JSON line on rank 0.

  value      pairs/s, inputs resident in HBM, CUDA-event timed, max over ranks; up to n = 2048 several CUDA-graph instances
             of the forward are in flight on their own streams (independent pairs overlap; GEOTX_INFLIGHT, default 6 up
             to n = 1024: measured 1390 / 1436 / 1447 / 1455 / 1454 pairs/s with 3 / 4 / 5 / 6 / 8; 4 at 2048, 3 at 4096)
  e2e        pairs/s through PipelinedGeometricTransformer.map with HOST (pinned) inputs and a host read of the outputs
  roofline   the dominant kernel of the default path, K3r (geotx_rpe_scores_cheb: low-rank embedding + folded-query
             scores, E never stored), timed by CUDA events captured inside the forward's graph: flops the kernel
             EXECUTES / time against the measured bf16 tensor rate, with its share of the legacy mma.sync pipe beside it
  kernels    every other kernel class of the step, timed the same way (launches, time, share of the step)
  variants   the same forward (a) with E stored: K2c writes it, K3a reads it (the HBM-bound formulation, both with their
             fraction of the measured copy bandwidth); (b) with the full-rank tcgen05 embedding K2 (north_star's "% of
             tensor-pipe peak": SURVEY.md 8d's F_min flops / time against the measured bf16 rate); (c) with the library
             GEMM glue of round 1
  sweep      bounded runs of 3dmatch_4096 (north_star's size) and kitti_2048 in the same process
  configs4   BASELINE.json configs[4]: 256 pairs in total sharded over the ranks, all outputs gathered (strong scaling)
  cpu_baseline  the reference's formulation on this box's host cores (the reference itself when its tree is on the box)

--impl reference times the reference's CPU implementation alone (rank 0 only): the UNMODIFIED reference classes when a
reference tree is present (GEOTX_REFERENCE_ROOT, baseline/_ref, /root/reference), else the oracle port; whole pairs."""
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
    '3dmatch_8192': ('3dmatch', 8192, 8192),
    '3dmatch_16384': ('3dmatch', 16384, 16384),
}
# dram__bytes_read.sum + dram__bytes_write.sum per launch from `ncu --set full` captures of the SHIPPED kernels
# (profiles/r02_forward_3dmatch_1024_ncu_full.txt; None where not captured).  K3r: 5.12 MB read (the folded queries, u
# and the geometry) and 0 B written INSIDE the launch: its 16.8 MB of scores stay in the 126 MB L2 until the attention
# kernel that follows has read them (that launch shows 19.9 MB of DRAM reads under ncu's cache flush, ~0 in the step).
NCU_TRAFFIC_BYTES = {
    ('gse_embed_cheb_kernel<SCORES> (geotx_rpe_scores_cheb, K3r)', '3dmatch_1024'): 5.12e6,
}
METRIC = 'registration pairs/sec (embed+RPE-attn) at N superpoints'
UNIT = 'pairs/s'
CONFIGS4_PAIRS = 256


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

    def wait_ready(self, timeout=5.0):
        """Blocks until nvidia-smi has printed its first sample: its start-up (fork, NVML initialisation: tens of ms, and it
        takes driver locks) must be over before the timed region begins."""
        t0 = time.time()
        while self.proc is not None and not self.lines and time.time() - t0 < timeout:
            time.sleep(0.01)

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
# CPU legs: test infrastructure used only as the timed baseline
# ------------------------------------------------------------------------------------------------------------
def find_reference_root():
    """A tree of the UNMODIFIED reference on this machine, if any (never present on a plain GPU box)."""
    for cand in (os.environ.get('GEOTX_REFERENCE_ROOT'), os.path.join(ROOT, 'baseline', '_ref'), '/root/reference'):
        if cand and os.path.isdir(os.path.join(cand, 'geotransformer', 'modules')):
            return cand
    return None


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


def cpu_forward_fn(workload):
    """-> (callable running ONE whole pair of the workload on the host cores, kind, cores, description).
    kind 'reference': the unmodified reference classes (geotransformer.py:114-155) imported from a reference tree, run
    whole when its (n, n, k, C) intermediates fit the host memory, else driven in query-row chunks through its own
    submodules; kind 'port': the oracle's row-chunked restatement (no reference tree on this machine)."""
    import torch
    oracle, cfg, sd, inputs, n, m, cores = cpu_setup(workload)
    root = find_reference_root()
    if root is not None:
        os.environ['GEOTX_REFERENCE_ROOT'] = root
        import importlib
        import load_reference
        importlib.reload(load_reference)
        GT, _ = load_reference.load_reference()
        torch.manual_seed(7351)
        model = GT(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['num_heads'], cfg['blocks'], cfg['sigma_d'],
                   cfg['sigma_a'], cfg['angle_k'], reduction_a=cfg.get('reduction_a', 'max')).eval()
        model.load_state_dict(sd, strict=True)
        whole = max(n, m) ** 2 * cfg['angle_k'] * cfg['hidden_dim'] * 4 * 5 < 0.5 * _host_memory_bytes()
        if whole:
            def run():
                with torch.no_grad():
                    return model(*inputs)
            how = 'unmodified reference GeometricTransformer.forward'
        else:
            sys.path.insert(0, os.path.join(ROOT, 'tests', 'golden'))
            from make_golden_large import chunked_forward

            def run():
                with torch.no_grad():
                    return chunked_forward(model, *inputs, 64)[:2]
            how = 'unmodified reference submodules driven in 64-row chunks (its (n,n,k,C) tensors do not fit this host)'
        return run, 'reference', cores, f'{how} from {root}'

    def run():
        return oracle.geometric_transformer_chunked(sd, cfg, *inputs, rows=128)
    return run, 'port', cores, 'oracle port (row-chunked restatement of the reference; no reference tree on this machine)'


def _host_memory_bytes():
    try:
        return os.sysconf('SC_PAGE_SIZE') * os.sysconf('SC_PHYS_PAGES')
    except (ValueError, OSError):
        return 32 << 30


def cpu_baseline(workload, max_pairs=3, budget_s=25.0):
    """Bounded CPU baseline beside the GPU number (rank 0, N=1 only): whole pairs, as many as fit ~25 s (at least one)."""
    run, kind, cores, how = cpu_forward_fn(workload)
    times = []
    t_all = time.perf_counter()
    while len(times) < max_pairs and (not times or time.perf_counter() - t_all + times[-1] < budget_s):
        t0 = time.perf_counter()
        run()
        times.append(time.perf_counter() - t0)
    best = min(times)
    return {'value': 1.0 / best, 'unit': UNIT, 'cores': cores, 'kind': kind,
            'sample': f'{len(times)} whole pair(s) of {workload}, best {best:.2f} s; {how}'}


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    shape, n, m = WORKLOADS[args.workload]
    run, kind, cores, how = cpu_forward_fn(args.workload)
    for _ in range(min(args.warmup, 1)):  # one untimed pair warms the thread pool / allocator (a pair takes seconds)
        run()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run()
    dt = (time.perf_counter() - t0) / args.steps
    value = 1.0 / dt
    sample = f'each step = 1 whole pair of {args.workload}; {how}'
    print(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': min(args.warmup, 1), 'ms_per_step': dt * 1e3, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': bench_config(args.workload, args.gpus),
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': kind, 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }))


def bench_config(workload, world):
    """The `config` object both arms print (identical on both, so that the driver's same_config check reads true); what
    is specific to the GPU arm goes under `setup`."""
    shape, n, m = WORKLOADS[workload]
    return {'workload': workload, 'n': n, 'm': m, 'pairs_per_step_per_gpu': 1, 'sharding': f'by pair over {world} GPU(s)',
            'l2': 'inputs larger than L2: distinct pairs rotated whose inputs together exceed the 126 MB L2 (count in setup.l2)'}


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


class InFlight:
    """`count` CUDA-graph instances of the forward, each on its own stream: step s replays instance s % count, so that
    the small kernels of one pair (the cross blocks keep few SMs busy) run under the GPU-filling kernels of the next.
    Independent pairs, identical kernels per pair; count = 1 is the plain replay loop.  `keep` (a (steps, n+m, Dout)
    tensor) receives every step's output features (what a multi-GPU job gathers at its end)."""

    def __init__(self, pkg, model, example, count):
        import torch
        self.torch = torch
        self.joint = pkg.modules._joint_view
        self.graphs = [pkg.GraphedGeometricTransformer(model, *example, throughput=count > 1) for _ in range(count)]
        self.streams = [torch.cuda.Stream() for _ in range(count)] if count > 1 else [None]

    def run(self, resident, n_pairs, steps, keep=None):
        torch = self.torch
        main = torch.cuda.current_stream()
        out = None

        def one(s):
            o = self.graphs[s % len(self.graphs)](*resident[s % n_pairs])
            if keep is not None:
                n = o[0].shape[1]
                dst = keep.local if hasattr(keep, 'pair_done') else keep
                both = self.joint(o[0], o[1])
                if both is not None:  # the two outputs are slices of one (1, n+m, C) buffer: one copy
                    dst[s].copy_(both[0], non_blocking=True)
                else:
                    dst[s, :n].copy_(o[0][0], non_blocking=True)
                    dst[s, n:].copy_(o[1][0], non_blocking=True)
                if hasattr(keep, 'pair_done'):
                    keep.pair_done(s)  # pkg.dist.OverlappedGather: its chunk travels while later steps compute
            return o
        if self.streams[0] is None:
            for s in range(steps):
                out = one(s)
            return out
        for st in self.streams:
            st.wait_stream(main)
        for s in range(steps):
            with torch.cuda.stream(self.streams[s % len(self.graphs)]):
                out = one(s)
        for st in self.streams:
            main.wait_stream(st)
        return out


class KernelTimers:
    """CUDA-event brackets around the calls of the C-ABI wrappers (`pkg.ops.<fn>`), usable inside a CUDA-graph capture
    (external events): the kernels are timed where they really run, inside the replayed step."""

    def __init__(self, torch):
        self.torch, self.external, self.records, self._restore = torch, False, {}, []

    def wrap(self, owner, attr, name, work=None):
        orig = getattr(owner, attr)
        timers = self

        def timed(*a, **k):
            kw = dict(enable_timing=True, external=True) if timers.external else dict(enable_timing=True)
            e0, e1 = timers.torch.cuda.Event(**kw), timers.torch.cuda.Event(**kw)
            e0.record()
            res = orig(*a, **k)
            e1.record()
            timers.records.setdefault(name, []).append((e0, e1, work(*a, **k) if work else 0.0))
            return res
        setattr(owner, attr, timed)
        self._restore.append((owner, attr, orig))

    def clear(self):
        self.records = {}

    def restore(self):
        for owner, attr, orig in self._restore:
            setattr(owner, attr, orig)
        self._restore = []

    def summary(self, records=None):
        """name -> (list of ms, list of work) over the given (or all) event pairs; synchronise first."""
        out = {}
        for name, evs in (records or self.records).items():
            out[name] = ([a.elapsed_time(b) for a, b, _ in evs], [w for _, _, w in evs])
        return out


def host_d_set(pkg, points, sigma_d):
    """The distance-expansion set K2c / K3r pick on the device for this cloud (largest distance / sigma_d against the
    sets' ranges), recomputed on the host for reporting the executed flops."""
    import torch
    p = points[0].double()
    d_max = float(torch.cdist(p, p).max()) / sigma_d if p.shape[0] <= 8192 else float((p.max(0)[0] - p.min(0)[0]).norm()) / sigma_d
    lib = pkg._abi.load()
    for s in range(6):
        if d_max <= float(lib.geotx_gse_cheb_d_range(s)):
            return s, d_max
    return 5, d_max


KD_TILES = (1, 2, 3, 4, 6, 8)


def measure_mma_sync_peak(pkg, torch, dev):
    """TFLOP/s of mma.sync.m16n8k16 (f16, fp32 accumulate) on register operands, all SMs busy (geotx_mma_sync_probe)."""
    import ctypes
    try:
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        ctas, iters = sms * 4, 4000
        out = torch.empty(ctas * 256, device=dev)
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        best = 0.0
        for rep in range(4):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            pkg._abi.call('geotx_mma_sync_probe', ctas, iters, ctypes.c_void_p(out.data_ptr()), st)
            b.record()
            torch.cuda.synchronize()
            if rep:
                best = max(best, ctas * 8.0 * iters * 16 * 4096 / (a.elapsed_time(b) * 1e-3) / 1e12)
        return best
    except Exception:
        return None


def default_inflight(n):
    """CUDA-graph instances of the forward kept in flight at cloud size n (measured, see run_b200)."""
    return 6 if n <= 1024 else 4 if n <= 2048 else 3 if n <= 4096 else 1


def timed_value(torch, fast, resident, n_pairs, steps, warm):
    fast.run(resident, n_pairs, warm)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    out = fast.run(resident, n_pairs, steps)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b), out


def probe_kernels(pkg, torch, model, timers, resident, n_pairs, steps, throughput=False):
    """Per-kernel-class times INSIDE a replayed CUDA graph of the forward: the wrappers record external event pairs
    during the capture; every replay refreshes them.  -> name -> dict(ms per launch, launches per step, work per launch)."""
    timers.clear()
    timers.external = True
    try:
        probe = pkg.GraphedGeometricTransformer(model, *resident[0], warmup=1, throughput=throughput)  # the timed step's schedule
        # the wrappers saw two forwards: the eager warm-up and the captured one; the capture's event pairs are the second
        # half of every list
        rec = {name: evs[len(evs) // 2:] for name, evs in timers.records.items()}
        acc = {name: ([], [w for _, _, w in evs]) for name, evs in rec.items()}
        for s in range(3 + steps):
            probe(*resident[s % n_pairs])
            torch.cuda.synchronize()
            if s >= 3:
                for name, evs in rec.items():
                    acc[name][0].append([a.elapsed_time(b) for a, b, _ in evs])
        del probe
        out = {}
        for name, (per_replay, work) in acc.items():
            launches = len(work)
            ms = [sum(r[i] for r in per_replay) / len(per_replay) for i in range(launches)]
            out[name] = {'launches_per_step': launches, 'ms_per_launch': ms, 'work_per_launch': work}
        return out, 'CUDA events recorded inside the CUDA graph of the forward, around every launch, averaged over the replays'
    finally:
        timers.external = False
        timers.clear()


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
    peaks = measured_peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def build(workload):
        shape, n, m = WORKLOADS[workload]
        cfg = dict(pkg.synth.CONFIGS[shape])
        torch.manual_seed(pkg.synth.SEED)
        model = pkg.GeometricTransformer(**cfg).to(dev).eval()  # default nn.Linear init under seed 7351 = oracle init
        return shape, n, m, cfg, model

    def make_pairs(shape, n, m, count, first_seed):
        host = [pkg.synth.make_pair(shape, n, m, seed=first_seed + p) for p in range(count)]
        host = [tuple(t.pin_memory() for t in pr) for pr in host]
        return host, [tuple(t.to(dev) for t in pr) for pr in host]

    shape, n, m, cfg, model = build(args.workload)
    C, k, H = cfg['hidden_dim'], cfg['angle_k'], cfg['num_heads']
    Ls = sum(b == 'self' for b in cfg['blocks'])
    big = max(n, m) >= 8192
    # inputs larger than L2: a rotation of distinct pairs per rank whose inputs together exceed the 126 MB L2 (on top of
    # that every step streams its n^2 H fp32 score matrices, 16.8 MB per 'self' block and cloud at n = 1024)
    pair_bytes = (n + m) * (cfg['input_dim'] + 3) * 4
    n_pairs = max(4, -(-140_000_000 // pair_bytes))  # distinct input pairs rotated: together larger than the 126 MB L2
    host, resident = make_pairs(shape, n, m, n_pairs, pkg.synth.SEED + rank * 101)
    h2d = sum(t.numel() * t.element_size() for t in host[0])
    d2h = (n + m) * cfg['output_dim'] * 4
    d_set = max(host_d_set(pkg, host[0][0], cfg['sigma_d'])[0], host_d_set(pkg, host[0][1], cfg['sigma_d'])[0])
    kd = 16 * KD_TILES[d_set]
    # the widest cloud of the rotation decides how many distance coefficients the embedding kernel must be prepared for
    # (the module widens itself on the FIRST call of a shape only; a later pair of the rotation may be wider)
    widest = max(host_d_set(pkg, pr[c], cfg['sigma_d'])[1] for pr in host for c in (0, 1))
    if widest > model.embedding.cheb_d_index_max:
        lib = pkg._abi.load()
        model.embedding.cheb_d_index_max = next(float(lib.geotx_gse_cheb_d_range(s)) for s in range(6)
                                                if widest <= float(lib.geotx_gse_cheb_d_range(s)) or s == 5)

    def k3r_flops(geom, coef, qt, *a, **kw):  # flops the kernel EXECUTES: expansions of the k angle slots, scores hi + lo,
        rows = geom.n                          # and the distance term (inside E, or split off in coefficient space)
        split = kw.get('u') is not None and H <= 4
        dist = 2.0 * kd * 8 * 2 if split else 2.0 * C * kd
        score = 2.0 * C * 8 * (1 if split else 2)  # split instances: hi and lo halves of the query share one MMA's 8 columns
        return geom.B * rows * geom.n * (2.0 * C * (k * 16 * coef.ka_tiles) + dist + score)

    def embed_bytes(geom, out_dtype, row_begin=0, row_count=None, out=None, qt=None):
        rows = geom.n - row_begin if row_count is None else row_count
        return float(geom.B * rows * geom.n * C * (4 if out_dtype == torch.float32 else 2))

    def k3a_bytes(qt, E, n=None, row_begin=0, out=None):
        return float(E.numel() * E.element_size() + E.numel() // E.shape[-1] * qt.shape[1] * 4)

    timers = KernelTimers(torch)
    timers.wrap(pkg.ops, 'rpe_scores_cheb', 'rpe_scores_cheb (K3r)', k3r_flops)
    timers.wrap(pkg.ops, 'rpe_scores', 'rpe_scores (K3a)', k3a_bytes)
    timers.wrap(model.embedding, 'embed', 'gse_embed (K2c / K2 / K2t)', embed_bytes)
    timers.wrap(pkg.ops, 'attention', 'attention (K3h)',
                lambda q, kk, v, heads, *a, **kw: 4.0 * q.shape[0] * q.shape[1] * kk.shape[1] * q.shape[2])
    timers.wrap(pkg.ops, 'post_attention', 'post_attention (K4f)',
                lambda hidden, *a, **kw: 2.0 * (hidden.numel() // hidden.shape[-1]) * 5 * hidden.shape[-1] ** 2)
    timers.wrap(pkg.ops, 'linear_f16', 'linear_f16 (K13)',
                lambda x, w, *a, **kw: 2.0 * (x.numel() // x.shape[-1]) * w.shape[0] * w.shape[1])
    timers.wrap(pkg.ops, 'gse_knn', 'gse_knn (K1)', lambda points, kk: float(points.shape[0] * points.shape[1] ** 2 * 16))
    timers.wrap(pkg.ops, 'add_layernorm', 'add_layernorm (K4)', None)

    def serial_ms(per_kernel):
        """One forward's kernels back to back (the probe replays ONE graph instance alone): the denominator of every
        share_of_step.  The timed region overlaps several instances, so ms_per_step is smaller than this sum."""
        return sum(sum(d['ms_per_launch']) for d in per_kernel.values())

    def kernel_report(per_kernel, step_ms):
        rep = {}
        whole = serial_ms(per_kernel)
        for name, d in sorted(per_kernel.items()):
            total = sum(d['ms_per_launch'])
            rep[name] = {'launches_per_step': d['launches_per_step'], 'ms_per_step': total,
                         'avg_launch_ms': total / max(d['launches_per_step'], 1),
                         'share_of_step': total / whole if whole else None}
        return rep

    def roof_tensor(name, d, step_ms, mma_peak, what):
        ms = sum(d['ms_per_launch']) / max(d['launches_per_step'], 1)
        fl = sum(d['work_per_launch']) / max(d['launches_per_step'], 1)
        ach = fl / (ms * 1e-3) / 1e12 if ms else 0.0
        r = {'kernel': name, 'bound': 'tensor', 'achieved': ach, 'peak': peaks['tensor_sustained'], 'unit': 'TFLOP/s',
             'frac': ach / peaks['tensor_sustained'], 'traffic': NCU_TRAFFIC_BYTES.get((name, args.workload)),
             'peak_source': peaks['source'] + ' bf16 sustained (cuBLAS on tcgen05; the kernel is timed inside the step)',
             'flops_per_launch': fl, 'flops': what, 'avg_launch_ms': ms, 'launches_per_step': d['launches_per_step'],
             'share_of_step': sum(d['ms_per_launch']) / step_ms if step_ms else None}
        if mma_peak:
            r['mma_sync_peak_tflops'] = mma_peak
            r['frac_of_mma_sync_peak'] = ach / mma_peak
            r['pipe'] = ('mma.sync.m16n8k16 (legacy tensor path): the accumulators must be max-reduced over the k angle slots '
                         'in registers; through tcgen05 they would have to be read back from TMEM at 64 B/clk/SM, 3x slower '
                         'than the contraction itself (DESIGN.md section 4)')
        return r

    def roof_hbm(name, d, step_ms, what):
        ms = sum(d['ms_per_launch']) / max(d['launches_per_step'], 1)
        by = sum(d['work_per_launch']) / max(d['launches_per_step'], 1)
        gbs = by / (ms * 1e-3) / 1e9 if ms else 0.0
        return {'kernel': name, 'bound': 'hbm', 'achieved': gbs, 'peak': peaks['hbm'], 'unit': 'GB/s',
                'frac': gbs / peaks['hbm'], 'traffic': NCU_TRAFFIC_BYTES.get((name, args.workload)),
                'peak_source': peaks['source'] + ' copy bandwidth', 'bytes_per_launch': by, 'algorithmic_bytes': what,
                'avg_launch_ms': ms, 'launches_per_step': d['launches_per_step'],
                'share_of_step': sum(d['ms_per_launch']) / step_ms if step_ms else None}

    # graph instances in flight (independent pairs overlap; tools/ablate_step.py --inflight): 6 up to n = 1024 (1436 ->
    # 1455 pairs/s against 4), 4 at 2048 (6: the same), 3 at 4096 (105.8 -> 108.0 against 2; 4: 108.0), 1 above
    many = max(n, m) <= 2048
    mid = not many and max(n, m) <= 4096
    inflight = max(1, int(os.environ.get('GEOTX_INFLIGHT', str(default_inflight(max(n, m)))))) if (many or mid) else 1

    def run_variant(tag, setup, undo, steps, want_kernels=True):
        """The same forward under another configuration: value (graph replay, CUDA events) + its kernel table."""
        try:
            setup()
            with torch.no_grad():
                out = [t.clone() for t in model(*resident[0])]  # compared with the default path on the SAME pair
                fast = InFlight(pkg, model, resident[0], inflight)
                ms, _ = timed_value(torch, fast, resident, n_pairs, steps, max(3, 2 * inflight))
                del fast
                res = {'value': steps / (ms * 1e-3), 'unit': UNIT, 'ms_per_step': ms / steps, 'e_storage': model.last_e_mode}
                if want_kernels:
                    per_kernel, _ = probe_kernels(pkg, torch, model, timers, resident, n_pairs, min(steps, 10), inflight > 1)
                    res['kernels'] = kernel_report(per_kernel, ms / steps)
                    res['_raw'] = per_kernel
            return res, out
        except Exception as exc:  # keeps the headline line alive
            return {'error': f'{type(exc).__name__}: {exc}'}, None
        finally:
            undo()
            torch.cuda.empty_cache()

    with torch.no_grad():
        # ---- pass 1 (eager): launch count through the C ABI ----
        for s in range(3):
            last = model(*resident[s % n_pairs])
        barrier()
        if args.launch_list:  # the forward's launches, one by one, for `ncu --metrics gpu__time_duration.sum`
            model.transformer.throughput = inflight > 1  # the kernels of the timed step's schedule
            model(*resident[0])
            torch.cuda.synchronize()
            torch.cuda.profiler.start()
            for s in range(args.steps):
                model(*resident[s % n_pairs])
            torch.cuda.synchronize()
            torch.cuda.profiler.stop()
            print(json.dumps({'launch_list': args.workload, 'forwards': args.steps, 'e_storage': model.last_e_mode,
                              'note': 'not a bench line: run under ncu --profile-from-start off'}))
            return
        launches0 = pkg._abi.launch_count
        e_start, e_stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e_start.record()
        for s in range(args.steps):
            last = model(*resident[s % n_pairs])
        e_stop.record()
        barrier()
        eager_ms = e_start.elapsed_time(e_stop)
        launches = pkg._abi.launch_count - launches0
        default_mode = model.last_e_mode
        default_out = [t.clone() for t in model(*resident[0])]  # what the variants below are compared with

        # ---- per-kernel times inside the replayed graph of the forward ----
        per_kernel, kernel_timing = probe_kernels(pkg, torch, model, timers, resident, n_pairs, args.steps, inflight > 1)
        mma_peak = measure_mma_sync_peak(pkg, torch, dev) if rank == 0 else None

        # ---- pass 2 (the timed region of `value`): the forward replayed as one CUDA graph per step ----
        fast = InFlight(pkg, model, resident[0], inflight)
        # N > 1: the outputs of every step of every rank are gathered inside the timed region, in chunks of 4 steps on a
        # communication stream behind the steps that produce them (pkg.dist.OverlappedGather)
        shape_out = (n + m, cfg['output_dim'])
        gather_chunk = max(1, int(os.environ.get('GEOTX_GATHER_CHUNK', '4')))
        if world > 1:  # warm-up gather: NCCL channel setup and buffer registration stay outside the timed region
            warm = pkg.dist.make_gather(args.steps * world, shape_out, dev, chunk=gather_chunk)
            fast.run(resident, n_pairs, args.steps, keep=warm)
            warm.finish()
            del warm
        last = fast.run(resident, n_pairs, max(args.warmup, 3, 2 * inflight))
        keep = pkg.dist.make_gather(args.steps * world, shape_out, dev, chunk=gather_chunk) if ((world > 1 and os.environ.get('GEOTX_KEEP') != '0') or os.environ.get('GEOTX_KEEP') == '1') else None
        # the clock sampler (rank 0 spawns nvidia-smi) is started, and has delivered its first sample, BEFORE the barrier that
        # opens the timed region: started after it, rank 0 entered the region ~0.7 ms late (and nvidia-smi initialised
        # inside it), and with a gather at the end every other rank's time included that wait
        sampler = ClockSampler(local) if rank == 0 else None
        if sampler:
            sampler.start()
            sampler.wait_ready()
        t_start, t_stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        t_start.record()
        last = fast.run(resident, n_pairs, args.steps, keep=keep)
        if keep is not None:  # the one collective of the job: every rank ends up with the features of all steps x ranks pairs
            gathered = keep.finish()
        t_stop.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        ms = t_start.elapsed_time(t_stop)
        gather_bytes = args.steps * world * (n + m) * cfg['output_dim'] * 4 if world > 1 else 0

        # ---- BASELINE.json configs[4]: 256 pairs in total, sharded by pair, every output gathered (strong scaling) ----
        configs4 = None
        if not big:
            try:
                mine = pkg.dist.shard_pairs(CONFIGS4_PAIRS, rank, world)
                keep4 = pkg.dist.make_gather(CONFIGS4_PAIRS, shape_out, dev, chunk=8)
                barrier()
                c_start, c_stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                c_start.record()
                fast.run(resident, n_pairs, len(mine), keep=keep4)
                all4 = keep4.finish()
                c_stop.record()
                barrier()
                configs4 = {'ms': c_start.elapsed_time(c_stop), 'pairs': CONFIGS4_PAIRS, 'pairs_per_rank': len(mine),
                            'gathered_shape': list(all4.shape), 'gather_bytes': all4.numel() * 4}
                del all4, keep4
            except Exception as exc:
                configs4 = {'error': f'{type(exc).__name__}: {exc}'}

        # ---- end to end: host (pinned) inputs -> device -> forward -> host outputs, every step, through the serving
        #      API (PipelinedGeometricTransformer: the upload of pair p+1 / download of pair p-1 overlap pair p) ----
        depth = 1 if big else (max(1, int(os.environ.get('GEOTX_PIPE_DEPTH', str(inflight + 1)))) if (many or mid) else 2)
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
        del pipe
        torch.cuda.empty_cache()

    tt = torch.tensor([ms, e2e_s * 1e3, configs4['ms'] if configs4 and 'ms' in configs4 else 0.0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms, e2e_ms, c4_ms = tt.tolist()

    variants, sweep, next_rows = {}, {}, None
    if rank == 0 and world == 1:
        emb = model.embedding

        def set_mode(e_mode=None, kernel=None, dense=None):
            def f():
                if e_mode:
                    model.e_mode = e_mode
                if kernel:
                    emb.embed_kernel = kernel
                if dense:
                    os.environ['GEOTX_DENSE'] = dense
            return f

        def reset():
            model.e_mode, emb.embed_kernel = 'auto', 'cheb'
            os.environ.pop('GEOTX_DENSE', None)
        vsteps = min(args.steps, 20)
        if not big and default_mode == 'scores':
            # (a) E stored: K2c writes it once per cloud, K3a reads it once per 'self' block -- the HBM-bound formulation
            v, out = run_variant('stored_e', set_mode(e_mode='store'), reset, vsteps)
            if 'error' not in v:
                raw = v.pop('_raw')
                v['rel_diff_of_outputs_vs_default'] = max(float((x - y).norm() / y.norm()) for x, y in zip(out, default_out))
                if 'gse_embed (K2c / K2 / K2t)' in raw:
                    v['embedding_write'] = roof_hbm('gse_embed_cheb_kernel (geotx_gse_embed_cheb, K2c)', raw['gse_embed (K2c / K2 / K2t)'],
                                                    serial_ms(raw), 'E written once: n*n*C*2 bytes per launch')
                if 'rpe_scores (K3a)' in raw:
                    v['scores_read'] = roof_hbm('rpe_scores_tc_kernel (geotx_rpe_scores, K3a)', raw['rpe_scores (K3a)'], serial_ms(raw),
                                                'E read once (n*n*C*2) + scores written (n*n*H*4) per launch')
            variants['stored_e'] = v
            # (b) the full-rank tcgen05 embedding: north_star's "% of tensor-pipe peak" on the formulation it names
            v, out = run_variant('tensor_embedding', set_mode(e_mode='store', kernel='tensor'), reset, vsteps)
            if 'error' not in v:
                raw = v.pop('_raw')
                v['rel_diff_of_outputs_vs_default'] = max(float((x - y).norm() / y.norm()) for x, y in zip(out, default_out))
                if 'gse_embed (K2c / K2 / K2t)' in raw:
                    d = dict(raw['gse_embed (K2c / K2 / K2t)'])
                    d['work_per_launch'] = [flops_gse(n if i % 2 == 0 else m, C, k) for i in range(d['launches_per_step'])]
                    v['embedding_tensor_roof'] = roof_tensor('gse_embed_tc_kernel (geotx_gse_embed, K2, tcgen05)', d, serial_ms(raw),
                                                             None, 'SURVEY.md 8d: 2 n^2 C^2 (1+k) per cloud')
                v['pair_tflops_min'] = flops_pair(n, m, cfg) / 1e12
                v['pair_tensor_frac'] = flops_pair(n, m, cfg) * v['value'] / 1e12 / peaks['tensor_sustained']
                v['pair_tensor_frac_note'] = 'F_min of SURVEY.md 8d (the flops this variant really executes) x pairs/s / measured bf16 sustained'
            variants['tensor_embedding'] = v
            # (c) round 1's dense glue (cuBLASLt TF32 GEMMs + K4) under the default embedding / scores
            v, out = run_variant('library_gemm_glue', set_mode(dense='cublas'), reset, vsteps, want_kernels=False)
            if 'error' not in v:
                v['rel_diff_of_outputs_vs_default'] = max(float((x - y).norm() / y.norm()) for x, y in zip(out, default_out))
            variants['library_gemm_glue'] = v

        # ---- bounded sweep points in the same process (driver-visible): north_star's size and KITTI ----
        if args.workload == '3dmatch_1024' and not args.no_sweep:
            for wl, ssteps in (('3dmatch_4096', 5), ('kitti_2048', 16)):
                try:
                    s_shape, s_n, s_m, s_cfg, s_model = build(wl)
                    _, s_res = make_pairs(s_shape, s_n, s_m, 2, pkg.synth.SEED)
                    entry = {}
                    with torch.no_grad():
                        for tag, e_mode, kernel in (('default', 'auto', 'cheb'), ('tensor_embedding', 'store', 'tensor')):
                            s_model.e_mode, s_model.embedding.embed_kernel = e_mode, kernel
                            s_model(*s_res[0])
                            s_inflight = default_inflight(max(s_n, s_m)) if inflight > 1 else 1  # the rule of the main line
                            sf = InFlight(pkg, s_model, s_res[0], s_inflight)
                            s_ms, _ = timed_value(torch, sf, s_res, 2, ssteps, max(3, 2 * s_inflight))
                            del sf
                            torch.cuda.empty_cache()
                            val = ssteps / (s_ms * 1e-3)
                            entry[tag] = {'value': val, 'unit': UNIT, 'ms_per_step': s_ms / ssteps, 'e_storage': s_model.last_e_mode,
                                          'graph_instances_in_flight': s_inflight}
                            if tag == 'tensor_embedding':
                                entry[tag]['pair_tflops_min'] = flops_pair(s_n, s_m, s_cfg) / 1e12
                                entry[tag]['pair_tensor_frac'] = flops_pair(s_n, s_m, s_cfg) * val / 1e12 / peaks['tensor_sustained']
                        entry['default']['fmin_equivalent_tensor_frac'] = (flops_pair(s_n, s_m, s_cfg) * entry['default']['value'] / 1e12 /
                                                                           peaks['tensor_sustained'])
                        entry['note'] = ('fmin_equivalent_tensor_frac = what running SURVEY.md 8d\'s F_min at this rate would need; the '
                                         'default path executes ~12x fewer flops (low-rank embedding), so it is a throughput '
                                         'equivalent, not a utilisation')
                    sweep[wl] = entry
                    del s_model, s_res
                    torch.cuda.empty_cache()
                except Exception as exc:
                    sweep[wl] = {'error': f'{type(exc).__name__}: {exc}'}

        with torch.no_grad():
            shape_name = WORKLOADS[args.workload][0]
            next_rows = {'superpoint_matching': time_matching(pkg, last),
                         'point_to_node_partition': time_partition(pkg, shape_name, n),
                         'log_optimal_transport': time_transport(pkg, shape_name),
                         'weighted_procrustes': time_procrustes(pkg, shape_name),
                         'patch_scores': time_patch_scores(pkg, shape_name),
                         'local_global_registration': time_registration(pkg, shape_name),
                         'collate_pyramid': time_pyramid(pkg)}
    timers.restore()

    if rank == 0:
        value = world * args.steps / (ms * 1e-3)
        step_ms = ms / args.steps
        kernels = kernel_report(per_kernel, step_ms)
        serial_step_ms = serial_ms(per_kernel)
        dom = 'rpe_scores_cheb (K3r)'
        if dom in per_kernel:
            roofline = roof_tensor('gse_embed_cheb_kernel<SCORES> (geotx_rpe_scores_cheb, K3r)', per_kernel[dom], serial_ms(per_kernel), mma_peak,
                                   f'executed: per (i,j) pair 2 C 16 k ka for the expansions of the {k} angle slots + 2 C 8 for the '
                                   f'folded-query scores (hi | lo halves in the 8 columns of one MMA; 2 C 8 2 without the split) + the distance term (distance set {d_set}: {kd} coefficients): '
                                   f'2 {kd} 8 2 in coefficient space (distance split) or 2 C {kd} inside E')
            dtype = 'f16'
        elif 'rpe_scores (K3a)' in per_kernel:
            roofline = roof_hbm('rpe_scores_tc_kernel (geotx_rpe_scores, K3a)', per_kernel['rpe_scores (K3a)'], serial_ms(per_kernel),
                                'E read once (n*n*C*2) + scores written (n*n*H*4) per launch')
            dtype = 'f16'
        else:
            roofline, dtype = None, 'f32'
        if roofline is not None:
            roofline['timing'] = kernel_timing + ('; share_of_step = this kernel\'s launches / all launches of one forward, one graph instance replayed alone '
                                                     '(what an ncu launch list sees); the timed region overlaps several instances, so '
                                                     'ms_per_step is smaller than the serial sum (serial_ms_per_step)')
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': max(args.warmup, 3), 'ms_per_step': step_ms, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': dtype, 'data': 'synthetic',
            'config': bench_config(args.workload, world),
            'setup': {
                'hidden_dim': C, 'num_heads': H, 'blocks': len(cfg['blocks']), 'angle_k': k,
                'launch': 'one CUDA graph replay per step' + (f'; {inflight} graph instances in flight on {inflight} streams (independent pairs overlap)' if inflight > 1 else ''),
                'e_storage': default_mode + (' (E never formed in memory: K3r rebuilds it tile by tile in registers for each of the '
                                             f'{Ls} self blocks)' if default_mode == 'scores' else ''),
                'embed_kernel': model.embedding.embed_kernel, 'dense_glue': os.environ.get('GEOTX_DENSE', 'own'),
                'l2': f'inputs larger than L2: {n_pairs} distinct pairs ({n_pairs * pair_bytes / 1e6:.0f} MB of inputs) rotated over '
                      f'{inflight} graph instance(s); every step also streams {2 * Ls * n * n * H * 4 / 1e6:.0f} MB of score matrices',
                'gather': f'outputs of every step of every rank gathered inside the timed region ({gather_bytes / 1e6:.0f} MB): chunks of 4 steps on a communication stream behind the steps that produce them ({type(keep).__name__}: ' + ('copy-engine writes into the peers\' buffers mapped through CUDA IPC, no collective kernel' if type(keep).__name__ == 'PeerGather' else 'NCCL all_gather per chunk') + '), after one warm-up pass' if world > 1 else 'single GPU: nothing to gather'},
            'e2e': {'value': world * args.steps / (e2e_ms * 1e-3), 'unit': UNIT, 'h2d_bytes_per_step': h2d,
                    'd2h_bytes_per_step': d2h,
                    'api': f'PipelinedGeometricTransformer.map over pinned host pairs, depth {depth}, wall clock'},
            'gpu_launches': launches,
            'gpu_launches_per_step': launches / args.steps,
            'roofline': roofline,
            'kernels': kernels,
            'serial_ms_per_step': serial_step_ms,
            'eager_ms_per_step': eager_ms / args.steps,
            'pair_tflops_executed': sum(sum(d['work_per_launch']) for nme, d in per_kernel.items()
                                        if nme.split(' ')[0] in ('rpe_scores_cheb', 'attention', 'post_attention', 'linear_f16')) / 1e12,
            'pair_tflops_min_reference_formulation': flops_pair(n, m, cfg) / 1e12,
            'fmin_equivalent_tensor_frac': flops_pair(n, m, cfg) * value / world / 1e12 / peaks['tensor_sustained'],
            'fmin_equivalent_note': 'throughput equivalent only: SURVEY.md 8d F_min x pairs/s / measured bf16 sustained; the default '
                                    'path does not execute F_min (see pair_tflops_executed); variants.tensor_embedding.pair_tensor_frac '
                                    'is the utilisation figure of the formulation north_star names',
            'clocks': clocks,
            'configs4': ({'value': CONFIGS4_PAIRS / (c4_ms * 1e-3), 'unit': UNIT, 'scaling': 'strong', 'ms': c4_ms,
                          **{kk: vv for kk, vv in configs4.items() if kk != 'ms'}} if configs4 and 'ms' in configs4 else configs4),
            'next_rows': next_rows,
            'variants': variants,
            'sweep': sweep,
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
    ap.add_argument('--no-sweep', action='store_true')
    ap.add_argument('--launch-list', action='store_true',
                    help='ncu pass (B200_PROFILING.md launch list): warm up, then --steps eager forwards between '
                         'cudaProfilerStart/Stop (ncu --profile-from-start off) and nothing else; prints no bench line')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)


if __name__ == '__main__':
    main()
