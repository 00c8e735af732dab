"""GPU timing of K3a (geotx_rpe_scores) and K3b (geotx_attention) alone, through the C ABI (eager calls: the attention
numbers include host launch overhead, use tools/k3b_bench.py for those).  usage: k3_bench.py [lib.so]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
if len(sys.argv) > 1:
    pkg._abi.LIB_PATH = os.path.abspath(sys.argv[1])
dev = 'cuda:0'


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for C, n in ((256, 1024), (256, 2048), (128, 2048)):
    H = 4
    torch.manual_seed(0)
    att = pkg.RPEMultiHeadAttention(C, H).to(dev)
    x = torch.randn(1, n, C, device=dev)
    Es = [torch.randn(1, n, n, C, device=dev).to(torch.float16) for _ in range(2)]  # rotate: each > L2
    with torch.no_grad():
        q = att.proj_q(x)
        k = att.proj_k(x)
        v = att.proj_v(x)
        qt = att.folded_query(q)
        sp = torch.empty(1, H, n, n, device=dev)
        it = [0]

        def f_scores():
            it[0] ^= 1
            pkg.ops.rpe_scores(qt, Es[it[0]], out=sp)
        ms = timeit(f_scores)
        gb = n * n * C * 2 / 1e9
        print(f'K3a rpe_scores C={C} n={n}: {ms * 1e3:.1f} us  {gb / ms * 1e3:.0f} GB/s', flush=True)
        ms = timeit(lambda: pkg.ops.attention(q, k, v, H, sp))
        print(f'K3b attention(+scores_p) C={C} n={n}: {ms * 1e3:.1f} us', flush=True)
        ms = timeit(lambda: pkg.ops.attention(q, k, v, H, None))
        print(f'K3b attention(cross)     C={C} n={n}: {ms * 1e3:.1f} us', flush=True)
