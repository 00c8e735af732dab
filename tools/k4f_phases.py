"""Phase clocks of one K4f CTA (variant library built with -DPA_PHASE_CLOCKS: tools/build_variants.sh k4f_clk
k4_post_attention.cu -DPA_PHASE_CLOCKS; run with GEOTX_LIB=tools/bin/k4f_clk.so)."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import pkgpath  # noqa: E402

pkg = pkgpath.load_package()
dev = torch.device('cuda', 0)
C, T = 256, int(sys.argv[1]) if len(sys.argv) > 1 else 2048
g = torch.Generator().manual_seed(3)
r = lambda *s: torch.randn(*s, generator=g).to(dev)  # noqa: E731
hidden, x = r(1, T, C), r(1, T, C)
w = [pkg.ops.pack_linear(r(C, C) / 16), pkg.ops.pack_linear(r(2 * C, C) / 16), pkg.ops.pack_linear(r(C, 2 * C) / 22)]
args = (hidden, x, w[0], r(C), r(C), r(C), 1e-5, w[1], r(2 * C), w[2], r(C), r(C), r(C), 1e-5)
names = ['ring init + sync', 'prologue (tile loads)', 'GEMM 0 (C x C)', 'epilogue 0 + sync', 'LayerNorm 1 + sync', 'GEMM 1 (C -> 2C)',
         'epilogue 1 + sync', 'GEMM 2 (2C -> C)', 'epilogue 2 + sync', 'LayerNorm 2 + store']
lib = ctypes.CDLL(pkg._abi.LIB_PATH)
for rep in range(4):
    pkg.ops.post_attention(*args)
    torch.cuda.synchronize()
    buf = (ctypes.c_longlong * 16)()
    assert lib.geotx_debug_k4f_phases(buf) == 0
    if rep >= 2:
        t = list(buf)[:10]
        print(f'T={T}: total {t[9] - t[0]} clk; ' + '; '.join(f'{names[i + 1]} {t[i + 1] - t[i]}' for i in range(9)))
