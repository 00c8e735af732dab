"""Helpers shared by the -m gpu parity tests.  Every GPU call goes through the C ABI (ops.py -> _abi.py -> the
in-tree libgeotx_b200.so); the oracle (oracle/geotx_oracle.py) and the committed golden vectors are the checkers."""
import numpy as np
import torch

DEV = 'cuda:0'

# tolerances (BASELINE.json north_star): indices bit-exact; fp32 CUDA-core path 1e-4; bf16 tensor-core path 2e-2
TOL_FP32 = 1e-4
TOL_BF16 = 2e-2


def rel(a, b):
    a = torch.as_tensor(np.asarray(a) if not isinstance(a, torch.Tensor) else a).detach().float().cpu()
    b = torch.as_tensor(np.asarray(b) if not isinstance(b, torch.Tensor) else b).detach().float().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def sd_from_npz(z):
    return {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith('sd.')}


def build_model(pkg, cfg, sd, precision='auto'):
    m = pkg.GeometricTransformer(**cfg, precision=precision).to(DEV).eval()
    m.load_state_dict(sd, strict=True)
    return m
