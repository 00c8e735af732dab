"""K8 weighted Procrustes: CPU tests pin the oracle restatement on golden vectors written by the
reference's own function (tests/golden/make_golden_procrustes.py); -m gpu tests compare the CUDA path (through the C
ABI) with those vectors and with the oracle.
Tolerance: |T_gpu - T_ref| <= 2e-5 (entries are O(1); both sides accumulate the covariance in fp32 in different
orders, the reference runs LAPACK's fp32 SVD, the kernel a fp64 Jacobi SVD; measured ~2e-6)."""
import numpy as np
import pytest
import torch

from load_reference import load_reference, reference_available

import importlib.util
import os

_spec = importlib.util.spec_from_file_location(
    'make_golden_procrustes', os.path.join(os.path.dirname(__file__), 'golden', 'make_golden_procrustes.py'))
CASES = None


def _cases():
    global CASES
    if CASES is None:
        mod = importlib.util.module_from_spec(_spec)
        _spec.loader.exec_module(mod)
        CASES = mod.CASES
    return CASES


TAGS = ['local', 'thresh', 'noweights', 'planar', 'reflect', 'global']
TOL = 2e-5


def _inputs(pkg, tag):
    kw, thresh, use_w = _cases()[tag]
    src, ref, w, T = pkg.synth.make_procrustes_inputs(**kw)
    return src, ref, (w if use_w else None), thresh, T


@pytest.mark.parametrize('tag', TAGS)
def test_oracle_vs_reference_golden(pkg, oracle, golden, tag):
    z = golden('procrustes_cases.npz')
    src, ref, w, thresh, _ = _inputs(pkg, tag)
    got = oracle.weighted_procrustes(src, ref, w, weight_thresh=thresh)
    want = torch.from_numpy(z[f'{tag}.transform'])
    assert torch.allclose(got, want, rtol=0, atol=1e-6 if torch.__version__ in str(z['meta']) else TOL)
    R = want[:, :3, :3]
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3).expand_as(R), atol=1e-5)
    assert torch.allclose(torch.det(R), torch.ones(R.shape[0]), atol=1e-5)  # proper rotations, also for 'reflect'


@pytest.mark.skipif(not reference_available(), reason='/root/reference not present (GPU box)')
def test_oracle_equals_live_reference(pkg, oracle):
    load_reference()
    from geotransformer.modules.registration.procrustes import weighted_procrustes
    src, ref, w, _ = pkg.synth.make_procrustes_inputs(batch=5, n=30, seed=77)
    saved = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        with torch.no_grad():
            want = weighted_procrustes(src, ref, w, weight_thresh=0.1, return_transform=True)
    finally:
        torch.Tensor.cuda = saved
    assert torch.equal(oracle.weighted_procrustes(src, ref, w, weight_thresh=0.1), want)


def test_host_api_without_gpu(pkg):
    with pytest.raises(RuntimeError):
        pkg.weighted_procrustes(torch.zeros(4, 3), torch.zeros(4, 3))  # CPU tensors: no fallback
    m = pkg.WeightedProcrustes(weight_thresh=0.1, eps=1e-4, return_transform=True)
    assert (m.weight_thresh, m.eps, m.return_transform) == (0.1, 1e-4, True) and len(m.state_dict()) == 0
    import ctypes
    lib = pkg._abi.load()
    p = ctypes.c_void_p(256)
    assert lib.geotx_weighted_procrustes(p, p, None, 0, 10, 0.0, 1e-5, p, None) == -1


# ---- GPU ----------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize('tag', TAGS)
def test_gpu_vs_reference_golden(pkg, oracle, golden, tag):
    from gpu_common import DEV
    z = golden('procrustes_cases.npz')
    src, ref, w, thresh, _ = _inputs(pkg, tag)
    got = pkg.weighted_procrustes(src.to(DEV), ref.to(DEV), None if w is None else w.to(DEV), weight_thresh=thresh,
                                  return_transform=True).cpu()
    want = torch.from_numpy(z[f'{tag}.transform'])
    assert got.shape == want.shape and got.dtype == torch.float32
    assert float((got - want).abs().max()) <= TOL, float((got - want).abs().max())
    R, t = pkg.weighted_procrustes(src.to(DEV), ref.to(DEV), None if w is None else w.to(DEV), weight_thresh=thresh)
    assert torch.equal(R.cpu(), got[:, :3, :3]) and torch.equal(t.cpu(), got[:, :3, 3])
    assert torch.equal(got[:, 3], torch.tensor([0., 0, 0, 1]).expand(got.shape[0], 4))


@pytest.mark.gpu
def test_gpu_full_size_and_properties(pkg, oracle):
    """256 patch problems x 64 / 128 points (config.py:93,100) and a 20000-point global problem against the oracle;
    noise-free inputs recover the true motion; rotating the targets rotates the solution (equivariance)."""
    from gpu_common import DEV
    for batch, n in ((256, 64), (256, 128), (1, 20000)):
        src, ref, w, _ = pkg.synth.make_procrustes_inputs(batch=batch, n=n, seed=5)
        got = pkg.ops.weighted_procrustes(src.to(DEV), ref.to(DEV), w.to(DEV)).cpu()
        want = oracle.weighted_procrustes(src, ref, w)
        assert float((got - want).abs().max()) <= TOL * (5 if n > 10000 else 1)
    src, ref, w, T = pkg.synth.make_procrustes_inputs(batch=64, n=64, seed=6, noise=0.0, outliers=0.0)
    got = pkg.ops.weighted_procrustes(src.to(DEV), ref.to(DEV), w.to(DEV)).cpu()
    assert float((got - T).abs().max()) <= 1e-4  # eps = 1e-5 in the weight normalisation shifts the centroids slightly
    Q = T[:1, :3, :3]
    rot = pkg.ops.weighted_procrustes(src.to(DEV), (ref @ Q.transpose(1, 2)).contiguous().to(DEV), w.to(DEV)).cpu()
    assert torch.allclose(rot[:, :3, :3], Q @ got[:, :3, :3], atol=1e-5)


@pytest.mark.gpu
def test_gpu_edge_cases(pkg):
    from gpu_common import DEV
    # 2-D inputs are a batch of one (procrustes.py:31-38); three points determine the motion
    src = torch.tensor([[0., 0, 0], [1, 0, 0], [0, 1, 0]])
    ref = torch.tensor([[1., 1, 1], [1, 2, 1], [0, 1, 1]])  # rotate 90 degrees about z, translate (1,1,1)
    T = pkg.weighted_procrustes(src.to(DEV), ref.to(DEV), return_transform=True).cpu()
    want = torch.tensor([[0., -1, 0, 1], [1, 0, 0, 1], [0, 0, 1, 1], [0, 0, 0, 1]])
    assert T.shape == (4, 4) and float((T - want).abs().max()) <= 1e-4
    R, t = pkg.WeightedProcrustes()(src.to(DEV), ref.to(DEV), torch.ones(3, device=DEV))
    assert R.shape == (3, 3) and t.shape == (3,)
    # all weights zero: H = 0; still a proper rotation (the reference's own result depends on LAPACK here)
    Z = pkg.ops.weighted_procrustes(src[None].to(DEV), ref[None].to(DEV), torch.zeros(1, 3, device=DEV)).cpu()[0]
    assert torch.allclose(Z[:3, :3] @ Z[:3, :3].t(), torch.eye(3), atol=1e-6) and abs(float(torch.det(Z[:3, :3])) - 1) < 1e-6
    with pytest.raises(RuntimeError):
        pkg.ops.weighted_procrustes(src[None].to(DEV), ref[None, :2].to(DEV))
