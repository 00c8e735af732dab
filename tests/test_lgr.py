"""K11 LocalGlobalRegistration (the consumer of the transport and Procrustes rows).  CPU tests pin the oracle
restatement on golden vectors written by the reference's own module (tests/golden/make_golden_lgr.py); -m gpu tests
compare the CUDA path (through the C ABI) with them.
Tolerances: the correspondence set is index work on exp(scores) -- the same pairs in the same (patch, i, j) order, scores
1e-6 relative (expf vs torch.exp); the transform <= 2e-4 abs (it goes through a hypothesis selection by inlier count and
five re-weighted Procrustes steps; a correspondence whose residual sits within rounding of the radius may flip)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from load_reference import load_reference, reference_available

_spec = importlib.util.spec_from_file_location(
    'make_golden_lgr', os.path.join(os.path.dirname(__file__), 'golden', 'make_golden_lgr.py'))
_mod = None
TAGS = ['3dmatch', 'kitti', 'small', 'nonmutual', 'global_score', 'degenerate']
T_TOL = 2e-4


def _cases():
    global _mod
    if _mod is None:
        _mod = importlib.util.module_from_spec(_spec)
        _spec.loader.exec_module(_mod)
    return _mod.CASES


@pytest.mark.parametrize('tag', TAGS)
def test_oracle_vs_reference_golden(pkg, oracle, golden, tag):
    z = golden('lgr_cases.npz')
    kw, mkw = _cases()[tag]
    inp = pkg.synth.make_lgr_inputs(**kw)
    rc, sc, s, T = oracle.local_global_registration(*inp[:6], **mkw)
    assert np.array_equal(rc.numpy(), z[f'{tag}.ref_corr_points'])
    assert np.array_equal(sc.numpy(), z[f'{tag}.src_corr_points'])
    assert np.allclose(s.numpy(), z[f'{tag}.corr_scores'], rtol=1e-6, atol=0)
    assert float(np.abs(T.numpy() - z[f'{tag}.transform']).max()) <= 1e-5


@pytest.mark.skipif(not reference_available(), reason='/root/reference not present (GPU box)')
def test_oracle_equals_live_reference(pkg, oracle):
    load_reference()
    from geotransformer.modules.geotransformer.local_global_registration import LocalGlobalRegistration
    inp = pkg.synth.make_lgr_inputs(num_patches=10, K=16, seed=9)
    saved = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        with torch.no_grad():
            want = LocalGlobalRegistration(3, 0.1)(*inp[:6])
    finally:
        torch.Tensor.cuda = saved
    got = oracle.local_global_registration(*inp[:6], k=3, acceptance_radius=0.1)
    assert all(torch.equal(a, b) for a, b in zip(got[:3], want[:3]))
    assert torch.allclose(got[3], want[3], atol=1e-6)


def test_host_api_without_gpu(pkg):
    """Constructor mirrors the reference; unsupported switches and CPU tensors raise instead of falling back."""
    mod = pkg.LocalGlobalRegistration(3, 0.1)
    assert (mod.k, mod.acceptance_radius, mod.mutual, mod.confidence_threshold, mod.correspondence_threshold,
            mod.num_refinement_steps) == (3, 0.1, True, 0.05, 3, 5)
    with pytest.raises(NotImplementedError):
        pkg.LocalGlobalRegistration(3, 0.1, use_dustbin=True)
    with pytest.raises(NotImplementedError):
        pkg.LocalGlobalRegistration(3, 0.1, correspondence_limit=100)
    inp = pkg.synth.make_lgr_inputs(num_patches=4, K=8, seed=1)
    with pytest.raises(RuntimeError):
        mod(*inp[:6])


def _gpu_inputs(pkg, kw, dev):
    inp = pkg.synth.make_lgr_inputs(**kw)
    return [t.to(dev) for t in inp[:6]], inp[6]


@pytest.mark.gpu
@pytest.mark.parametrize('tag', TAGS)
def test_gpu_vs_reference_golden(pkg, golden, tag):
    from gpu_common import DEV
    z = golden('lgr_cases.npz')
    kw, mkw = _cases()[tag]
    inp, _ = _gpu_inputs(pkg, kw, DEV)
    rc, sc, s, T = pkg.LocalGlobalRegistration(**mkw)(*inp)
    assert np.array_equal(rc.cpu().numpy(), z[f'{tag}.ref_corr_points'])
    assert np.array_equal(sc.cpu().numpy(), z[f'{tag}.src_corr_points'])
    assert np.allclose(s.cpu().numpy(), z[f'{tag}.corr_scores'], rtol=1e-6, atol=0)
    assert float(np.abs(T.cpu().numpy() - z[f'{tag}.transform']).max()) <= T_TOL


@pytest.mark.gpu
def test_gpu_strided_score_view_and_oracle(pkg, oracle):
    """The model passes matching_scores[:, :-1, :-1] (experiments/*/model.py:195-196): a view with row stride K+1."""
    from gpu_common import DEV
    for seed, P, K, k in ((61, 32, 48, 3), (62, 200, 64, 2), (63, 7, 100, 3)):
        inp = pkg.synth.make_lgr_inputs(num_patches=P, K=K, seed=seed)
        padded = torch.full((P, K + 1, K + 1), -20.0)
        padded[:, :K, :K] = inp[4]
        dev_in = [t.to(DEV) for t in inp[:6]]
        dev_in[4] = padded.to(DEV)[:, :-1, :-1]
        assert not dev_in[4].is_contiguous()
        got = pkg.LocalGlobalRegistration(k, 0.1)(*dev_in)
        want = oracle.local_global_registration(*inp[:6], k=k, acceptance_radius=0.1)
        assert torch.equal(got[0].cpu(), want[0]) and torch.equal(got[1].cpu(), want[1])
        assert torch.allclose(got[2].cpu(), want[2], rtol=1e-6, atol=0)
        assert float((got[3].cpu() - want[3]).abs().max()) <= T_TOL
        assert float((got[3].cpu() - inp[6]).abs().max()) < 5e-2  # and it recovers the planted motion


@pytest.mark.gpu
def test_gpu_full_size_properties(pkg):
    """Full 3DMatch size (256 patch pairs x 64 points, experiments/*3DMatch*/config.py:88,116-125) through
    size-independent properties: counts within the top-k bound, every correspondence made of valid patch points, the
    result invariant to a rigid motion applied to the source side (the transform composes), deterministic."""
    from gpu_common import DEV
    P, K, k = 256, 64, 3
    inp, T_true = _gpu_inputs(pkg, dict(num_patches=P, K=K, seed=71), DEV)
    mod = pkg.LocalGlobalRegistration(k, 0.1)
    rc, sc, s, T = mod(*inp)
    rc2, sc2, s2, T2 = mod(*inp)
    assert torch.equal(rc, rc2) and torch.equal(sc, sc2) and torch.equal(s, s2) and torch.equal(T, T2)
    C = rc.shape[0]
    assert 0 < C <= P * K * k
    assert bool((s > 0.05).all()) and bool((s <= 1.0 + 1e-6).all())
    valid_ref = inp[0][inp[2]]
    valid_src = inp[1][inp[3]]
    exact = dict(compute_mode='donot_use_mm_for_euclid_dist')
    assert bool((torch.cdist(rc, valid_ref, **exact).min(dim=1).values == 0).all())
    assert bool((torch.cdist(sc, valid_src, **exact).min(dim=1).values == 0).all())
    R = T[:3, :3]
    assert float((R @ R.T - torch.eye(3, device=DEV)).abs().max()) < 1e-5 and abs(float(torch.det(R)) - 1) < 1e-5
    assert float((T.cpu() - T_true).abs().max()) < 5e-2
    # move the source side by a known motion G: the estimate must become T @ G^-1
    ang = 0.7
    G = torch.eye(4, device=DEV)
    G[:3, :3] = torch.tensor([[np.cos(ang), -np.sin(ang), 0], [np.sin(ang), np.cos(ang), 0], [0, 0, 1]], device=DEV)
    G[:3, 3] = torch.tensor([0.3, -0.2, 0.5], device=DEV)
    moved = inp[1] @ G[:3, :3].T + G[:3, 3]
    _, _, s3, T3 = mod(inp[0], moved, *inp[2:])
    assert torch.equal(s3, s)
    assert float((T3 @ G - T).abs().max()) < 1e-3


@pytest.mark.gpu
def test_gpu_edge_cases(pkg):
    from gpu_common import DEV
    mod = pkg.LocalGlobalRegistration(3, 0.1)
    inp, _ = _gpu_inputs(pkg, dict(num_patches=4, K=16, seed=81), DEV)
    # nothing above the confidence threshold: no correspondences, a finite transform, no crash
    low = torch.full_like(inp[4], -10.0)
    rc, sc, s, T = mod(inp[0], inp[1], inp[2], inp[3], low, inp[5])
    assert rc.shape == (0, 3) and sc.shape == (0, 3) and s.shape == (0,) and bool(torch.isfinite(T).all())
    # all points masked on one side
    none = torch.zeros_like(inp[2])
    rc, _, _, T = mod(inp[0], inp[1], none, inp[3], inp[4], inp[5])
    assert rc.shape[0] == 0 and bool(torch.isfinite(T).all())
    # limits of the kernel are reported, not silently clipped
    big = pkg.LocalGlobalRegistration(9, 0.1)
    with pytest.raises(RuntimeError):
        big(*inp)
    with pytest.raises(RuntimeError):
        mod(inp[0].cpu(), *inp[1:])
