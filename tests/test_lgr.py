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
