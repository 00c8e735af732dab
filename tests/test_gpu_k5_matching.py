"""-m gpu: K5 coarse matching (F.normalize + SuperPointMatching), through the C ABI.

Checkers: golden vectors written by the reference's own SuperPointMatching (tests/golden/make_golden_spm.py), the
oracle restatement, and -- for the selection, which is integer work -- torch.topk on the kernel's own score matrix.
Tolerances: scores 1e-5 relative (fp32 arithmetic on both sides, different summation orders); correspondence
indices equal to the reference's wherever the reference's neighbouring scores differ by more than 2e-5 relative,
and otherwise a pick whose reference score is within 2e-5 of the reference's pick at that rank."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from gpu_common import DEV

pytestmark = pytest.mark.gpu
RTOL = 1e-5
TIE = 2e-5
CASES = ['small_dual', 'small_plain', 'small_masked', 'small_allpairs', 'mid_dual']


def _check_against_reference(oracle, ref, src, rm, sm, K, dual, got_ri, got_si, got_sc, exp_ri, exp_si, exp_sc):
    """got_* from the GPU, exp_* from the reference (golden)."""
    assert len(got_sc) == len(exp_sc)
    assert torch.all(got_sc[:-1] >= got_sc[1:]), 'scores must be descending'
    flat = got_ri * src.shape[0] + got_si
    assert flat.unique().numel() == flat.numel(), 'a correspondence was returned twice'
    assert torch.allclose(got_sc, exp_sc, rtol=RTOL, atol=0), (got_sc - exp_sc).abs().max()
    # the reference's score of every pick (full matrix from the oracle, original indexing)
    rn, sn = F.normalize(ref, p=2, dim=1), F.normalize(src, p=2, dim=1)
    keep_r, keep_s = torch.nonzero(rm, as_tuple=True)[0], torch.nonzero(sm, as_tuple=True)[0]
    sub = oracle.superpoint_matching_scores(rn[keep_r], sn[keep_s], dual)
    full = torch.full((ref.shape[0], src.shape[0]), -1.0)
    full[keep_r[:, None], keep_s[None, :]] = sub
    assert rm[got_ri].all() and sm[got_si].all(), 'a removed superpoint was selected'
    mine = full[got_ri, got_si]
    same = (got_ri == exp_ri) & (got_si == exp_si)
    near = (mine - exp_sc).abs() <= TIE * exp_sc
    assert torch.all(same | near), f'{int((~(same | near)).sum())} picks are neither the reference\'s nor near-ties'
    # ranks whose reference score is isolated from both neighbours must agree exactly
    gap = (exp_sc[:-1] - exp_sc[1:]) / exp_sc[:-1]
    isolated = torch.ones(len(exp_sc), dtype=torch.bool)
    isolated[:-1] &= gap > TIE
    isolated[1:] &= gap > TIE
    kth_next = sub.flatten().topk(min(K + 1, sub.numel()))[0]
    if len(kth_next) > len(exp_sc):
        isolated[-1] &= bool((exp_sc[-1] - kth_next[-1]) / exp_sc[-1] > TIE)
    assert torch.all(same[isolated]), 'an isolated rank differs from the reference'
    return int(same.sum()), int(isolated.sum())


@pytest.mark.parametrize('tag', CASES)
def test_matching_vs_reference_golden(pkg, oracle, golden, tag):
    z = golden('spm_cases.npz')
    cfg = eval(str(z[f'{tag}.cfg']))
    ref, src = torch.from_numpy(z[f'{tag}.ref_feats']), torch.from_numpy(z[f'{tag}.src_feats'])
    rm, sm = torch.from_numpy(z[f'{tag}.ref_masks']), torch.from_numpy(z[f'{tag}.src_masks'])
    matcher = pkg.SuperPointMatching(cfg['K'], cfg['dual'])
    ru, su, ri, si, sc = matcher.match_normalized(ref.to(DEV), src.to(DEV), rm.to(DEV), sm.to(DEV))
    assert ri.dtype == torch.int64 and si.dtype == torch.int64 and sc.dtype == torch.float32
    assert torch.allclose(ru.cpu(), F.normalize(ref, p=2, dim=1), rtol=1e-6, atol=1e-9)
    assert torch.allclose(su.cpu(), F.normalize(src, p=2, dim=1), rtol=1e-6, atol=1e-9)
    same, isolated = _check_against_reference(
        oracle, ref, src, rm, sm, cfg['K'], cfg['dual'], ri.cpu(), si.cpu(), sc.cpu(),
        torch.from_numpy(z[f'{tag}.ref_idx']), torch.from_numpy(z[f'{tag}.src_idx']), torch.from_numpy(z[f'{tag}.scores']))
    assert isolated >= len(sc) // 2, 'the fixture should pin most ranks exactly'
    # the reference's call signature with already-normalised rows gives the same answer
    ri2, si2, sc2 = matcher(ru, su, rm.to(DEV), sm.to(DEV))
    assert torch.equal(ri2, ri) and torch.equal(si2, si) and torch.equal(sc2, sc)


def test_matching_on_the_reference_transformer_output(pkg, oracle, golden):
    """BASELINE configs[1]: the matcher fed with the reference GeometricTransformer's own outputs at N = M = 1024."""
    z, zin = golden('spm_cases.npz'), golden('3dmatch_n1024_out.npz')
    ref, src = torch.from_numpy(zin['out_ref'][0]), torch.from_numpy(zin['out_src'][0])
    ones_r, ones_s = torch.ones(ref.shape[0], dtype=torch.bool), torch.ones(src.shape[0], dtype=torch.bool)
    ri, si, sc = pkg.SuperPointMatching(256, True)(ref.to(DEV), src.to(DEV), normalize=True)
    _check_against_reference(oracle, ref, src, ones_r, ones_s, 256, True, ri.cpu(), si.cpu(), sc.cpu(),
                             torch.from_numpy(z['3dmatch_1024.ref_idx']), torch.from_numpy(z['3dmatch_1024.src_idx']),
                             torch.from_numpy(z['3dmatch_1024.scores']))


@pytest.mark.parametrize('n,m,d,K,dual,masked', [
    (1024, 1024, 256, 256, True, False), (1000, 777, 256, 256, True, True), (333, 2050, 64, 1000, False, True),
    (4096, 4096, 256, 256, True, False), (65, 3, 7, 4096, True, False), (1, 1, 4, 5, True, False)])
def test_selection_is_exact_on_its_own_matrix(pkg, n, m, d, K, dual, masked):
    """Integer work: given the kernel's score matrix, the K returned entries are exactly torch.topk's (values bit-equal,
    positions equal), at full size, for ragged shapes, masks, K > n*m; and the matrix equals an fp32 torch evaluation."""
    g = torch.Generator().manual_seed(n * 7 + m)
    ref, src = torch.randn(n, d, generator=g).to(DEV), torch.randn(m, d, generator=g).to(DEV)
    rm = (torch.rand(n, generator=g) > 0.2).to(DEV) if masked else None
    sm = (torch.rand(m, generator=g) > 0.2).to(DEV) if masked else None
    out = pkg.ops.superpoint_matching(ref, src, K, dual, rm, sm, normalize=True, return_matrix=True)
    mat = out['matrix']
    removed = mat.view(torch.int32) == -1
    if masked:
        assert torch.equal(removed, ~(rm[:, None] & sm[None, :]))
    else:
        assert not removed.any()
    count = int(out['count'].item())
    assert count == min(K, int((~removed).sum().item()))
    valid = torch.where(removed, torch.full_like(mat, -1.0), mat)
    top_v, top_i = valid.flatten().topk(count)
    assert torch.equal(out['scores'][:count], top_v), 'scores are not the top-K values of the matrix'
    flat = out['ref_idx'][:count] * m + out['src_idx'][:count]
    distinct = top_v.unique().numel() == count
    if distinct:
        assert torch.equal(flat, top_i)
    else:
        assert torch.equal(valid.flatten()[flat], top_v)
    assert torch.all(out['ref_idx'][count:] == -1) and torch.all(out['src_idx'][count:] == -1)
    # the matrix itself against plain torch fp32 on the same device
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        rn, sn = F.normalize(ref, p=2, dim=1), F.normalize(src, p=2, dim=1)
        s = torch.exp(-(2.0 - 2.0 * rn @ sn.T).clamp(min=0.0))
        if masked:
            s = s * (rm[:, None] & sm[None, :])
        if dual:
            s = (s / s.sum(1, keepdim=True)) * (s / s.sum(0, keepdim=True))
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    ok = ~removed
    assert torch.allclose(mat[ok], s[ok], rtol=2e-5, atol=0)


def test_ties_go_to_the_lowest_flat_index_and_runs_are_identical(pkg):
    ref = torch.ones(40, 32, device=DEV)
    src = torch.ones(50, 32, device=DEV)
    out = pkg.ops.superpoint_matching(ref, src, 128, True, normalize=True)
    flat = out['ref_idx'] * 50 + out['src_idx']
    assert torch.equal(flat.cpu(), torch.arange(128))
    assert torch.all(out['scores'] == out['scores'][0])
    g = torch.Generator().manual_seed(5)
    a, b = torch.randn(700, 256, generator=g).to(DEV), torch.randn(900, 256, generator=g).to(DEV)
    first = pkg.ops.superpoint_matching(a, b, 256, True, normalize=True)
    for _ in range(3):
        again = pkg.ops.superpoint_matching(a, b, 256, True, normalize=True)
        for key in ('ref_idx', 'src_idx', 'scores'):
            assert torch.equal(first[key], again[key])


def test_everything_removed_and_argument_errors(pkg):
    ref, src = torch.randn(16, 8, device=DEV), torch.randn(12, 8, device=DEV)
    none = torch.zeros(16, dtype=torch.bool, device=DEV)
    ri, si, sc = pkg.SuperPointMatching(32)(ref, src, none, torch.ones(12, dtype=torch.bool, device=DEV), normalize=True)
    assert ri.numel() == 0 and si.numel() == 0 and sc.numel() == 0
    with pytest.raises(RuntimeError):
        pkg.SuperPointMatching(32)(ref.cpu(), src.cpu())
    with pytest.raises(RuntimeError):
        pkg.ops.superpoint_matching(ref, torch.randn(12, 9, device=DEV), 32)
    with pytest.raises(RuntimeError):
        pkg.ops.superpoint_matching(ref, src, 5000)  # num_corr above the kernel's limit
    with pytest.raises(RuntimeError):
        pkg.ops.superpoint_matching(ref, src, 8, ref_masks=torch.ones(3, dtype=torch.bool, device=DEV))


def test_forward_then_matching_end_to_end(pkg, oracle, golden):
    """The path and its consumer together (model.py:135-157) at a reduced 3DMatch size, against the oracle of both."""
    cfg = dict(pkg.synth.CONFIGS['3dmatch'])
    rp, sp, rf, sf, _ = pkg.synth.make_tie_free_pair('3dmatch', 192, 160, seed=pkg.synth.SEED)
    sd = oracle.init_state_dict(cfg['input_dim'], cfg['output_dim'], cfg['hidden_dim'], cfg['blocks'])
    o0, o1 = oracle.geometric_transformer(sd, cfg, rp, sp, rf, sf)
    exp = oracle.superpoint_matching(o0[0], o1[0], 64, True, normalize=True)
    model = pkg.GeometricTransformer(**cfg).to(DEV).eval()
    model.load_state_dict(sd, strict=True)
    with torch.no_grad():
        g0, g1 = model(rp.to(DEV), sp.to(DEV), rf.to(DEV), sf.to(DEV))
        ri, si, sc = pkg.SuperPointMatching(64, True)(g0[0], g1[0], normalize=True)
    # the features carry the path's half-precision tolerance (2e-2), so: scores close, and every pick is, by the ORACLE's
    # own score matrix, at least as good as the oracle's last pick up to that tolerance
    assert torch.allclose(sc.cpu(), exp[2], rtol=5e-2)
    full = oracle.superpoint_matching_scores(F.normalize(o0[0], p=2, dim=1), F.normalize(o1[0], p=2, dim=1), True)
    assert torch.all(full[ri.cpu(), si.cpu()] >= exp[2][-1] * (1 - 5e-2))
