"""The stages after the transformer, chained as experiments/*/model.py:98-108,141-205 chains them: point-to-node
partition (K6) of both clouds, coarse matching (K5), the patch gathers and patch scores (K12), optimal transport (K7) and
local-to-global registration (K11).  The glue below is model.py's own (index_select / einsum); the CPU chain runs the
oracle restatements (each pinned on reference-written golden vectors in its own test file), the GPU chain the CUDA
kernels through the drop-in modules.  Near-ties may pick different node correspondences on the two sides, so the chains
are compared on what the pipeline is for: the estimated transform (<= 5e-3 between the chains, <= 2e-2 from the
planted motion) and the size of the correspondence set (within 5 %)."""
import pytest
import torch
import torch.nn.functional as F

NUM_CORR, POINT_LIMIT, TOPK, RADIUS = 128, 32, 3, 0.1


@torch.no_grad()
def run_tail(d, p2n, spm, ot, lgr, patch_scores=None):
    """model.py:98-108 and :141-205 with the four stage functions passed in (inference: engine/single_tester.py:42-43
    disables grad, and the B200 modules are forward-only)."""
    _, ref_node_masks, ref_knn_idx, ref_knn_masks = p2n(d['ref_points_f'], d['ref_points_c'], POINT_LIMIT)
    _, src_node_masks, src_knn_idx, src_knn_masks = p2n(d['src_points_f'], d['src_points_c'], POINT_LIMIT)
    pad = lambda x: torch.cat([x, torch.zeros_like(x[:1])], dim=0)  # noqa: E731
    ref_knn_points = pad(d['ref_points_f'])[ref_knn_idx]
    src_knn_points = pad(d['src_points_f'])[src_knn_idx]
    ref_idx, src_idx, node_scores = spm(F.normalize(d['ref_feats_c'], p=2, dim=1),
                                        F.normalize(d['src_feats_c'], p=2, dim=1), ref_node_masks, src_node_masks)
    if patch_scores is None:  # model.py:169-170,176-179,187-188 as written
        ref_feats = pad(d['ref_feats_f'])[ref_knn_idx[ref_idx]]
        src_feats = pad(d['src_feats_f'])[src_knn_idx[src_idx]]
        scores = torch.einsum('bnd,bmd->bnm', ref_feats, src_feats) / d['ref_feats_f'].shape[1] ** 0.5
    else:  # K12
        scores = patch_scores(d['ref_feats_f'], d['src_feats_f'], ref_knn_idx, src_knn_idx, ref_idx, src_idx)
    rm, sm = ref_knn_masks[ref_idx], src_knn_masks[src_idx]
    plan = ot(scores, rm, sm)
    return lgr(ref_knn_points[ref_idx], src_knn_points[src_idx], rm, sm, plan[:, :-1, :-1], node_scores)


def oracle_tail(oracle, d):
    return run_tail(
        d, oracle.point_to_node_partition,
        lambda a, b, ma, mb: oracle.superpoint_matching(a, b, NUM_CORR, True, ma, mb),
        lambda s, rm, sm: oracle.log_optimal_transport(s, 1.0, 100, rm, sm),
        lambda *a: oracle.local_global_registration(*a, k=TOPK, acceptance_radius=RADIUS))


def test_oracle_chain_recovers_the_motion(pkg, oracle):
    d = pkg.synth.make_tail_inputs(num_nodes=128, points_per_node=16, seed=11)
    rc, sc, s, T = oracle_tail(oracle, d)
    assert rc.shape[0] > 200
    assert float((T - d['transform']).abs().max()) < 2e-2


@pytest.mark.gpu
@pytest.mark.parametrize('num_nodes,seed', [(128, 11), (512, 12)])
def test_gpu_chain_vs_oracle_chain(pkg, oracle, num_nodes, seed):
    from gpu_common import DEV
    d = pkg.synth.make_tail_inputs(num_nodes=num_nodes, points_per_node=16, seed=seed)
    want = oracle_tail(oracle, d)
    g = {k: v.to(DEV) for k, v in d.items()}
    got = run_tail(g, pkg.point_to_node_partition, pkg.SuperPointMatching(NUM_CORR, True),
                   pkg.LearnableLogOptimalTransport(100).to(DEV), pkg.LocalGlobalRegistration(TOPK, RADIUS),
                   pkg.patch_matching_scores)
    T = got[3].cpu()
    assert float((T - d['transform']).abs().max()) < 2e-2
    assert float((T - want[3]).abs().max()) < 5e-3
    assert abs(got[0].shape[0] - want[0].shape[0]) <= 0.05 * want[0].shape[0]
    inl = (torch.linalg.norm(got[0] - (got[1] @ got[3][:3, :3].T + got[3][:3, 3]), dim=1) < RADIUS).float().mean()
    assert float(inl) > 0.8
