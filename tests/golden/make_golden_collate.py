"""Golden vectors for SURVEY.md 8(f)-4 (radius_neighbors), written by the reference's own C++ operator compiled from
/root/reference (oracle/build_ref.sh -> oracle/_ref/libgeotx_ref_cpu.so).  Runs only in the build container:

    python tests/golden/make_golden_collate.py

Cases (inputs regenerated from seeds by synth.make_dense_cloud; stored are the reference's neighbour lists truncated
to `limit` columns, as modules/ops/radius_search.py:24-27 does, plus the untruncated width):
  rn_small   two clouds of 700 / 500 points, self search, radius 0.25, limit 10 (most lists are cut)
  rn_cross   queries = every 4th point (a "subsampling" search, utils/data.py:49-57), radius 0.3, limit 16
  rn_sparse  radius so small that most queries only find themselves; limit 8 (> max_count: narrower result)
  rn_stage   a pyramid-stage-sized search: 8000 + 7000 points, radius 0.125, limit 12 (about half of the lists are cut)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import pkgpath  # noqa: E402
import ref_cpu  # noqa: E402

CASES = {
    # tag: (sizes, extent, seed, query stride, radius, limit)
    'rn_small': ((700, 500), 3.0, 41, 1, 0.25, 10),
    'rn_cross': ((900, 800), 3.0, 42, 4, 0.3, 16),
    'rn_sparse': ((400, 300), 3.0, 43, 1, 0.02, 8),
    'rn_stage': ((8000, 7000), 3.0, 44, 1, 0.125, 12),
}


def case_inputs(synth, tag):
    import torch
    sizes, extent, seed, stride, radius, limit = CASES[tag]
    clouds = [synth.make_dense_cloud(n, extent, seed + 100 * b) for b, n in enumerate(sizes)]
    queries = [c[::stride].contiguous() for c in clouds]
    return (torch.cat(queries), torch.cat(clouds), [q.shape[0] for q in queries], [c.shape[0] for c in clouds],
            radius, limit)


def main():
    synth = pkgpath.load_package().synth
    out = {}
    for tag in CASES:
        q, s, ql, sl, radius, limit = case_inputs(synth, tag)
        full = ref_cpu.radius_neighbors(q.numpy(), s.numpy(), ql, sl, radius)
        out[f'{tag}.neighbors'] = full[:, :limit].astype(np.int32)
        out[f'{tag}.max_count'] = np.int64(full.shape[1])
        print(tag, full.shape, 'kept', out[f'{tag}.neighbors'].shape)
    np.savez_compressed(os.path.join(HERE, 'collate_cases.npz'), **out)
    print(os.path.getsize(os.path.join(HERE, 'collate_cases.npz')))


if __name__ == '__main__':
    main()
