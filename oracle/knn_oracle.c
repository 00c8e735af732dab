/* Plain-C bit-exact restatement of the integer/index part of the hot path.
 *
 * TEST INFRASTRUCTURE ONLY (never linked into the product library).
 *
 * Follows /root/reference/geotransformer/modules/ops/pairwise_distance.py:22-30 and
 * /root/reference/geotransformer/modules/geotransformer/geotransformer.py:38-42 as
 * executed by torch-CPU fp32 (measured recipe):
 *     xy  = fma(x2, y2, fma(x1, y1, x0*y0))          (MKL sgemm, K=3)
 *     sq  = (x0*x0 + x1*x1) + x2*x2                  (each square rounded, no fma)
 *     d2  = max(0, (sq_i - 2*xy) + sq_j)             (left to right)
 *     d   = sqrtf(d2);  d_idx = d / sigma_d      (IEEE; torch-CPU's vectorised sqrt is NOT correctly
 *           rounded: measured 0.57 % of entries are 1 ulp off IEEE sqrtf, never more, and it stays
 *           monotone -- so indices are compared bit-exactly, d_idx to <= 1 ulp of d)
 *     knn = the k+1 smallest of row i by (d2, index), first one dropped
 * Ties: torch.topk leaves tie order unspecified; this file (and the CUDA kernel)
 * break ties by the smaller index.  Parity is only defined on tie-free inputs
 * (tests assert that).
 *
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared knn_oracle.c -o _build/libknn_oracle.so -lm
 */
#include <math.h>
#include <stdint.h>

static float sq_norm(const float* p) {
    volatile float a = p[0] * p[0];
    volatile float b = p[1] * p[1];
    volatile float c = p[2] * p[2];
    volatile float ab = a + b;
    return ab + c;
}

float geotx_oracle_sqdist(const float* pi, const float* pj) {
    volatile float m0 = pi[0] * pj[0];
    float xy = fmaf(pi[2], pj[2], fmaf(pi[1], pj[1], m0));
    volatile float t = sq_norm(pi) - 2.0f * xy;
    volatile float d2 = t + sq_norm(pj);
    return d2 < 0.0f ? 0.0f : d2;
}

/* points: n*3 floats; knn_out: n*k int64; d_idx_out: n*n floats or NULL */
void geotx_oracle_knn(const float* points, int n, int k, float sigma_d, int64_t* knn_out, float* d_idx_out) {
    int kk = k + 1;
    for (int i = 0; i < n; ++i) {
        float bv[16];
        int bi[16];
        int cnt = 0;
        for (int j = 0; j < n; ++j) {
            /* selection key is the clamped d2: sqrt is monotone, so wherever the reference's order over
             * sqrt(d2) is strict (tie-free rows) the two orders coincide */
            float d = geotx_oracle_sqdist(points + 3 * i, points + 3 * j);
            if (d_idx_out) d_idx_out[(int64_t)i * n + j] = sqrtf(d) / sigma_d;
            /* insertion into the sorted list of the kk smallest, strict < keeps the smaller index first */
            int pos = cnt;
            while (pos > 0 && d < bv[pos - 1]) --pos;
            if (pos < kk) {
                int last = cnt < kk ? cnt : kk - 1;
                for (int q = last; q > pos; --q) { bv[q] = bv[q - 1]; bi[q] = bi[q - 1]; }
                bv[pos] = d; bi[pos] = j;
                if (cnt < kk) ++cnt;
            }
        }
        for (int q = 0; q < k; ++q) knn_out[(int64_t)i * k + q] = bi[q + 1];
    }
}

/* SURVEY.md 8(f)-2 -- /root/reference/geotransformer/modules/ops/pointcloud_partition.py:82-98, torch-independent:
 * d2(node, point) with x = node, y = point (pairwise_distance(nodes, points)); every point goes to the node with the
 * smallest d2 (first one on ties); every node lists its own points by ascending (d2, index), at most point_limit of
 * them, then the sentinel num_points.
 *   point_to_node: N int64; node_sizes: M int64; knn_indices: M*K int64; knn_masks: M*K bytes; best_d2: N floats scratch */
void geotx_oracle_point_to_node(const float* points, int N, const float* nodes, int M, int K, int64_t* point_to_node,
                                int64_t* node_sizes, int64_t* knn_indices, uint8_t* knn_masks, float* best_d2) {
    for (int m = 0; m < M; ++m) node_sizes[m] = 0;
    for (int p = 0; p < N; ++p) {
        float best = INFINITY;
        int bi = 0;
        for (int m = 0; m < M; ++m) {
            float d = geotx_oracle_sqdist(nodes + 3 * m, points + 3 * p);
            if (d < best) { best = d; bi = m; }
        }
        point_to_node[p] = bi;
        best_d2[p] = best;
        node_sizes[bi]++;
    }
    for (int m = 0; m < M; ++m) {
        int cnt = 0;
        int64_t* idx = knn_indices + (int64_t)m * K;
        for (int p = 0; p < N; ++p) {
            if (point_to_node[p] != m) continue;
            /* insertion into the sorted list of the K smallest; p ascends, strict < keeps the smaller index first */
            int pos = cnt;
            while (pos > 0 && best_d2[p] < best_d2[idx[pos - 1]]) --pos;
            if (pos < K) {
                int last = cnt < K ? cnt : K - 1;
                for (int q = last; q > pos; --q) idx[q] = idx[q - 1];
                idx[pos] = p;
                if (cnt < K) ++cnt;
            }
        }
        for (int q = 0; q < K; ++q) {
            knn_masks[(int64_t)m * K + q] = q < cnt;
            if (q >= cnt) idx[q] = N;
        }
    }
}
