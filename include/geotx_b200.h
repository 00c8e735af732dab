/* geotx_b200 -- C ABI of the B200-native GeoTransformer geometric-attention hot path.
 *
 * Drop-in boundary.  The reference's own native boundary is a pybind11 torch extension
 * (/root/reference/setup.py:7-22, geotransformer/extensions/pybind.cpp:6-18) whose functions take tensors,
 * validate them with CHECK_* macros (extensions/common/torch_helper.h:6-35), allocate nothing on the caller's
 * behalf beyond their outputs and run on the caller's thread.  This library keeps those conventions but with a
 * plain C signature: device pointers + sizes + a CUDA stream; no torch types.  The caller owns every buffer.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless named host_*; tensors are dense row-major (contiguous);
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); all work is enqueued on it,
 *     nothing synchronises, nothing is allocated;
 *   - return value: 0 on success, a negative GEOTX_E_* code otherwise; geotx_last_error() gives the text
 *     (thread-local).  There is no CPU fallback: a machine without an sm_100 GPU gets GEOTX_E_CUDA.
 *
 * Each entry point names the reference code it replaces (paths relative to /root/reference/geotransformer/).
 */
#ifndef GEOTX_B200_H
#define GEOTX_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GEOTX_OK 0
#define GEOTX_E_ARG (-1)   /* invalid argument (shape, k, alignment, unsupported hidden_dim ...) */
#define GEOTX_E_CUDA (-2)  /* CUDA runtime error (text in geotx_last_error) */

#define GEOTX_MAX_ANGLE_K 7
#define GEOTX_MAX_NEIGHBOR_LIMIT 512 /* geotx_radius_neighbors: neighbours kept per query */
#define GEOTX_MAX_POINT_LIMIT 8192 /* geotx_point_to_node_partition: points kept per node */

/* dtype tags for buffers that may be stored in either precision */
#define GEOTX_F32 0
#define GEOTX_BF16 1
#define GEOTX_F16 2 /* IEEE half: the default storage format of E (11-bit significand, finite-saturated) */

const char* geotx_last_error(void);
int geotx_abi_version(void);
/* No reference counterpart: an asynchronous device-to-device copy (copy engines, no kernel) between two pointers valid in
 * this process, local or peer memory mapped through CUDA IPC.  The multi-GPU gather of the pair outputs
 * (dist.PeerGather) moves its chunks with it, so that no collective kernel competes with the path's kernels for SMs. */
int geotx_copy_async(void* dst, const void* src, int64_t bytes, void* stream);

/* Bytes of caller-provided scratch needed by geotx_gse_knn for B clouds of n points and k neighbours
 * (packed float4 points + k reference vectors per point).  The scratch is an OUTPUT of geotx_gse_knn and an
 * INPUT of geotx_gse_indices / geotx_gse_embed (it holds p and p_knn - p in the kernels' layout). */
int64_t geotx_gse_geometry_bytes(int B, int n, int k);  /* includes a 16-byte trailer: the clouds' largest squared distance */

/* modules/geotransformer/geotransformer.py:38,42 (+ modules/ops/pairwise_distance.py:4-31):
 * knn[b,i,:] = indices of the k+1 smallest sqrt(d2(i,:)), first dropped.  Bit-exact with the reference's
 * torch-CPU fp32 result on tie-free rows; exact ties resolve to the smaller index.
 *   points   f32 [B,n,3]        knn_out  i64 [B,n,k]        geometry  scratch, geotx_gse_geometry_bytes() */
int geotx_gse_knn(const float* points, int B, int n, int k, int64_t* knn_out, void* geometry, void* stream);

/* geotransformer.py:25-55 get_embedding_indices:  d_idx f32 [B,n,n], a_idx f32 [B,n,n,k]. */
int geotx_gse_indices(const void* geometry, int B, int n, int k, float sigma_d, float factor_a,
                      float* d_idx, float* a_idx, void* stream);

/* One-time (per weight update) repack of proj_d / proj_a into the bf16, 128B-swizzled, K-chunked shared-memory
 * image the tcgen05 kernel bulk-copies.  w_d, w_a f32 [C,C] (nn.Linear layout [out,in]); b_d, b_a f32 [C].
 * packed: geotx_gse_packed_weight_bytes(C) bytes. */
int64_t geotx_gse_packed_weight_bytes(int C);
int geotx_gse_pack_weights(const float* w_d, const float* b_d, const float* w_a, const float* b_a, int C,
                           void* packed, void* stream);

/* geotransformer.py:57-72 GeometricStructureEmbedding.forward + positional_embedding.py:18-34:
 *   E[b,i,j,:] = W_d s(d_ij/sigma_d) + b_d + reduce_k (W_a s(angle_ijk*factor_a) + b_a),  reduce = max | mean
 * with the sinusoids generated in registers and contracted on tcgen05 (bf16 operands, fp32 accumulation);
 * the (B,n,n,k,C) intermediate of the reference is never formed.
 *   div_term f32 [C/2] (the module's registered buffer)      out  [B,n,n,C] in out_dtype (GEOTX_F32|GEOTX_BF16)
 * C must be 128 or 256 for this entry point.  row_begin/row_count select a slab of query rows i of every cloud
 * (whole cloud: 0, n); out then has shape [B,row_count,n,C].  Slabs let a caller bound the size of E when
 * n is large (E is n*n*C*2 bytes per cloud) and stream slab by slab through geotx_rpe_scores. */
int geotx_gse_embed(const void* geometry, int B, int n, int k, int row_begin, int row_count,
                    const float* div_term, const void* packed_weights, int C, float sigma_d, float factor_a,
                    int reduce_mean, int out_dtype, void* out, void* stream);

/* geotx_gse_embed that ALSO emits the additive attention scores of the first 'self' block from the accumulator,
 * before E is rounded for storage (rpe_transformer.py:54-56 with proj_p folded, see geotx_rpe_scores):
 *   scores_p[b,h,i,j] = sum_c qt[b,h,i,c] * E[b,i,j,c]      for the slab's rows i, all j
 * This removes one of the block-wise HBM reads of E (the first block's folded query depends only on the input
 * features, not on E, so it is available before E exists).  qt is addressed exactly as in geotx_rpe_scores
 * (element strides over batch, head, row; 16-byte aligned, strides multiples of 4); H <= 4; scores_p f32 [B,H,n,n].
 * Fused when n is a multiple of 128 and k <= 3 (a tile of 128 pairs then shares its query row, whose folded query is
 * staged in shared memory); otherwise the call runs geotx_gse_embed followed by geotx_rpe_scores. */
int geotx_gse_embed_scores(const void* geometry, int B, int n, int k, int row_begin, int row_count,
                           const float* div_term, const void* packed_weights, int C, float sigma_d, float factor_a,
                           int reduce_mean, int out_dtype, void* out, const float* qt, int64_t qt_stride_b,
                           int64_t qt_stride_h, int64_t qt_stride_n, int H, float* scores_p, void* stream);

/* Diagnostic, no reference counterpart: ctas x 8 warps x iters x 16 mma.sync.m16n8k16 (f16 x f16 -> f32, 4096 flop each)
 * on register operands; the caller times the launch.  bench.py uses it to state the peak of the legacy tensor path that
 * K2c / K3r / K4f / K13 run on, next to the measured tcgen05 rate.  out f32 [ctas * 256]. */
int geotx_mma_sync_probe(int ctas, int iters, float* out, void* stream);

/* Debug aid (synchronises the device): the tcgen05 kernel's mbarrier waits are bounded; a wait that times out
 * records {0x80000000 | role<<24 | ring position, barrier smem address, parity, block} and the launch finishes with
 * undefined output instead of hanging the GPU.  Copies the record to host_out4 (all zero = healthy) and clears it. */
int geotx_watchdog_status(uint32_t* host_out4);

/* Same contract, any even C, fp32 CUDA-core arithmetic with accurate sin/cos (no tensor cores): the
 * high-precision path (<=1e-5 of the reference) used for small hidden sizes and as the on-device cross-check. */
int geotx_gse_embed_fp32(const void* geometry, int B, int n, int k, int row_begin, int row_count,
                         const float* div_term,
                         const float* w_d, const float* b_d, const float* w_a, const float* b_a,
                         int C, float sigma_d, float factor_a, int reduce_mean, int out_dtype, void* out,
                         void* stream);

/* modules/transformer/rpe_transformer.py:54-56 with proj_p folded into the query:
 *   scores_p[b,h,i,j] = sum_c qt[b,h,i,c] * E[b,i,j,c],   qt[b,h,i,:] = W_p[h]^T q[b,h,i,:]
 * (the q . b_p term is constant over j and cancels in the softmax).  E is streamed from HBM exactly once.
 *   qt f32, element (b,h,i,c) at qt[b*qt_stride_b + h*qt_stride_h + i*qt_stride_n + c] (strides in elements; a dense
 *      [B,H,n,C] tensor has qt_stride_b = H*n*C, qt_stride_h = n*C, qt_stride_n = C; the strided form lets qt be a
 *      column block of one fused projection GEMM output)
 *   E [B,row_count,m,C] (e_dtype): the slab of query rows [row_begin,row_begin+row_count)
 *   scores_p f32 [B,H,n,m] (only the slab's rows are written)
 * Tensor-core path: e_dtype bf16, C in {128,256}, H <= 8; anything else takes a slower generic kernel. */
int geotx_rpe_scores(const float* qt, int64_t qt_stride_b, int64_t qt_stride_h, int64_t qt_stride_n, const void* E, int e_dtype,
                     int B, int H, int n, int m, int C, int row_begin, int row_count, float* scores_p,
                     void* stream);

/* rpe_transformer.py:57-70 / vanilla_transformer.py:54-68: softmax((q k^T + scores_p) * scale [masked]) v.
 *   q f32 [B,n,H*c] (token-major, heads interleaved as in 'b n (h c)'), k,v f32 [B,m,H*c]; ld_q / ld_k / ld_v are the
 *   token (row) strides in floats (H*c when dense; larger when q,k,v are column blocks of a fused projection output),
 *   scores_p f32 [B,H,n,m] or NULL, key_mask u8 [B,m] or NULL (non-zero = ignore),
 *   out f32 [B,n,H*c].  c (per-head dim) in {16, 32, 64}.
 *   split_keys != 0: when the grid would otherwise be small (n about a thousand gives only 64 (query tile, head)
 *   pairs for 148 SMs) the key range is split over the 2, 4 or 8 CTAs of a thread-block cluster, which merge their
 *   partial softmax states through distributed shared memory (no workspace, no second launch).  split_keys == 0: one
 *   CTA walks all keys of its query tile.  Same result up to fp32 rounding.
 *   Arithmetic: q, k, v and the softmax weights enter the two contractions as IEEE half (11-bit significand, values
 *   saturated to +-65504); accumulation, logits, scores_p, softmax and the output are fp32. */
int geotx_attention(const float* q, const float* k, const float* v, const float* scores_p,
                    const uint8_t* key_mask, int B, int H, int n, int m, int c, int ld_q, int ld_k, int ld_v,
                    float scale, float* out, int split_keys, void* stream);

/* geotx_attention for q, k, v that are already IEEE half (geotx_linear_f16_mixed): same contract and arithmetic (the
 * fp32 entry rounds q, k, v to half itself), ld_q / ld_k / ld_v in halves (multiples of 8), q/k/v 16-byte aligned.
 * K and V tiles and the additive-score tile of the CTA's queries arrive through one 3- or 4-stage cp.async pipeline
 * (one block barrier per 64-key tile, no conversion pass).  scale must be positive. */
int geotx_attention_f16(const void* q, const void* k, const void* v, const float* scores_p, const uint8_t* key_mask,
                        int B, int H, int n, int m, int c, int ld_q, int ld_k, int ld_v, float scale, float* out,
                        int split_keys, void* stream);

/* out = LayerNorm(x + residual) over the last dimension (rpe_transformer.py:103, vanilla_transformer.py:101,
 * output_layer.py:20; torch.nn.LayerNorm semantics: biased variance, eps inside the root).  residual may be NULL.
 *   x, residual, out f32 [rows, C]; gamma, beta f32 [C]; C a multiple of 4, <= 1024. */
int geotx_add_layernorm(const float* x, const float* residual, const float* gamma, const float* beta, int64_t rows,
                        int C, float eps, float* out, void* stream);

/* K13: y = x W^T + b (torch.nn.Linear) for the token-wise projections of the path that are not part of
 * geotx_post_attention's chain: in_proj / out_proj (geotransformer.py:140-141,152-153) and the q | k | v | folded-query
 * projections (rpe_transformer.py:51-53, vanilla_transformer.py:50-52; several Linears of a block are concatenated along
 * N by the caller).  x f32 [T,K] with row stride ld_x (floats), rounded to half on its way into shared memory; w_f16 the
 * PACKED half image of the nn.Linear weight [N][K] (geotx_pack_linear_weight: K chunks of 32, chunk-major, 16-byte
 * units XOR-swizzled; streamed by TMA bulk copies); mma.sync.m16n8k16, fp32 accumulation; y f32 [T,N] with row stride
 * ld_y; bias f32 [N] or NULL; relu != 0 applies max(., 0).  K % 32 == 0, K <= 2048, N % 64 == 0.
 * geotx_f32_to_f16: plain element-wise conversion (same layout). */
int geotx_f32_to_f16(const float* src, void* dst, int64_t count, void* stream);
int geotx_pack_linear_weight(const float* w, int N, int K, void* packed_f16, void* stream);
int geotx_linear_f16(const float* x, int64_t ld_x, int64_t T, int K, const void* w_f16, const float* bias, int N,
                     int relu, float* y, int64_t ld_y, void* stream);
/* The same Linear with a two-part result: output columns [0, half_cols) are written as IEEE half (round to nearest,
 * saturated to +-65504) to y_half [T, half_cols] (row stride ld_yh halves, multiple of 8), columns [half_cols, N) as fp32
 * to y [T, N - half_cols] (y may be NULL when half_cols == N).  half_cols % 64 == 0.  The q | k | v blocks of a fused
 * projection are consumed by geotx_attention_f16 only, which contracts in half anyway: writing them as half removes the
 * attention kernel's conversion pass and halves their traffic, with bit-identical operands. */
int geotx_linear_f16_mixed(const float* x, int64_t ld_x, int64_t T, int K, const void* w_f16, const float* bias, int N,
                           int relu, void* y_half, int64_t ld_yh, int half_cols, float* y, int64_t ld_y, void* stream);

/* K4f: everything a transformer layer does after its attention core, in one kernel (rpe_transformer.py:101-103 /
 * vanilla_transformer.py:99-101 followed by output_layer.py:15-21):
 *   x1  = LayerNorm(hidden W_lin^T + b_lin + x_in; gamma1, beta1, eps1)
 *   out = LayerNorm(x1 + (ReLU(x1 W_expand^T + b_expand)) W_squeeze^T + b_squeeze; gamma2, beta2, eps2)
 * hidden, x_in, out f32 [T,C] with row strides in floats (multiples of 4); C = 128 or 256; weights as the PACKED half
 * images (geotx_pack_linear_weight) of the nn.Linear weights w_lin [C][C], w_expand [2C][C], w_squeeze [C][2C]; biases
 * and LayerNorm parameters f32.
 * A CTA keeps a tile of 32 tokens in shared memory through the whole chain; activations enter the contractions as
 * halves (saturated), accumulation / residuals / LayerNorm are fp32 (torch.nn.LayerNorm semantics). */
int geotx_post_attention(const float* hidden, int64_t ld_hidden, const float* x_in, int64_t ld_x, int64_t T, int C,
                         const void* w_lin_f16, const float* b_lin, const float* gamma1, const float* beta1, float eps1,
                         const void* w_expand_f16, const float* b_expand, const void* w_squeeze_f16,
                         const float* b_squeeze, const float* gamma2, const float* beta2, float eps2, float* out,
                         int64_t ld_out, void* stream);

/* Coarse matching behind the path: F.normalize of both clouds' output features
 * (experiments/<cfg>/model.py:141-142) followed by SuperPointMatching.forward
 * (geotransformer/modules/geotransformer/superpoint_matching.py:13-51, with pairwise_distance(normalized=True),
 * modules/ops/pairwise_distance.py:24-25,30): scores = exp(-max(0, 2 - 2 x.y)); with dual_normalization
 * scores = (s / rowsum) * (s / colsum); the num_corr largest entries of the (n,m) matrix in descending order.
 *   ref_feats f32 [n,d], src_feats f32 [m,d]
 *   ref_keep, src_keep   u8 [n], [m] or NULL: nonzero = keep (the reference's masks: "False if empty"); removed rows and
 *                        columns take no part in the sums or the selection; indices returned are ORIGINAL row numbers
 *   normalize != 0       rows are first divided by max(||row||_2, 1e-12); the unit rows are also written to
 *                        ref_unit_out / src_unit_out (f32 [n,d] / [m,d]) when these are not NULL
 *   ref_idx_out, src_idx_out i64 [num_corr], score_out f32 [num_corr], count_out i32 [1] (all device memory):
 *                        count = min(num_corr, kept_rows * kept_cols) entries are valid, the rest are -1 / 0
 *   score_matrix         optional f32 [n,m]: receives the final score matrix (removed entries hold the bit pattern
 *                        0xFFFFFFFF); NULL keeps it inside the workspace
 *   workspace            geotx_spm_workspace_bytes(n, m, d, num_corr) bytes, 256-byte aligned
 * Equal scores are ordered by ascending flat index i*m + j (torch.topk leaves that order unspecified).
 * num_corr <= 4096; n*m < 2^32.  fp32 arithmetic throughout. */
int64_t geotx_spm_workspace_bytes(int n, int m, int d, int num_corr);
int geotx_superpoint_matching(const float* ref_feats, const float* src_feats, int n, int m, int d,
                              const uint8_t* ref_keep, const uint8_t* src_keep, int normalize, float* ref_unit_out,
                              float* src_unit_out, int dual_normalization, int num_corr, int64_t* ref_idx_out,
                              int64_t* src_idx_out, float* score_out, int32_t* count_out, float* score_matrix,
                              void* workspace, void* stream);

/* Point-to-node partition in front of the path: modules/ops/pointcloud_partition.py:61-107
 * (`point_to_node_partition`, called at experiments/<cfg>/model.py:98-103) plus the patch gather of model.py:105-108.
 * Every point is assigned to its nearest node (pairwise_distance(nodes, points), ops/pairwise_distance.py:22-30, same
 * fp32 operation order; argmin over nodes); every node keeps the point_limit nearest of ITS OWN points, ascending.
 *   points f32 [num_points,3], nodes f32 [num_nodes,3]
 *   point_to_node    i64 [num_points]           index of the nearest node
 *   node_sizes       i64 [num_nodes] or NULL    points assigned to each node (return_count=True)
 *   node_masks       u8  [num_nodes]            1 if the node owns at least one point
 *   node_knn_indices i64 [num_nodes,point_limit] own points by ascending distance, then the sentinel num_points
 *   node_knn_masks   u8  [num_nodes,point_limit] 1 for real entries
 *   node_knn_points  f32 [num_nodes,point_limit,3] or NULL: points[node_knn_indices] with zeros at the sentinel
 *   workspace        geotx_p2n_workspace_bytes(...) bytes, 256-byte aligned
 * Equal distances are ordered by the smaller index (torch.min / torch.topk leave that order unspecified).
 * Like torch.topk, point_limit > num_points is an error.  point_limit <= GEOTX_MAX_POINT_LIMIT. */
int64_t geotx_p2n_workspace_bytes(int num_points, int num_nodes, int point_limit);
int geotx_point_to_node_partition(const float* points, int num_points, const float* nodes, int num_nodes,
                                  int point_limit, int64_t* point_to_node, int64_t* node_sizes, uint8_t* node_masks,
                                  int64_t* node_knn_indices, uint8_t* node_knn_masks, float* node_knn_points,
                                  void* workspace, void* stream);

/* Allocation-free variant of geotx_point_to_node_partition.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same inputs, outputs, arithmetic and results as geotx_point_to_node_partition, but
 * the workspace is a file-scope `static __device__` array of GEOTX_P2N_STATIC_SCRATCH_BYTES inside the library (every
 * GPU that loads the library holds it until the process exits); the call allocates nothing (no cudaMalloc / cudaFree /
 * allocating library routine) and takes no workspace argument.  Sizes whose workspace
 * (geotx_p2n_workspace_bytes(num_points, num_nodes, point_limit)) exceeds GEOTX_P2N_STATIC_SCRATCH_BYTES are rejected
 * with GEOTX_E_ARG -- e.g. 65 536 points x 4096 nodes (1.36 MB) fits, 131 072 points do not.
 * All calls on a device share the array, which keeps the last call's contents (each call clears the counters it needs):
 * calls on ONE stream are ordered and safe; calls that overlap on the GPU (different streams, or this entry point
 * replayed from two graphs at once) RACE and must be serialised by the caller.  The device must be current. */
#define GEOTX_P2N_STATIC_SCRATCH_BYTES (2 * 1024 * 1024)
int64_t geotx_p2n_static_scratch_bytes(void);
int geotx_point_to_node_partition_static(const float* points, int num_points, const float* nodes, int num_nodes,
                                         int point_limit, int64_t* point_to_node, int64_t* node_sizes,
                                         uint8_t* node_masks, int64_t* node_knn_indices, uint8_t* node_knn_masks,
                                         float* node_knn_points, void* stream);

/* Optimal transport behind the coarse matching: modules/sinkhorn/learnable_sinkhorn.py:13-66
 * (`LearnableLogOptimalTransport.forward`, called at experiments/<cfg>/model.py:187-189): scores padded with a dustbin
 * row and column holding alpha, masked rows / columns filled with -inf_value, num_iterations log-domain Sinkhorn
 * iterations (two logsumexp each), result (scores + u + v) - norm.  One CTA per problem; the padded matrix stays in
 * shared memory for the whole loop.
 *   scores f32 [B,M,N]; row_masks u8 [B,M] / col_masks u8 [B,N] or NULL, nonzero = VALID (the reference's masks);
 *   alpha f32 [1] (device: the module's parameter); out f32 [B,M+1,N+1].
 * (M+1) x (N+1) must fit shared memory (about 230 x 230).  fp32 arithmetic in the reference's operation order. */
int geotx_log_optimal_transport(const float* scores, const uint8_t* row_masks, const uint8_t* col_masks,
                                const float* alpha, int num_iterations, float inf_value, int B, int M, int N,
                                float* out, void* stream);

/* Weighted Procrustes / Kabsch with the 3x3 SVD on the device: modules/registration/procrustes.py:6-73
 * (`weighted_procrustes`, called from modules/geotransformer/local_global_registration.py:169,181-191).  The reference
 * moves the covariances to the host for torch.svd (procrustes.py:53-54); this entry point never leaves the device.
 *   src_points, ref_points f32 [B,N,3]; weights f32 [B,N] or NULL (all ones)
 *   weights below weight_thresh count as 0; the rest are divided by (their sum + eps)
 *   transform_out f32 [B,4,4] row-major [R t; 0 0 0 1] with R, t mapping src onto ref
 * (return_transform=False of the reference is R = T[:3,:3], t = T[:3,3]). */
int geotx_weighted_procrustes(const float* src_points, const float* ref_points, const float* weights, int B, int N,
                              float weight_thresh, float eps, float* transform_out, void* stream);

/* Radius neighbours in stack mode: extensions/cpu/radius_neighbors/radius_neighbors_cpu.cpp:3-104
 * (exported as ext.radius_neighbors, extensions/pybind.cpp) followed by the truncation of modules/ops/radius_search.py:
 * 24-27; called 3 times per pyramid stage from utils/data.py:33-72.  For every query point: the support points of the
 * SAME batch element with ((qx-sx)^2 + (qy-sy)^2) + (qz-sz)^2 < radius^2 (fp32, the reference's own arithmetic),
 * nearest first, at most `limit`, then the sentinel sum(s_lengths).
 *   q_points f32 [sum(q_lengths),3], s_points f32 [sum(s_lengths),3] (device; batch elements stacked)
 *   host_q_lengths, host_s_lengths i64 [batch] (HOST memory, as in the reference where they are CPU tensors)
 *   neighbors i64 [sum(q_lengths), limit]   indices into the stacked s_points
 *   counts    i32 [sum(q_lengths)]          in-radius support points of every query BEFORE truncation
 *   max_count i32 [1]                       their maximum = the width of the reference's untruncated result; the
 *                                           reference returns min(max_count, limit) columns
 * Equal distances are ordered by the smaller index (std::sort leaves ties unspecified).  limit <= 512. */
int geotx_radius_neighbors(const float* q_points, const float* s_points, const int64_t* host_q_lengths,
                           const int64_t* host_s_lengths, int batch, float radius, int limit, int64_t* neighbors,
                           int32_t* counts, int32_t* max_count, void* stream);

/* The same search through per-element cell lists (uniform grid of cells of 1.001 x radius, 27 cells per query) instead
 * of the exhaustive scan: identical results (same distance arithmetic, keys and selection), ~10-50x less work at the
 * pyramid's sizes.  workspace: geotx_radius_neighbors_workspace_bytes(max_support_points, max_cells) bytes, 256-byte
 * aligned; max_support_points >= sum(host_s_lengths); max_cells is per batch element (8 grids are kept, for the 8
 * elements one set of launches searches): when an element's grid would exceed it the cells are made larger (still
 * exact, only slower).  radius > 0. */
int64_t geotx_radius_neighbors_workspace_bytes(int64_t max_support_points, int64_t max_cells);
int geotx_radius_neighbors_grid(const float* q_points, const float* s_points, const int64_t* host_q_lengths,
                                const int64_t* host_s_lengths, int batch, float radius, int limit, int64_t* neighbors,
                                int32_t* counts, int32_t* max_count, void* workspace, int64_t max_support_points,
                                int64_t max_cells, void* stream);

/* Grid (voxel barycentre) subsampling in stack mode: extensions/cpu/grid_subsampling/
 * grid_subsampling_cpu.cpp:3-79 (ext.grid_subsampling; modules/ops/grid_subsample.py:9-22; utils/data.py:22-30).
 * Per batch element: voxels of size voxel_size anchored at floor(min_corner / voxel) * voxel (the reference's fp32
 * sequence, bit for bit), every occupied voxel replaced by the barycentre of its points.
 *   points f32 [sum(lengths),3] (device), host_lengths i64 [batch] (HOST; every element needs >= 1 point)
 *   s_points  f32 [sum(lengths),3] capacity; the first sum(s_lengths) rows are written, batch elements stacked
 *   s_lengths i64 [batch] (device): occupied voxels per element, or -1 if the element's bounding grid has more than
 *             max_cells cells (nothing is written for it)
 *   workspace geotx_grid_subsample_workspace_bytes(max_cells) bytes, 256-byte aligned; max_cells <= 2^31
 * Output order: ascending voxel index (x fastest) inside every element; the reference's order is that of a
 * std::unordered_map.  Barycentres are summed in fp64 and rounded once (the reference: fp32, input order). */
int64_t geotx_grid_subsample_workspace_bytes(int64_t max_cells);
int geotx_grid_subsample(const float* points, const int64_t* host_lengths, int batch, float voxel_size, float* s_points,
                         int64_t* s_lengths, void* workspace, int64_t max_cells, void* stream);

/* K11 LocalGlobalRegistration on the device.  Replaces
 * geotransformer/modules/geotransformer/local_global_registration.py:49-235 (use_dustbin = False,
 * correspondence_limit = None, as in every shipped config).
 *   ref_knn_points, src_knn_points  (P, K, 3) f32 patch points        ref_knn_masks, src_knn_masks  (P, K) u8
 *   score_mat        f32 log-likelihoods, element (b, i, j) at b*score_batch_stride + i*score_row_stride + j (so the
 *                    reference's matching_scores[:, :-1, :-1] view is passed without a copy)
 *   global_scores    (P,) f32 or NULL (use_global_score = False)
 *   topk             the module's k (<= 8), K <= 128
 *   outputs          ref_corr_points / src_corr_points (cap, 3) f32 and corr_scores (cap,) f32 with
 *                    cap = 2 * P * K * topk, filled in torch.nonzero order (patch, i, j) for count_out[0] entries;
 *                    transform_out 16 f32 (row-major 4x4)
 *   workspace        geotx_lgr_workspace_bytes(P, K) bytes, 256-byte aligned
 * No host synchronisation; equal scores inside a top-k are ordered by index. */
int64_t geotx_lgr_workspace_bytes(int num_patches, int K);
int geotx_local_global_registration(const float* ref_knn_points, const float* src_knn_points,
                                    const uint8_t* ref_knn_masks, const uint8_t* src_knn_masks, const float* score_mat,
                                    int64_t score_batch_stride, int64_t score_row_stride, const float* global_scores,
                                    int num_patches, int K, int topk, float acceptance_radius, int mutual,
                                    float confidence_threshold, int correspondence_threshold, int num_refinement_steps,
                                    float* ref_corr_points, float* src_corr_points, float* corr_scores,
                                    int32_t* count_out, float* transform_out, void* workspace, void* stream);

/* K12 patch matching scores: the glue between coarse matching and optimal transport.  Replaces
 * experiments/<any>/model.py:169-170,176-179,187-188 (zero-row padding of the fine features, two (P,K,C) gathers,
 * einsum('bnd,bmd->bnm') and the division by sqrt(C)).
 *   ref_feats (n_ref, C), src_feats (n_src, C) f32, C % 4 == 0
 *   ref_node_knn_indices (M_ref, K), src_node_knn_indices (M_src, K) int64 as point_to_node_partition returns them
 *                    (n_ref / n_src marks a padded slot: its feature row is zero); K <= 128
 *   ref_node_corr_indices, src_node_corr_indices (P,) int64 node correspondences (superpoint matching)
 *   scores           (P, K, K) f32 out */
int geotx_patch_scores(const float* ref_feats, int64_t n_ref, const float* src_feats, int64_t n_src, int C,
                       const int64_t* ref_node_knn_indices, const int64_t* src_node_knn_indices, int K,
                       const int64_t* ref_node_corr_indices, const int64_t* src_node_corr_indices, int P,
                       float* scores, void* stream);

/* K2c, the default embedding kernel: geotransformer.py:57-72 + positional_embedding.py:18-34 as a LOW-RANK contraction.
 * Per output channel, proj_x(sinus(x)) is a band-limited function of one scalar on a bounded interval, so it has a
 * short Chebyshev expansion f_c(x) = sum_{m<K} G[m,c] T_m(2x/L - 1): K = 16 * ka_tiles for the angle term (L = pi *
 * factor_a: 16 coefficients up to 13 index units, 32 up to 37) and K_d = 16 .. 128 for the distance term, in six
 * prebuilt sets for ranges up to geotx_gse_cheb_d_range(0..5) = 13, 37, 64, 93, 150, 210 index units.
 *   geotx_gse_cheb_ka_tiles(pi * factor_a)   1 or 2; 0 = the angle range is not covered (use geotx_gse_embed)
 *   geotx_gse_build_cheb   once per weight set: G in fp64 (Gauss-Chebyshev quadrature of the module's own weights and
 *                          div_term), rounded to half, in the kernel's register-image layout (geotx_gse_cheb_bytes bytes);
 *                          stats f32 [7][4] (device; set 0 = angle, 1..6 = distance sets): {largest |expansion - f| found
 *                          at 256 check points per channel, largest |f|, largest |G|, 0} -- the caller's accuracy guard;
 *                          plain (optional, else NULL): f32 [geotx_gse_cheb_plain_rows(ka_tiles)][C], the unrounded G of
 *                          the angle set (16 ka_tiles rows) and of the six distance sets (16, 32, 48, 64, 96, 128 rows),
 *                          row-major, for host-side folds (see geotx_rpe_scores_cheb: u)
 *   geotx_gse_embed_cheb   E[b,i,j,:] (f16 or bf16) for a slab of query rows, any k <= GEOTX_MAX_ANGLE_K, 'max' or 'mean'.
 *                          The basis values are generated per pair in registers (three-term recurrence, fp32), stored as
 *                          halves in shared memory and contracted with mma.sync.m16n8k16 (fp32 accumulation); the max
 *                          over k and the sum with the distance term happen in the accumulator registers.
 *                          The kernel selects the distance set from the clouds' largest distance, which geotx_gse_knn
 *                          leaves in the geometry scratch; max_d_set caps the choice (it sizes the shared memory).
 *                          *overflow is set to 1 (never cleared) if an argument fell outside its range (distance beyond
 *                          set max_d_set; the value is then clamped to the end of the range). */
int geotx_gse_cheb_ka_tiles(float angle_index_max);
float geotx_gse_cheb_d_range(int d_set);
int64_t geotx_gse_cheb_bytes(int C, int ka_tiles);
int64_t geotx_gse_cheb_plain_rows(int ka_tiles);
int geotx_gse_build_cheb(const float* div_term, const float* w_d, const float* b_d, const float* w_a, const float* b_a,
                         int C, float angle_index_max, int ka_tiles, void* coef, float* stats, float* plain,
                         void* stream);
int geotx_gse_embed_cheb(const void* geometry, int B, int n, int k, int row_begin, int row_count, const void* coef,
                         int C, float angle_index_max, int ka_tiles, int max_d_set, float sigma_d, float factor_a,
                         int reduce_mean, int out_dtype, void* out, int32_t* overflow, void* stream);

/* K3r: geotransformer.py:57-72 fused with rpe_transformer.py:54-56 (proj_p folded into the query, see geotx_rpe_scores):
 *   scores_p[b,h,i,j] = sum_c qt[b,h,i,c] * E[b,i,j,c]      for the slab's rows i, all j
 * WITHOUT E ever being written to or read from memory: the kernel of geotx_gse_embed_cheb rebuilds E for 128 keys of one
 * query row at a time in its accumulator registers (rounded to half, exactly what a stored half E holds), feeds them
 * back as the A operand of a second mma.sync contraction with that row's folded query (hi + lo halves, as in
 * geotx_rpe_scores), adds the 8 channel blocks' partial scores in shared memory and writes 512-byte rows of scores_p.
 * One launch per 'self' block and cloud replaces one embedding pass (n^2 C 2 B written) plus one HBM read of E per block.
 *   qt f32 addressed as in geotx_rpe_scores (16-byte aligned, strides multiples of 4); H <= 8; scores_p f32 [B,H,n,n]
 *   (only the slab's rows are written); coef / angle_index_max / ka_tiles / max_d_set / overflow as in geotx_gse_embed_cheb.
 *   u (optional, else NULL): the DISTANCE SPLIT.  The distance term of E is linear in its Chebyshev basis, so its share
 *   of the scores is sum_m u[b,h,i,m] T_m(d_ij) with u[b,h,i,:] = G_d qt[b,h,i,:] -- a K_d-term product per (h,i,j)
 *   instead of a C-term one, and one third fewer MMAs in the kernel (the embedding contraction then covers the angle
 *   slots only).  u f32, element (b,h,i,m) at u[b*u_stride_b + h*u_stride_h + i*u_stride_n + m]; per (b,h,i) the 32
 *   coefficients of distance set 1 (G rows of `plain`), then the 48 of set 2 (80 values: max_d_set <= 2), then the 64 of
 *   set 3 (144 values: REQUIRED when max_d_set >= 3); 8-byte aligned, even strides.  The caller obtains u from the same
 *   projection that yields qt (G_d folded into the weights).  With u the kernel runs its split instances (H <= 4,
 *   k = 3, 'max', clouds up to 93 index units: a consumer warp owns 16 keys, walks the channel blocks, keeps the scores
 *   in registers and writes scores_p itself); other shapes ignore u (distance term inside E as without it). */
int geotx_rpe_scores_cheb(const void* geometry, int B, int n, int k, int row_begin, int row_count, const void* coef,
                          int C, float angle_index_max, int ka_tiles, int max_d_set, float sigma_d, float factor_a,
                          int reduce_mean, const float* qt, int64_t qt_stride_b, int64_t qt_stride_h, int64_t qt_stride_n,
                          int H, const float* u, int64_t u_stride_b, int64_t u_stride_h, int64_t u_stride_n,
                          float* scores_p, int32_t* overflow, void* stream);

/* K2t (superseded by K2c, kept as embed_kernel='table'): the embedding by tabulation.  proj_x(sinus(x)) + bias is, per
 * channel, a function of one scalar; geotx_gse_build_tables samples the angle term every h_a index units (row r at
 * x = r h_a, rows_a >= pi*factor_a/h_a + 2; read by linear interpolation) and the distance term every h_d (row r at
 * x = (r-1) h_d, rows_d >= max d_index/h_d + 4; cubic), as halves (geotx_gse_table_bytes bytes), and
 * geotx_gse_embed_table interpolates them out of shared memory instead of contracting on the tensor cores.  Same geometry scratch and E layout (f16) as geotx_gse_embed; k = 3, C = 128 or 256, reduction 'max'.
 * A distance beyond the table is evaluated exactly from div_term, w_d, b_d (slow, meant to be rare); *overflow is set to
 * 1 only if an ANGLE argument fell outside its table (cannot happen for rows_a as above); never cleared. */
int64_t geotx_gse_table_bytes(int C, int rows_a, int rows_d);
int geotx_gse_build_tables(const float* div_term, const float* w_d, const float* b_d, const float* w_a,
                           const float* b_a, int C, float h_a, float h_d, int rows_a, int rows_d, void* tables,
                           void* stream);
int geotx_gse_embed_table(const void* geometry, int B, int n, int k, int row_begin, int row_count, const void* tables,
                          int C, int rows_a, int rows_d, float h_a, float h_d, float sigma_d, float factor_a,
                          const float* div_term, const float* w_d, const float* b_d, void* out_f16, int32_t* overflow,
                          void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GEOTX_B200_H */
