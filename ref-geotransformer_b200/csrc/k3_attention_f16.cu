// K3h: K3b (k3_attention.cu) for q, k, v that arrive as IEEE half -- the form K13 writes them in for the fused schedule
// (geotx_linear_f16_mixed).  Same arithmetic as K3b (which rounds its fp32 inputs to half before the two contractions):
//   out = softmax((q k^T + scores_p) * scale  [key_mask -> -inf]) v
// rpe_transformer.py:57-70 (self blocks, scores_p from K3r) / vanilla_transformer.py:54-68 (cross blocks, no scores_p).
//
// What changed against K3b, and why.  K3b's step over one 64-key tile was: wait for the raw fp32 K/V tile, barrier,
// convert it to half in a second buffer, barrier, Q K^T, softmax, P V -- with the additive scores of the tile loaded
// straight from global memory into registers one tile ahead.  Measured inside the pipelined step (tools/ablate_step.py)
// the attention launches cost 17 % of the step at every cloud size while running at 23 % of the mma.sync peak.  Here
//   * K and V tiles are copied as they are (cp.async, 16 bytes) into padded half tiles: no conversion, half the bytes;
//   * the additive scores of the CTA's queries x the tile's keys travel through the SAME rings into a padded fp32
//     tile, and seed the accumulators of Q K^T (C operand) instead of being added afterwards;
//   * two 3-slot rings (K + scores + mask bias | V; 6 slots without scores): the copies of a tile are requested two
//     steps before its Q K^T, ONE block barrier per tile, the per-thread copy pattern worked out once outside the loop
//     (the first version spent 420 of its 850 instructions per tile on copy addressing);
//   * software pipelining inside the warp: a step issues Q K^T of tile t+1 and then softmax + P V of tile t, two
//     independent instruction streams in one basic block -- ptxas interleaves the HMMAs with the MUFU.EX2 / FFMA /
//     FMNMX of the softmax (two tiles' logits live: ~240 registers, two 4-warp CTAs per SM);
//   * the softmax works on raw logits (max first, then one FFMA + EX2 per weight).
// Stand-alone (tools/att_time.py, graph replays, us per launch, with / without scores), H = 4, c = 64:
//            n = 1024            n = 2048         n = 4096
//   K3b      29.4 / 25.8         58.0 / 49.1      166.9 / 118.6      (one CTA per query tile walks all keys)
//   K3h      16.1 / 13.0         30.3 / 23.8       80.6 /  64.9      (213 / 265 TFLOP/s at n = 4096)
//   K3h, keys split over clusters (one forward at a time): 11.1 / 10.3 at n = 1024.
// ncu at n = 1024 without key split (64 CTAs, one warp per scheduler): 493 instructions per warp and tile at 0.27 per
// clock -- fixed-latency dependency stalls of a lone warp; 2.4 us of the launch are fixed, 0.87 us goes per tile.
// Not kept (profiles/r02_k3h_notes.txt): deeper rings (2, 3 and 5 slots time the same), two warp groups per CTA on
// alternate tiles (15.3 us against 16.0, slower inside the step), the 128-query / 8-warp CTA shape.
// Key ranges are split over a thread-block cluster exactly as in K3b (partials merged through distributed shared memory).
#include <cstdlib>
#include <cuda_fp16.h>
#include "geotx_common.cuh"
#include "sm100_ptx.cuh"
#include "host_util.h"

namespace geotx {

constexpr int AH_BK = 64;        // keys per tile
constexpr int AH_SPITCH = 72;    // floats per row of the staged score tile: 8-byte fragment reads are conflict-free
constexpr int AH_MAX_SPLITS = 8;
#ifndef AH_SLOTS_SCORES
#define AH_SLOTS_SCORES 3  // ring depth with a score tile per slot (27.9 KB + 9.2 KB per slot at c = 64, 64 queries)
#endif
#ifndef AH_SLOTS_PLAIN
#define AH_SLOTS_PLAIN 6   // ring depth without scores (9.5 KB + 9.2 KB per slot)
#endif

__device__ __forceinline__ void ah_mma(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ah_ldsm4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ah_ldsm2(uint32_t addr, uint32_t& r0, uint32_t& r1) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}
__device__ __forceinline__ void ah_ldsm4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ float ah_exp2(float x) {  // one MUFU.EX2; arguments are <= 0, -inf -> 0
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// async copies global -> shared; src_bytes = 0 zero-fills (keys past the end of the cloud)
__device__ __forceinline__ void ah_cp16(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void ah_cp4(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void ah_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void ah_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int DH, int WARPS, bool SCORES>
struct AhSmem {
    static constexpr int BQ = 16 * WARPS;
    static constexpr int HP = DH + 8;  // halves per K / V row: 16-byte aligned, pitch / 16 B odd -> ldmatrix conflict-free
    static constexpr int KV_BYTES = AH_BK * HP * 2;
    static constexpr int S_BYTES = SCORES ? BQ * AH_SPITCH * 4 : 0;
    // K ring slot: K tile | score tile | mask bias; SLOTS of them, then the V ring's SLOTS slots
    static constexpr int OFF_S = KV_BYTES, OFF_MB = OFF_S + S_BYTES;
    static constexpr int KS_STAGE = OFF_MB + AH_BK * 4;
    static constexpr int SLOTS = SCORES ? AH_SLOTS_SCORES : AH_SLOTS_PLAIN;  // K ring and V ring depth
    static constexpr int OFF_VRING = SLOTS * KS_STAGE;
    static constexpr int RING_BYTES = OFF_VRING + SLOTS * KV_BYTES;
    static constexpr int PO_PITCH = DH + 4;  // key-split merge: partial O [BQ][PO_PITCH] + (max, sum)[BQ] reuse the rings
    static constexpr int MERGE_BYTES = (BQ * PO_PITCH + BQ * 2) * 4;
    static constexpr int BYTES = RING_BYTES > MERGE_BYTES ? RING_BYTES : MERGE_BYTES;
    static_assert(KS_STAGE % 16 == 0, "slot alignment");
};

struct AhParams {
    const __half* q;
    const __half* k;
    const __half* v;
    const float* scores_p;
    const uint8_t* key_mask;
    float* out;
    int H, n, m, ldq, ldk, ldv;
    float scale_log2e;
    int tiles_per_split;
};

// CTAs per SM the launch bounds ask for: two of the 4-warp shape (111 KB of shared memory each with scores; ~220
// registers: two tiles' logits are live at a time), one of the 8-warp shape
__host__ __device__ constexpr int ah_ctas_per_sm(int warps, bool scores) {
    return warps <= 4 && (scores ? AH_SLOTS_SCORES : AH_SLOTS_PLAIN) <= (scores ? 3 : 6) ? 2 : 1;
}

template <int DH, int WARPS, bool SCORES>
__global__ void __launch_bounds__(32 * WARPS, ah_ctas_per_sm(WARPS, SCORES))
attention_f16_kernel(const __grid_constant__ AhParams prm) {
    using SM = AhSmem<DH, WARPS, SCORES>;
    constexpr int BQ = SM::BQ, THREADS = 32 * WARPS, HP = SM::HP;
    constexpr int KS = DH / 16;  // k16 steps of q.k^T
    constexpr int NT = DH / 8;   // 8-wide n-tiles of the output
    constexpr int CH = DH / 8;   // 16-byte chunks per K / V row
    extern __shared__ __align__(16) uint8_t ah_smem[];
    const int tid = threadIdx.x;
    const uint32_t smem0 = smem_u32(ah_smem);

    const int warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int H = prm.H, n = prm.n, m = prm.m;
    const int bh = blockIdx.y;
    const int b = bh / H, h = bh - b * H;
    const int qbase = blockIdx.x * BQ;
    const int q0 = qbase + warp * 16;
    const int ld = H * DH;
    const int r0 = q0 + g, r1 = q0 + g + 8;  // the two query rows this thread holds fragments of
    const int num_tiles = (m + AH_BK - 1) / AH_BK;
    const int tile_first = blockIdx.z * prm.tiles_per_split;
    const int tile_end = min(num_tiles, tile_first + prm.tiles_per_split);
    const bool sp_vec = (m & 3) == 0;  // 16-byte copies need 16-byte aligned score rows

    // Tile requests, kept lean (one or two warps per scheduler run this between two tiles' math): every thread copies the
    // same (row, 16-byte chunk) positions of every tile, so the positions are worked out once, here.
    constexpr int KV_CHUNKS = AH_BK * CH;                         // 16-byte chunks of one K (or V) tile
    constexpr int KV_ROWS_PER_IT = THREADS / CH;                  // rows between a thread's successive chunks
    constexpr int KV_IT = (KV_CHUNKS + THREADS - 1) / THREADS;
    constexpr int S_ROWS_PER_IT = THREADS / (AH_BK / 4);          // score tile: 16 chunks per row
    constexpr int S_IT = BQ / S_ROWS_PER_IT;                      // = 8
    static_assert(THREADS % CH == 0 && BQ % S_ROWS_PER_IT == 0, "copy pattern");
    const int kv_c8 = tid % CH, kv_row = tid / CH;
    const __half* const k_col = prm.k + (size_t)b * m * prm.ldk + h * DH + 8 * kv_c8;
    const __half* const v_col = prm.v + (size_t)b * m * prm.ldv + h * DH + 8 * kv_c8;
    const uint32_t kv_dst = (kv_row * HP + 8 * kv_c8) * 2;
    const int s_c4 = tid % (AH_BK / 4), s_row = tid / (AH_BK / 4);
    const float* const sp_head = SCORES ? prm.scores_p + ((size_t)b * H + h) * n * (size_t)m : nullptr;
    const uint32_t s_dst = SM::OFF_S + (s_row * AH_SPITCH + 4 * s_c4) * 4;

    // one K (or V) tile: rows past the end of the cloud are zero-filled
    auto issue_rows = [&](const __half* col, int ld_rows, int tile, uint32_t dst) {
        const int j0 = tile * AH_BK;
        const int last = m - 1 - j0;  // tile-local index of the cloud's last key (>= 63 for a full tile)
        const __half* const src = col + (size_t)j0 * ld_rows;
#pragma unroll
        for (int it = 0; it < KV_IT; ++it) {
            const int row = kv_row + it * KV_ROWS_PER_IT;
            if (KV_CHUNKS % THREADS == 0 || row < AH_BK)
                ah_cp16(dst + kv_dst + it * (KV_ROWS_PER_IT * HP * 2), src + min(row, last) * ld_rows, row <= last ? 16u : 0u);
        }
    };
    // K tile + the additive scores of the CTA's queries against its keys + the keys' mask bias -> slot of the K ring
    auto issue_ks = [&](int tile, int slot) {
        if (tile >= tile_end) return;
        const uint32_t base = smem0 + slot * SM::KS_STAGE;
        const int j0 = tile * AH_BK;
        const int last = m - 1 - j0;
        issue_rows(k_col, prm.ldk, tile, base);
        if (SCORES) {
            if (sp_vec) {
                const int jc = 4 * s_c4;  // m % 4 == 0: a 4-key chunk is inside the cloud or outside, never across its end
                const float* const sp_tile = sp_head + j0 + min(jc, last - 3);
                const uint32_t nbytes = jc <= last ? 16u : 0u;
#pragma unroll
                for (int it = 0; it < S_IT; ++it) {
                    const int qr = min(qbase + s_row + it * S_ROWS_PER_IT, n - 1);
                    ah_cp16(base + s_dst + it * (S_ROWS_PER_IT * AH_SPITCH * 4), sp_tile + (size_t)qr * m, nbytes);
                }
            } else {
                for (int idx = tid; idx < BQ * AH_BK; idx += THREADS) {
                    const int row = idx / AH_BK, c = idx - row * AH_BK;
                    ah_cp4(base + SM::OFF_S + (row * AH_SPITCH + c) * 4,
                           sp_head + (size_t)min(qbase + row, n - 1) * m + j0 + min(c, last), c <= last ? 4u : 0u);
                }
            }
        }
        if (tid < AH_BK) {
            const bool dead = tid > last || (prm.key_mask && prm.key_mask[(size_t)b * m + j0 + tid]);
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(base + SM::OFF_MB + tid * 4), "f"(dead ? -INFINITY : 0.f) : "memory");
        }
    };
    auto issue_v = [&](int tile, int slot) {
        if (tile < tile_end) issue_rows(v_col, prm.ldv, tile, smem0 + SM::OFF_VRING + slot * SM::KV_BYTES);
    };

    // cp.async groups (N = SLOTS): G(g) = {K/S(g), V(g)} for g <= N-2, G(N-1) = {K/S(N-1)}, then iteration i commits
    // G(N+i) = {K/S(i+N), V(i+N-1)}: everything iteration i needs (K/S(i+1), V(i)) is in groups <= i+1, and N+i groups
    // have been committed by then: wait_group N-2
    constexpr int N = SM::SLOTS;
#pragma unroll
    for (int g2 = 0; g2 < N; ++g2) {
        issue_ks(tile_first + g2, g2);
        if (g2 < N - 1) issue_v(tile_first + g2, g2);
        ah_commit();
    }

    // Q fragments (A operand, 16 x 16 per k-step): a0 (g, 2t..) a1 (g+8, 2t..) a2 (g, 2t+8..) a3 (g+8, 2t+8..)
    uint32_t qa[KS][4];
    {
        const __half* q_r0 = prm.q + ((size_t)b * n + min(r0, n - 1)) * prm.ldq + h * DH + 2 * t;
        const __half* q_r1 = prm.q + ((size_t)b * n + min(r1, n - 1)) * prm.ldq + h * DH + 2 * t;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
            qa[ks][0] = __ldg((const uint32_t*)(q_r0 + 16 * ks));
            qa[ks][1] = __ldg((const uint32_t*)(q_r1 + 16 * ks));
            qa[ks][2] = __ldg((const uint32_t*)(q_r0 + 16 * ks + 8));
            qa[ks][3] = __ldg((const uint32_t*)(q_r1 + 16 * ks + 8));
        }
    }
    float o[NT][4];
#pragma unroll
    for (int i = 0; i < NT; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    float mx0 = -INFINITY, mx1 = -INFINITY, sum0 = 0.f, sum1 = 0.f;  // running max in RAW logit units (before the scale)
    const float scale_log2e = prm.scale_log2e;
    const bool any_mask = prm.key_mask != nullptr;

    // ---- S = scores_p tile + Q K^T : 8 key n-tiles x KS k-steps.  The staged scores seed the accumulators (C operand).
    //      B frag (k = channel, n = key) straight from the row-major half tile: ldmatrix matrix j = channels 8j..8j+7 of
    //      keys 8nt..8nt+7 ----
    auto qk = [&](float (&s)[AH_BK / 8][4], uint32_t ks_base) {
#pragma unroll
        for (int nt = 0; nt < AH_BK / 8; ++nt) {
            if (SCORES) {  // C frag: (g, 2t) (g, 2t+1) (g+8, 2t) (g+8, 2t+1)
                const uint32_t a = ks_base + SM::OFF_S + ((warp * 16 + g) * AH_SPITCH + nt * 8 + 2 * t) * 4;
                asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(s[nt][0]), "=f"(s[nt][1]) : "r"(a));
                asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(s[nt][2]), "=f"(s[nt][3]) : "r"(a + 8 * AH_SPITCH * 4));
            } else {
                s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
            }
        }
        if (KS >= 2) {  // k-step outer, key n-tile inner: eight independent accumulator chains in flight
#pragma unroll
            for (int ks2 = 0; ks2 < KS / 2; ++ks2) {
#pragma unroll
                for (int nt = 0; nt < AH_BK / 8; ++nt) {
                    uint32_t b0, b1, b2, b3;
                    ah_ldsm4(ks_base + ((nt * 8 + (lane & 7)) * HP + 32 * ks2 + 8 * (lane >> 3)) * 2, b0, b1, b2, b3);
                    ah_mma(s[nt], qa[2 * ks2], b0, b1);
                    ah_mma(s[nt], qa[(2 * ks2 + 1) % KS], b2, b3);
                }
            }
        } else {
#pragma unroll
            for (int nt = 0; nt < AH_BK / 8; ++nt) {
                uint32_t b0, b1;
                ah_ldsm2(ks_base + ((nt * 8 + (lane & 7)) * HP + 8 * ((lane >> 3) & 1)) * 2, b0, b1);
                ah_mma(s[nt], qa[0], b0, b1);
            }
        }
    };
    // dead keys -> -inf (only tiles that have some pay for it)
    auto mask_add = [&](float (&s)[AH_BK / 8][4], uint32_t ks_base, int tile) {
        if (tile < tile_end && (any_mask || tile * AH_BK + AH_BK > m)) {
#pragma unroll
            for (int nt = 0; nt < AH_BK / 8; ++nt) {
                float mb0, mb1;
                asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(mb0), "=f"(mb1) : "r"(ks_base + SM::OFF_MB + (nt * 8 + 2 * t) * 4));
                s[nt][0] += mb0; s[nt][1] += mb1; s[nt][2] += mb0; s[nt][3] += mb1;
            }
        }
    };

    // One tile step.  `cur` holds the raw logits of tile `tile` (computed one step earlier); the step computes the logits
    // of tile + 1 into `nxt` -- independent tensor-pipe work that overlaps this tile's softmax (FMNMX / FFMA / MUFU.EX2)
    // in the same basic block -- then O += P V.  The K ring holds K/S(tile+1 .. tile+N-1) [the later ones in flight], the
    // slot of K/S(tile) is refilled with tile + N; the V ring holds V(tile .. tile+N-2) and refills V(tile-1)'s slot.
    auto step = [&](float (&cur)[AH_BK / 8][4], float (&nxt)[AH_BK / 8][4], int tile, int r) {
        const int r1 = r == N - 1 ? 0 : r + 1, rp = r == 0 ? N - 1 : r - 1;
        ah_wait<N - 2>();  // this thread's copies of K/S(next tile) and V(tile) have landed ...
        __syncthreads();   // ... everybody's have; and every warp is done with K/S(tile) and V(previous tile)
        issue_ks(tile + N, r);
        issue_v(tile + N - 1, rp);
        ah_commit();
        const uint32_t ks_next = smem0 + r1 * SM::KS_STAGE;
        const uint32_t Vh = smem0 + SM::OFF_VRING + r * SM::KV_BYTES;
        qk(nxt, ks_next);  // past the last tile this reads a stale slot: finite work on garbage, discarded

        // ---- running max on the raw logits, weights, running sum ----
        float tmax0 = -INFINITY, tmax1 = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < AH_BK / 8; ++nt) {
            tmax0 = fmaxf(tmax0, fmaxf(cur[nt][0], cur[nt][1]));
            tmax1 = fmaxf(tmax1, fmaxf(cur[nt][2], cur[nt][3]));
        }
        tmax0 = fmaxf(tmax0, __shfl_xor_sync(0xffffffffu, tmax0, 1));
        tmax0 = fmaxf(tmax0, __shfl_xor_sync(0xffffffffu, tmax0, 2));
        tmax1 = fmaxf(tmax1, __shfl_xor_sync(0xffffffffu, tmax1, 1));
        tmax1 = fmaxf(tmax1, __shfl_xor_sync(0xffffffffu, tmax1, 2));
        const float nm0 = fmaxf(mx0, tmax0), nm1 = fmaxf(mx1, tmax1);
        // a row whose keys so far are all masked keeps max = -inf; use 0 as the subtrahend to avoid inf - inf
        const float sub0 = nm0 == -INFINITY ? 0.f : nm0 * scale_log2e, sub1 = nm1 == -INFINITY ? 0.f : nm1 * scale_log2e;
        const float corr0 = ah_exp2(fmaf(mx0, scale_log2e, -sub0)), corr1 = ah_exp2(fmaf(mx1, scale_log2e, -sub1));
        mx0 = nm0; mx1 = nm1;
        float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
        for (int nt = 0; nt < AH_BK / 8; ++nt) {
            cur[nt][0] = ah_exp2(fmaf(cur[nt][0], scale_log2e, -sub0)); cur[nt][1] = ah_exp2(fmaf(cur[nt][1], scale_log2e, -sub0));
            cur[nt][2] = ah_exp2(fmaf(cur[nt][2], scale_log2e, -sub1)); cur[nt][3] = ah_exp2(fmaf(cur[nt][3], scale_log2e, -sub1));
            ps0 += cur[nt][0] + cur[nt][1];
            ps1 += cur[nt][2] + cur[nt][3];
        }
        sum0 = sum0 * corr0 + ps0;
        sum1 = sum1 * corr1 + ps1;
#pragma unroll
        for (int j = 0; j < NT; ++j) { o[j][0] *= corr0; o[j][1] *= corr0; o[j][2] *= corr1; o[j][3] *= corr1; }

        // ---- O += P V.  The C fragments of two adjacent key n-tiles are exactly the A fragment of one k16 step.
        //      B frag (k = key, n = channel) needs V transposed: ldmatrix.trans on the row-major half tile;
        //      matrix j: keys 16kk + 8(j&1).., channels 16nt2 + 8(j>>1).. ----
#pragma unroll
        for (int kk = 0; kk < AH_BK / 16; ++kk) {
            uint32_t pf[4];
            {
                const __half2 h0 = __floats2half2_rn(cur[2 * kk][0], cur[2 * kk][1]);
                const __half2 h1 = __floats2half2_rn(cur[2 * kk][2], cur[2 * kk][3]);
                const __half2 h2 = __floats2half2_rn(cur[2 * kk + 1][0], cur[2 * kk + 1][1]);
                const __half2 h3 = __floats2half2_rn(cur[2 * kk + 1][2], cur[2 * kk + 1][3]);
                pf[0] = *(const uint32_t*)&h0; pf[1] = *(const uint32_t*)&h1;
                pf[2] = *(const uint32_t*)&h2; pf[3] = *(const uint32_t*)&h3;
            }
            const int key = 16 * kk + 8 * ((lane >> 3) & 1) + (lane & 7);
#pragma unroll
            for (int nt2 = 0; nt2 < NT / 2; ++nt2) {
                uint32_t b0, b1, b2, b3;
                ah_ldsm4_trans(Vh + (key * HP + 16 * nt2 + 8 * (lane >> 4)) * 2, b0, b1, b2, b3);
                ah_mma(o[2 * nt2], pf, b0, b1);
                ah_mma(o[2 * nt2 + 1], pf, b2, b3);
            }
        }
        mask_add(nxt, ks_next, tile + 1);
    };

    float sA[AH_BK / 8][4], sB[AH_BK / 8][4];
    ah_wait<N - 1>();  // G0: K/S and V of the first tile
    __syncthreads();
    qk(sA, smem0);
    mask_add(sA, smem0, tile_first);
    {
        int r = 0;
        for (int tile = tile_first; tile < tile_end; tile += 2) {
            step(sA, sB, tile, r);
            r = r == N - 1 ? 0 : r + 1;
            if (tile + 1 >= tile_end) break;
            step(sB, sA, tile + 1, r);
            r = r == N - 1 ? 0 : r + 1;
        }
    }
    ah_wait<0>();

    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1);
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
    if (gridDim.z == 1) {
        // ---- normalise and store: 'b h n c -> b n (h c)' ----
        const float inv0 = 1.f / sum0, inv1 = 1.f / sum1;  // all keys masked -> 0/0 = NaN, as torch.softmax gives
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int d = h * DH + nt * 8 + 2 * t;
            if (r0 < n) *(float2*)(prm.out + ((size_t)b * n + r0) * ld + d) = make_float2(o[nt][0] * inv0, o[nt][1] * inv0);
            if (r1 < n) *(float2*)(prm.out + ((size_t)b * n + r1) * ld + d) = make_float2(o[nt][2] * inv1, o[nt][3] * inv1);
        }
        return;
    }

    // ---- split keys: partial O and (max, sum) stay in this CTA's shared memory; the cluster merges them ----
    constexpr int POP = SM::PO_PITCH;
    float* po = (float*)ah_smem;       // [BQ][POP]   (the stages are no longer needed)
    float* pml = po + BQ * POP;        // [BQ][2]     max in log2 units (raw max x scale), sum
    __syncthreads();                   // every warp is done with the tiles this overwrites
    {
        const int lr0 = warp * 16 + g, lr1 = lr0 + 8;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int d = nt * 8 + 2 * t;
            *(float2*)(po + lr0 * POP + d) = make_float2(o[nt][0], o[nt][1]);
            *(float2*)(po + lr1 * POP + d) = make_float2(o[nt][2], o[nt][3]);
        }
        if (t == 0) {
            *(float2*)(pml + lr0 * 2) = make_float2(mx0 * scale_log2e, sum0);  // -inf stays -inf (scale > 0)
            *(float2*)(pml + lr1 * 2) = make_float2(mx1 * scale_log2e, sum1);
        }
    }
    ptx::cluster_sync();  // all partials of this (query tile, head) are in place, cluster-wide
    {
        // out = sum_z 2^(m_z - M) O_z / sum_z 2^(m_z - M) l_z.  CTA z of the cluster merges rows [z, z+1) * BQ/splits;
        // one thread per (row, 4 channels), peers read through shared::cluster addresses.
        const int splits = gridDim.z;
        const int rows_per_cta = BQ / splits;
        const uint32_t po_u32 = smem_u32(po), pml_u32 = smem_u32(pml);
        for (int item = threadIdx.x; item < rows_per_cta * (DH / 4); item += THREADS) {
            const int lr = blockIdx.z * rows_per_cta + item / (DH / 4);
            const int c4 = item % (DH / 4);
            const int row = qbase + lr;
            float mz[AH_MAX_SPLITS], lz[AH_MAX_SPLITS];
            float M = -INFINITY;
#pragma unroll
            for (int z = 0; z < AH_MAX_SPLITS; ++z) {
                if (z < splits) {
                    const uint32_t a = ptx::map_to_cta(pml_u32 + lr * 8, z);
                    asm volatile("ld.shared::cluster.v2.f32 {%0,%1}, [%2];" : "=f"(mz[z]), "=f"(lz[z]) : "r"(a));
                    M = fmaxf(M, mz[z]);
                }
            }
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float den = 0.f;
#pragma unroll
            for (int z = 0; z < AH_MAX_SPLITS; ++z) {
                if (z < splits) {
                    const float w = ah_exp2(mz[z] - M);  // every split masked: -inf - -inf = NaN, as torch.softmax gives
                    den += w * lz[z];
                    const uint32_t a = ptx::map_to_cta(po_u32 + (lr * POP + 4 * c4) * 4, z);
                    float4 x;
                    asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];"
                                 : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w) : "r"(a));
                    acc.x += w * x.x; acc.y += w * x.y; acc.z += w * x.z; acc.w += w * x.w;
                }
            }
            if (row < n) {
                const float inv = 1.f / den;
                *(float4*)(prm.out + ((size_t)b * n + row) * ld + h * DH + 4 * c4) =
                    make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
            }
        }
    }
    ptx::cluster_sync();  // nobody leaves while a peer may still read its shared memory
}

}  // namespace geotx

using namespace geotx;

static int env_int(const char* name, int dflt, int lo, int hi) {
    const char* e = getenv(name);
    if (!e || !*e) return dflt;
    const int v = atoi(e);
    return v < lo ? lo : v > hi ? hi : v;
}

// Key-range splits (= cluster size): the largest power of two <= 8 that keeps the grid within one wave and leaves
// every CTA at least one 64-key tile (the rule of K3b).
static void ah_plan(int B, int H, int n, int m, int warps, bool scores, int allow_split, int* splits, int* tiles_per_split) {
    const int bq = 16 * warps;
    const long long base = (long long)((n + bq - 1) / bq) * B * H;
    const int tiles = (m + AH_BK - 1) / AH_BK;
    const int wave = sm_count() * ah_ctas_per_sm(warps, scores);
    static const int max_splits = env_int("GEOTX_ATT_MAX_SPLITS", AH_MAX_SPLITS, 1, AH_MAX_SPLITS);  // timing experiments
    int s = 1;
    if (allow_split)
        while (s < max_splits && base * (2 * s) <= wave && 2 * s <= tiles) s *= 2;
    *tiles_per_split = (tiles + s - 1) / s;
    while (s > 1 && (s - 1) * *tiles_per_split >= tiles) {  // no CTA of the cluster may be left without keys
        s /= 2;
        *tiles_per_split = (tiles + s - 1) / s;
    }
    *splits = s;
}

template <int DH, int WARPS, bool SCORES>
static int ah_launch(AhParams prm, int B, int allow_split, cudaStream_t st) {
    using SM = AhSmem<DH, WARPS, SCORES>;
    constexpr int SMEM_BYTES = SM::BYTES;
    static bool opted_in[64] = {};
    int dev = 0;
    int rc = check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
    if (rc) return rc;
    if (dev < 0 || dev >= 64 || !opted_in[dev]) {
        rc = check_cuda(cudaFuncSetAttribute(attention_f16_kernel<DH, WARPS, SCORES>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES),
                        "cudaFuncSetAttribute(attention_f16_kernel)");
        if (rc) return rc;
        if (dev >= 0 && dev < 64) opted_in[dev] = true;
    }
    int splits, tps;
    ah_plan(B, prm.H, prm.n, prm.m, WARPS, SCORES, allow_split, &splits, &tps);
    prm.tiles_per_split = tps;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((prm.n + 16 * WARPS - 1) / (16 * WARPS), B * prm.H, splits);
    cfg.blockDim = dim3(32 * WARPS, 1, 1);
    cfg.dynamicSmemBytes = SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = splits;
    cfg.attrs = attr;
    cfg.numAttrs = splits > 1 ? 1 : 0;  // a plain launch when nothing is split
    rc = check_cuda(cudaLaunchKernelEx(&cfg, attention_f16_kernel<DH, WARPS, SCORES>, prm),
                    "cudaLaunchKernelEx(attention_f16_kernel)");
    if (rc) return rc;
    return launch_status("geotx_attention_f16");
}

template <int DH>
static int ah_launch_shape(const AhParams& prm, int B, int allow_split, cudaStream_t st) {
    // One CTA shape: 64 queries, 4 warps, two CTAs per SM.  K3b's 128-query / 8-warp shape (every K/V tile fetched once
    // per 128 queries) lost here: stand-alone at n = 4096, 92.6 us against 80.6 us (tools/att_time.py) -- K and V are
    // half the bytes and L2-resident, while two independent CTAs per SM desynchronise their barrier / softmax phases.
    return prm.scores_p ? ah_launch<DH, 4, true>(prm, B, allow_split, st) : ah_launch<DH, 4, false>(prm, B, allow_split, st);
}

extern "C" int geotx_attention_f16(const void* q, const void* k, const void* v, const float* scores_p, const uint8_t* key_mask,
                                   int B, int H, int n, int m, int c, int ld_q, int ld_k, int ld_v, float scale, float* out,
                                   int split_keys, void* stream) {
    GEOTX_REQUIRE(q && k && v && out, "geotx_attention_f16: null pointer");
    GEOTX_REQUIRE(B > 0 && H > 0 && n > 0 && m > 0, "geotx_attention_f16: bad sizes");
    GEOTX_REQUIRE(c == 16 || c == 32 || c == 64, "geotx_attention_f16: per-head dim %d not supported (16, 32, 64)", c);
    GEOTX_REQUIRE((((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)out) & 15) == 0,
                  "geotx_attention_f16: q/k/v/out must be 16-byte aligned");
    GEOTX_REQUIRE(scores_p == nullptr || ((uintptr_t)scores_p & 15) == 0, "geotx_attention_f16: scores_p must be 16-byte aligned");
    GEOTX_REQUIRE((long long)B * H <= 65535, "geotx_attention_f16: B*H too large");
    GEOTX_REQUIRE(ld_q >= H * c && ld_k >= H * c && ld_v >= H * c && ld_q % 8 == 0 && ld_k % 8 == 0 && ld_v % 8 == 0,
                  "geotx_attention_f16: row strides must be >= H*c and multiples of 8 halves");
    GEOTX_REQUIRE(scale > 0.f, "geotx_attention_f16: scale must be positive");
    AhParams prm;
    prm.q = (const __half*)q; prm.k = (const __half*)k; prm.v = (const __half*)v;
    prm.scores_p = scores_p; prm.key_mask = key_mask; prm.out = out;
    prm.H = H; prm.n = n; prm.m = m; prm.ldq = ld_q; prm.ldk = ld_k; prm.ldv = ld_v;
    prm.scale_log2e = scale * 1.4426950408889634f;
    prm.tiles_per_split = 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (c == 64) return ah_launch_shape<64>(prm, B, split_keys, st);
    if (c == 32) return ah_launch_shape<32>(prm, B, split_keys, st);
    return ah_launch_shape<16>(prm, B, split_keys, st);
}
