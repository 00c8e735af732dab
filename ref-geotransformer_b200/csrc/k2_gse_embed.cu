// K2: fused GeometricStructureEmbedding on tcgen05 / TMEM, CTA-pair (cta_group::2) version  (sm_100a).
//
// Replaces /root/reference/geotransformer/modules/geotransformer/geotransformer.py:57-72 together with
// /root/reference/geotransformer/modules/transformer/positional_embedding.py:18-34:
//     E[i,j,:] = W_d s(d_ij/sigma_d) + b_d + reduce_kk ( W_a s(angle_ijkk*factor_a) + b_a )
// The reference materialises s(.) as an (n,n,k,C) fp32 tensor (51 GB at n=4096) plus several copies.  Here a
// cluster of two CTAs (one SM pair) owns 256 consecutive (i,j) pairs, 128 per CTA.  For each of the k+1 "slots"
// (k angles, 1 distance) the producer warps of each CTA generate their 128 x C sinusoid rows in registers
// (MUFU sin/cos), pack them to bf16 and store them into a 128B-swizzled K-major shared-memory stage.  The weights
// never move after the prologue: each CTA keeps ITS HALF of the output channels of W_a and W_d resident in shared
// memory (2 x C*C bytes), which is exactly how tcgen05.mma.cta_group::2 wants the B operand split, so one thread of
// the leader CTA issues M=256 x N=C x K=16 MMAs whose A rows come from both CTAs' stages and whose B rows come from
// both CTAs' resident weights; each CTA's TMEM receives its own 128 rows x C columns.  Accumulators ping-pong between
// two TMEM buffers; the epilogue warps pull each finished accumulator out of TMEM once, keep the running max over k in
// registers, add the distance term and the biases and write E (bf16 or fp32) straight to HBM.  Nothing of size
// n*n*k*C ever exists, and no weight byte is re-read from L2 after the first microseconds.
//
// Warp roles (896 threads per CTA): warps 0-15 producers, 16-23 epilogue (TMEM lane quarter = warp % 4, column half =
// (warp-16)/4), 24 MMA issuer (leader CTA only), 25 weight loader (one bulk copy at start), 26 TMEM allocator, 27 idle.
// Register budget (setmaxnreg, 64512 in the CTA's pool): producers 64, epilogue 104, the rest 40.
// Cross-CTA signalling: producers / epilogue warps of the peer CTA arrive remotely on the leader's mbarriers
// (mapa + mbarrier.arrive.release.cluster); tcgen05.commit multicasts "stage free" / "accumulator ready" to both.
#include <cuda_fp16.h>
#include "geotx_common.cuh"
#include "sm100_ptx.cuh"
#include "host_util.h"

namespace geotx {

constexpr int TILE_M = 128;                    // pairs per CTA per tile (256 per cluster)
constexpr int KCHUNK = 64;                     // bf16 elements per 128-byte swizzled row
constexpr int A_STAGE_BYTES = TILE_M * 128;    // 16 KB
constexpr int NUM_PRODUCER_WARPS = 16;  // 4 per SM sub-partition: the per-chunk fence / barrier latencies overlap
constexpr int ARG_PARTS = NUM_PRODUCER_WARPS * 32 / 128;  // producer threads per row of a tile (argument split)
// The producer warps form KCH groups (KCH = C / 64 K-chunks per slot); group g builds K-chunk g of EVERY slot.  A thread
// therefore always works on the same frequencies and keeps its div_term values in registers for the whole kernel:
// fetching them from shared memory per chunk (LDS.128 broadcasts, two wavefronts each, with the multiplies waiting on
// them) cost 15 % of the kernel at C = 256 and 33 % at C = 128.  A warp's fixed per-chunk latencies (empty-barrier
// wait, proxy fence, arrive) are paid once per KCH chunk periods and overlap the MUFU work of the other groups.
constexpr int NUM_EPILOGUE_WARPS = 8;
constexpr int WARP_MMA = NUM_PRODUCER_WARPS + NUM_EPILOGUE_WARPS, WARP_LOADER = WARP_MMA + 1, WARP_ALLOC = WARP_MMA + 2;
constexpr int K2_THREADS = (WARP_MMA + 4) * 32;
#ifndef K2_MUFU_GROUPS
#define K2_MUFU_GROUPS 8  // of a thread's 16-byte pieces per chunk: this many on MUFU, the rest on the FMA pipe (experiment)
#endif

template <int C>
struct K2Cfg {
    static constexpr int KCH = C / KCHUNK;               // K chunks per slot
    static constexpr int W_CHUNK_BYTES = (C / 2) * 128;  // this CTA's C/2 weight rows x 64 bf16
    static constexpr int W_HALF_BYTES = KCH * W_CHUNK_BYTES;  // = C*C bytes: one matrix, this CTA's half of the rows
    static constexpr int NSTAGE = (C == 256) ? 5 : 8;
    static constexpr int PGROUPS = KCH;                                // producer groups
    static constexpr int WARPS_PER_PGROUP = NUM_PRODUCER_WARPS / PGROUPS;
    static constexpr int PARTS = WARPS_PER_PGROUP * 32 / 128;          // threads of one group per row (1 or 2)
    static constexpr int GROUPS_PER_THREAD = 8 / PARTS;                // 16-byte pieces of its 128-byte row a thread builds
#ifndef K2_PREGEN8
#define K2_PREGEN8 4  // C = 256 (8 pieces per thread): how many are built before the stage wait
#endif
    static constexpr bool PREGEN = GROUPS_PER_THREAD <= 4;             // build the whole share before the stage wait
    static_assert(PARTS >= 1 && NUM_PRODUCER_WARPS % PGROUPS == 0, "producer groups");
    static constexpr int TMEM_COLS = 2 * C;              // two fp32 accumulators of C columns (per CTA)
    static constexpr int COLS_PER_THREAD = C / 2;
    static constexpr int NBAR = 2 * NSTAGE + 5;
    static constexpr int SMEM_BYTES = 2 * W_HALF_BYTES + NSTAGE * A_STAGE_BYTES + C * 4 /*bias*/ +
                                      (C / 2) * 4 /*div_term*/ + 2 * TILE_M * 8 * 4 /*pair arguments*/ +
                                      TILE_M * 4 * 4 /*score partials*/ + NBAR * 8 + 16 +
                                      1024 /*alignment slack*/;
};

// ------------------------------------------------------------------------------------------------------------
// Weight repack: W (C x C fp32, [out][in]) -> bf16 image  [half = out/(C/2)][kc = in/64][row = out%(C/2)][64 in]
// with the 16-byte chunk index of every 128-byte row XOR-ed by (row & 7)  (the SWIZZLE_128B pattern tcgen05 expects
// for a K-major tile whose base is 1024-byte aligned).  CTA rank r of a pair bulk-copies half r of both matrices.
// Image order in `packed`: W_a, W_d, then fp32 (b_d + b_a)[C].
// ------------------------------------------------------------------------------------------------------------
__global__ void pack_weights_kernel(const float* __restrict__ w_d, const float* __restrict__ b_d,
                                    const float* __restrict__ w_a, const float* __restrict__ b_a, int C,
                                    uint8_t* __restrict__ packed) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    int total = 2 * C * C;
    if (t < total) {
        int which = t / (C * C);  // 0: W_a, 1: W_d
        int r = t - which * C * C;
        int o = r / C, i = r - o * C;
        const float* w = which == 0 ? w_a : w_d;
        int hc = C / 2;
        int half = o / hc, row = o - half * hc;
        int kc = i >> 6, kk = i & 63, g = kk >> 3, e = kk & 7;
        size_t off = (size_t)which * C * C * 2 + (size_t)half * C * C + (size_t)kc * hc * 128 + (size_t)row * 128 +
                     ((g ^ (row & 7)) << 4) + e * 2;
        *(__nv_bfloat16*)(packed + off) = __float2bfloat16_rn(w[(size_t)o * C + i]);
    }
    if (t < C) ((float*)(packed + (size_t)4 * C * C))[t] = b_d[t] + b_a[t];
}

struct K2Params {
    const float4* pts4;
    const float4* refv;
    const float* div_term;
    const uint8_t* packed;
    void* out;
    long long pairs_total;  // B * row_count * n
    int n, k, row_begin, row_count;
    int num_tile_pairs;     // ceil(ceil(pairs_total / 128) / 2)
    float sigma_d, factor_a;
    int reduce_mean, out_dtype;
    // optional fused first-block scores (geotx_gse_embed_scores): scores[b,h,i,j] = qt[b,h,i,:] . E[b,i,j,:]
    const float* qt;
    long long qt_sb, qt_sh, qt_sn;  // element strides of qt over (batch, head, query row)
    float* scores;                  // [B,H,n,n] fp32
    int H;
};

// watchdog tags: role << 24 | ring position
constexpr uint32_t WD_PRODUCER = 1u << 24, WD_EPILOGUE = 2u << 24, WD_MMA_ACC = 3u << 24, WD_MMA_FULL = 4u << 24,
                   WD_WEIGHTS = 6u << 24;

template <int C, bool MEAN, bool SCORES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(K2_THREADS, 1)
gse_embed_tc_kernel(const __grid_constant__ K2Params prm) {
    using Cfg = K2Cfg<C>;
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment: required by the 128B swizzle atoms of the UMMA descriptors.  The dynamic smem base is the
    // same in both CTAs of the pair, so every offset below is valid in the peer as well.
    // (offset arithmetic on smem_raw, not an integer round trip, so that ptxas keeps LDS/STS instead of generic LD/ST)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* s_w = smem;                                   // [W_a half][W_d half], each KCH x (C/2 rows x 128 B)
    uint8_t* stages = smem + 2 * Cfg::W_HALF_BYTES;        // NSTAGE x 16 KB
    float* s_bias = (float*)(stages + Cfg::NSTAGE * A_STAGE_BYTES);
    float* s_div = s_bias + C;
    float* s_args = s_div + C / 2;                         // [2][slots][128 rows] pair arguments (8 KB reserved: k <= 7)
    float* s_qt = s_args + TILE_M * 8;                     // [4 heads][C] folded query row of the tile (fused scores;
                                                           //  upper half of the same 8 KB: needs slots <= 4)
    float* s_xch = s_args + 2 * TILE_M * 8;                // [128 rows][4 heads] score partials of the upper column half
    uint64_t* bars = (uint64_t*)(s_xch + TILE_M * 4);
    uint32_t* s_tmem = (uint32_t*)(bars + Cfg::NBAR);

    const uint32_t bar_full = smem_u32(bars);                       // [NSTAGE]  leader: 16 warp arrivals
    const uint32_t bar_empty = bar_full + 8 * Cfg::NSTAGE;          // [NSTAGE]  each CTA: 1 (multicast commit)
    const uint32_t bar_acc_full = bar_empty + 8 * Cfg::NSTAGE;      // [2]       each CTA: 1 (multicast commit)
    const uint32_t bar_acc_empty = bar_acc_full + 16;               // [2]       leader: 16 warp arrivals
    const uint32_t bar_w = bar_acc_empty + 16;                      // [1]       each CTA: own weight copy
    const uint32_t stages_u32 = smem_u32(stages);
    const uint32_t w_u32 = smem_u32(s_w);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1;
    const int num_clusters = gridDim.x >> 1;
    const int k = prm.k;
    const int slots = k + 1;

    // ---- one-time setup ----
    for (int t = threadIdx.x; t < C; t += K2_THREADS)
        s_bias[t] = ((const float*)(prm.packed + (size_t)4 * C * C))[t];
    for (int t = threadIdx.x; t < C / 2; t += K2_THREADS) s_div[t] = prm.div_term[t];
    if (threadIdx.x == 0) {
        for (int s = 0; s < Cfg::NSTAGE; ++s) {
            ptx::mbar_init(bar_full + 8 * s, 2 * Cfg::WARPS_PER_PGROUP);
            ptx::mbar_init(bar_empty + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            ptx::mbar_init(bar_acc_full + 8 * b, 1);
            ptx::mbar_init(bar_acc_empty + 8 * b, 2 * NUM_EPILOGUE_WARPS);
        }
        ptx::mbar_init(bar_w, 1);
        ptx::fence_barrier_init();
    }
    if (warp == WARP_ALLOC) ptx::tmem_alloc_pair(smem_u32(s_tmem), Cfg::TMEM_COLS);
    __syncthreads();
    if (warp == WARP_LOADER && lane == 0) {
        // resident weights: this CTA's half of the output channels of W_a and W_d (TMA unit bulk copies)
        ptx::mbar_arrive_expect_tx(bar_w, 2 * Cfg::W_HALF_BYTES);
#pragma unroll 1
        for (int which = 0; which < 2; ++which) {
            const uint8_t* src = prm.packed + (size_t)which * 2 * C * C + (size_t)rank * Cfg::W_HALF_BYTES;
#pragma unroll 1
            for (int kc = 0; kc < Cfg::KCH; ++kc)
                ptx::bulk_g2s(w_u32 + which * Cfg::W_HALF_BYTES + kc * Cfg::W_CHUNK_BYTES,
                              src + (size_t)kc * Cfg::W_CHUNK_BYTES, Cfg::W_CHUNK_BYTES, bar_w);
        }
    }
    ptx::mbar_wait(bar_w, 0, WD_WEIGHTS);
    __syncwarp();  // barrier.cluster.*.aligned needs converged warps
    ptx::tc_fence_before_sync();
    ptx::cluster_sync();  // peer's barriers initialised, peer's weights resident, both TMEM allocations done
    ptx::tc_fence_after_sync();
    const uint32_t tmem_base = *s_tmem;

    if (warp < NUM_PRODUCER_WARPS) {
        // =============================== PRODUCERS: sinusoid rows -> swizzled smem ===============================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
        const int row = threadIdx.x & (TILE_M - 1);
        const int argpart = threadIdx.x >> 7;     // which of its row's k+1 arguments this thread evaluates
        constexpr int PGROUPS = Cfg::PGROUPS, PARTS = Cfg::PARTS, GROUPS_PER_THREAD = Cfg::GROUPS_PER_THREAD;
        const int pgroup = argpart / PARTS;       // producer group = the K-chunk (of every slot) it builds
        const int part = argpart - pgroup * PARTS;  // which GROUPS_PER_THREAD of the 8 sixteen-byte pieces of a row
        const long long per_b = (long long)prm.row_count * prm.n;
        const float4* s_div4 = (const float4*)s_div;
        const uint32_t full_leader = ptx::map_to_cta(bar_full, 0);
        const uint32_t a_row = stages_u32 + row * 128;
        const int sw = row & 7;  // 128-byte swizzle: 16-byte piece index XOR (row mod 8)
        // stage ring position of this group's next chunk and the parity of its pass
        uint32_t stage = pgroup % Cfg::NSTAGE, phase = 0;

        // The k+1 scalar arguments of a pair: the four producer threads of a row split them (atan2f is ~1/3 of a
        // producer's instructions if every thread evaluates all k of them) and exchange them through a
        // double-buffered table [buffer][slot][row] (conflict-free both ways): one named barrier per tile.
        int xbuf = 0;
        const float* x_row = s_args + row;  // + slot * TILE_M: arguments of the tile being generated
        auto load_x = [&](int tp) {
            float* xs = s_args + xbuf * (TILE_M * slots);
            const long long p = ((long long)tp * 2 + rank) * TILE_M + row;
            if (p < prm.pairs_total) {
                const long long b = p / per_b;
                const long long rem = p - b * per_b;
                const int i = prm.row_begin + (int)(rem / prm.n);
                const int j = (int)(rem % prm.n);
                const float4 pi = __ldg(prm.pts4 + b * prm.n + i);
                const float4 pj = __ldg(prm.pts4 + b * prm.n + j);
                for (int s = argpart; s < slots; s += ARG_PARTS)
#ifdef K2_DBG_NO_ARGS  // timing experiment: no atan2 / sqrt (wrong values)
                    xs[s * TILE_M + row] = pi.x * pj.y + (float)s;
#else
                    xs[s * TILE_M + row] = pair_argument_one(pi, pj, prm.refv + (b * prm.n + i) * k, k, s,
                                                             prm.sigma_d, prm.factor_a);
#endif
            } else {
                for (int s = argpart; s < slots; s += ARG_PARTS) xs[s * TILE_M + row] = 0.f;
            }
            asm volatile("bar.sync 1, %0;" ::"n"(NUM_PRODUCER_WARPS * 32) : "memory");
            x_row = xs + row;
            xbuf ^= 1;
        };

        // this thread's frequencies, for the whole kernel (chunk pgroup of every slot, pieces part*GPT .. +GPT-1)
        float4 w[GROUPS_PER_THREAD];
#pragma unroll
        for (int gg = 0; gg < GROUPS_PER_THREAD; ++gg) {
            w[gg] = s_div4[pgroup * 8 + part * GROUPS_PER_THREAD + gg];
        }
        // one 16-byte piece: 4 frequencies -> (sin, cos) pairs as bf16
        // positional_embedding.py:28-31: omega = x * div_term (fp32 product), then sin / cos
        auto piece = [&](float xv, int gg, uint32_t* o) {
            const float o0 = __fmul_rn(xv, w[gg].x), o1 = __fmul_rn(xv, w[gg].y);
            const float o2 = __fmul_rn(xv, w[gg].z), o3 = __fmul_rn(xv, w[gg].w);
#ifdef K2_DBG_NO_SINCOS  // timing experiments only (tools/k2_variants.sh)
            o[0] = pack_bf16x2(o0, o0 + 1.f); o[1] = pack_bf16x2(o1, o1 + 1.f);
            o[2] = pack_bf16x2(o2, o2 + 1.f); o[3] = pack_bf16x2(o3, o3 + 1.f);
#else
            if (gg < K2_MUFU_GROUPS) {  // MUFU.SIN / MUFU.COS (XU pipe, 16 results / clk / SM)
                o[0] = pack_bf16x2(__sinf(o0), __cosf(o0));
                o[1] = pack_bf16x2(__sinf(o1), __cosf(o1));
                o[2] = pack_bf16x2(__sinf(o2), __cosf(o2));
                o[3] = pack_bf16x2(__sinf(o3), __cosf(o3));
            } else {                    // polynomial on the FMA pipe (experiment: not faster)
                float s0, c0, s1, c1, s2, c2, s3, c3;
                sincos_fma(o0, s0, c0); sincos_fma(o1, s1, c1); sincos_fma(o2, s2, c2); sincos_fma(o3, s3, c3);
                o[0] = pack_bf16x2(s0, c0); o[1] = pack_bf16x2(s1, c1);
                o[2] = pack_bf16x2(s2, c2); o[3] = pack_bf16x2(s3, c3);
            }
#endif
        };
        auto store_piece = [&](uint32_t a_dst, int gg, const uint32_t* o) {
#if !defined(K2_DBG_NO_PRODUCE) && !defined(K2_DBG_NO_STS)
            const uint32_t addr = a_dst + (((part * GROUPS_PER_THREAD + gg) ^ sw) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3])
                         : "memory");
#endif
        };

        for (int tp = cluster_id; tp < prm.num_tile_pairs; tp += num_clusters) {
            load_x(tp);
#pragma unroll 1
            for (int s = 0; s < slots; ++s) {  // chunk (slot s, K-chunk pgroup)
                const float xv = x_row[s * TILE_M];  // angle s for s < k, distance at s == k
                const uint32_t a_dst = a_row + stage * A_STAGE_BYTES;
#if defined(K2_DBG_NO_PRODUCE) || defined(K2_DBG_NO_GEN)
                ptx::mbar_wait(bar_empty + 8 * stage, phase ^ 1, WD_PRODUCER | (uint32_t)s);
                uint32_t z[4] = {0u, 0u, 0u, 0u};
#pragma unroll
                for (int gg = 0; gg < GROUPS_PER_THREAD; ++gg) store_piece(a_dst, gg, z);
#else
                if (Cfg::PREGEN) {
                    // the thread's whole share is built in registers BEFORE the wait: the stage is normally free by then
                    uint32_t v[GROUPS_PER_THREAD][4];
#pragma unroll
                    for (int gg = 0; gg < GROUPS_PER_THREAD; ++gg) piece(xv, gg, v[gg]);
                    ptx::mbar_wait(bar_empty + 8 * stage, phase ^ 1, WD_PRODUCER | (uint32_t)s);
#pragma unroll
                    for (int gg = 0; gg < GROUPS_PER_THREAD; ++gg) store_piece(a_dst, gg, v[gg]);
                } else {
                    // 8 pieces do not fit next to the resident frequencies: the first half is built before the wait
                    constexpr int HALF = K2_PREGEN8;
                    uint32_t v[HALF > 0 ? HALF : 1][4];
#pragma unroll
                    for (int gg = 0; gg < HALF; ++gg) piece(xv, gg, v[gg]);
                    ptx::mbar_wait(bar_empty + 8 * stage, phase ^ 1, WD_PRODUCER | (uint32_t)s);
#pragma unroll
                    for (int gg = 0; gg < HALF; ++gg) store_piece(a_dst, gg, v[gg]);
#pragma unroll
                    for (int gg = HALF; gg < GROUPS_PER_THREAD; ++gg) {
                        uint32_t o[4];
                        piece(xv, gg, o);
                        store_piece(a_dst, gg, o);
                    }
                }
#endif
#ifndef K2_DBG_NO_FENCE
                ptx::fence_proxy_async_smem();  // make the st.shared visible to the tensor core (async proxy)
#endif
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive_cluster(full_leader + 8 * stage);
                stage += PGROUPS;
                if (stage >= Cfg::NSTAGE) { stage -= Cfg::NSTAGE; phase ^= 1; }
            }
        }
    } else if (warp < NUM_PRODUCER_WARPS + NUM_EPILOGUE_WARPS) {
        // =============================== EPILOGUE: TMEM -> registers -> running max -> HBM =======================
        // The running reduction over the k angle slots is kept as packed f16x2 (11-bit mantissa: 4x finer than the bf16
        // the result is stored in), which halves its register footprint -- the registers go to the producers.
        asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");
        constexpr int NCOL = Cfg::COLS_PER_THREAD;
        const int quarter = warp & 3;                       // TMEM lanes 32*quarter .. +31 belong to this warp
        const int colhalf = (warp - NUM_PRODUCER_WARPS) >> 2;
        const int row = quarter * 32 + lane;
        const int groups = MEAN ? 2 : slots;
        const float inv_k = 1.0f / (float)k;  // 'mean': the k angle slots are summed by the MMA into one accumulator
        const uint32_t acc_empty_leader = ptx::map_to_cta(bar_acc_empty, 0);
        __half2 run[NCOL / 2];
        uint32_t gc = 0;  // running accumulation-group counter (TMEM buffer ring position)
        for (int tp = cluster_id; tp < prm.num_tile_pairs; tp += num_clusters) {
            const long long p = ((long long)tp * 2 + rank) * TILE_M + row;
            // fused scores (host guarantees n % 128 == 0 and slots <= 4): the 128 pairs of a tile share the query row
            // (b, i); its folded query [H][C] is staged in shared memory by the 8 epilogue warps at tile start.
            float sc[4] = {0.f, 0.f, 0.f, 0.f};
            long long s_off = 0;
            if (SCORES) {
                const long long p0 = ((long long)tp * 2 + rank) * TILE_M;
                if (p0 < prm.pairs_total) {
                    const long long per_b = (long long)prm.row_count * prm.n;
                    const long long b = p0 / per_b;
                    const long long rem = p0 - b * per_b;
                    const int i = prm.row_begin + (int)(rem / prm.n);
                    const int j = (int)(rem % prm.n) + row;
                    s_off = ((b * prm.H) * prm.n + i) * (long long)prm.n + j;
                    const int te = (warp - NUM_PRODUCER_WARPS) * 32 + lane;  // 0..255
                    const int h = te / (C / 4), c4 = te - h * (C / 4);
                    if (h < prm.H)
                        *(float4*)(s_qt + h * C + 4 * c4) =
                            __ldg((const float4*)(prm.qt + b * prm.qt_sb + h * prm.qt_sh + (long long)i * prm.qt_sn) + c4);
                }
            }
            for (int g = 0; g < groups; ++g, ++gc) {
                const uint32_t buf = gc & 1, phase = (gc >> 1) & 1;
                ptx::mbar_wait_backoff(bar_acc_full + 8 * buf, phase, WD_EPILOGUE | (gc & 0xFFFFFF));
                ptx::tc_fence_after_sync();
                const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + buf * C + colhalf * NCOL;
                if (g < groups - 1) {
                    // angle groups: running max (or the single pre-summed group of 'mean')
#ifndef K2_DBG_NO_EPI_LD
#pragma unroll
                    for (int c0 = 0; c0 < NCOL; c0 += 16) {
                        uint32_t v[16];
                        ptx::tmem_ld16(taddr + c0, v);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int c = 0; c < 16; c += 2) {
                            const __half2 a = MEAN ? __floats2half2_rn(__uint_as_float(v[c]) * inv_k,
                                                                       __uint_as_float(v[c + 1]) * inv_k)
                                                   : __floats2half2_rn(__uint_as_float(v[c]), __uint_as_float(v[c + 1]));
                            run[(c0 + c) / 2] = (g == 0) ? a : __hmax2(run[(c0 + c) / 2], a);
                        }
                    }
#else
                    if (g == 0) {
#pragma unroll
                        for (int c = 0; c < NCOL / 2; ++c) run[c] = __floats2half2_rn((float)(c + lane), 1.f);
                    }
#endif
                } else {
                    // distance group: E = (W_d s_d) + reduce_k(W_a s_a) + (b_d + b_a), straight to HBM in 32-byte pieces
                    const float4* bias4 = (const float4*)(s_bias + colhalf * NCOL);
                    if (SCORES) asm volatile("bar.sync 6, 256;" ::: "memory");  // the tile's folded query is staged
#ifdef K2_DBG_NO_EPI_STORE
                    const bool do_store = p < prm.pairs_total && __low2float(run[0]) == 123.456f;
#else
                    const bool do_store = p < prm.pairs_total;
#endif
#pragma unroll
                    for (int c0 = 0; c0 < NCOL; c0 += 16) {
                        float e[16];
#ifndef K2_DBG_NO_EPI_LD
                        uint32_t v[16];
                        ptx::tmem_ld16(taddr + c0, v);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int c = 0; c < 16; ++c) e[c] = __uint_as_float(v[c]);
#else
#pragma unroll
                        for (int c = 0; c < 16; ++c) e[c] = 1.f;
#endif
#pragma unroll
                        for (int c = 0; c < 16; c += 4) {
                            const float4 bb = bias4[(c0 + c) / 4];
                            const float2 r0 = __half22float2(run[(c0 + c) / 2]);
                            const float2 r1 = __half22float2(run[(c0 + c) / 2 + 1]);
                            e[c] += r0.x + bb.x; e[c + 1] += r0.y + bb.y;
                            e[c + 2] += r1.x + bb.z; e[c + 3] += r1.y + bb.w;
                        }
                        if (SCORES) {
                            // broadcast reads (every lane the same address): one wavefront each
#pragma unroll
                            for (int h = 0; h < 4; ++h) {
                                if (h < prm.H) {
                                    const float4* q4 = (const float4*)(s_qt + h * C + colhalf * NCOL + c0);
#pragma unroll
                                    for (int u = 0; u < 4; ++u) {
                                        const float4 qv = q4[u];
                                        sc[h] = fmaf(qv.x, e[4 * u], sc[h]);
                                        sc[h] = fmaf(qv.y, e[4 * u + 1], sc[h]);
                                        sc[h] = fmaf(qv.z, e[4 * u + 2], sc[h]);
                                        sc[h] = fmaf(qv.w, e[4 * u + 3], sc[h]);
                                    }
                                }
                            }
                        }
                        if (do_store) {
                            if (prm.out_dtype == GEOTX_F16) {  // default storage format of E: 11-bit significand
                                __half2 o[8];
#pragma unroll
                                for (int c = 0; c < 8; ++c) o[c] = __floats2half2_rn(e[2 * c], e[2 * c + 1]);
                                st_global_v8((__half*)prm.out + p * C + colhalf * NCOL + c0, (const uint32_t*)o);
                            } else if (prm.out_dtype == GEOTX_BF16) {
                                uint32_t o[8];
#pragma unroll
                                for (int c = 0; c < 8; ++c) o[c] = pack_bf16x2(e[2 * c], e[2 * c + 1]);
                                st_global_v8((__nv_bfloat16*)prm.out + p * C + colhalf * NCOL + c0, o);
                            } else {
                                float* dst = (float*)prm.out + p * C + colhalf * NCOL + c0;
                                st_global_v8(dst, (const uint32_t*)e);
                                st_global_v8(dst + 8, (const uint32_t*)e + 8);
                            }
                        }
                    }
                }
                ptx::tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive_cluster(acc_empty_leader + 8 * buf);
            }
            if (SCORES) {
                // the two warps of a lane quarter hold the two column halves of the same 32 rows: add them through
                // shared memory (named barrier 2 + quarter over those 64 threads) and write scores[b,h,i,j]
                float4* x4 = (float4*)(s_xch + row * 4);
                if (colhalf == 1) *x4 = make_float4(sc[0], sc[1], sc[2], sc[3]);
                asm volatile("bar.sync %0, 64;" ::"r"(2 + quarter) : "memory");
                if (colhalf == 0 && p < prm.pairs_total) {
                    const float4 o = *x4;
                    const float tot[4] = {sc[0] + o.x, sc[1] + o.y, sc[2] + o.z, sc[3] + o.w};
                    const long long head_stride = (long long)prm.n * prm.n;
#pragma unroll
                    for (int h = 0; h < 4; ++h)
                        if (h < prm.H) prm.scores[s_off + h * head_stride] = tot[h];
                }
                asm volatile("bar.sync 6, 256;" ::: "memory");  // partials and the staged query consumed: reusable
            }
        }
    } else {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == WARP_MMA && rank == 0) {
            // =============================== MMA ISSUER (leader CTA) ============================================
            // The whole warp walks the loop (uniform control flow, operands in uniform registers); one elected lane
            // issues the tcgen05 instructions.
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(2 * TILE_M, C);
            uint32_t stage = 0, phase = 0, gc = 0;
            for (int tp = cluster_id; tp < prm.num_tile_pairs; tp += num_clusters) {
                for (int s = 0; s < slots; ++s) {
                    const bool first = MEAN ? (s == 0 || s == k) : true;
                    const bool last = MEAN ? (s == k - 1 || s == k) : true;
                    const uint32_t buf = gc & 1;
                    if (first) {
                        ptx::mbar_wait(bar_acc_empty + 8 * buf, ((gc >> 1) & 1) ^ 1, WD_MMA_ACC | (gc & 0xFFFFFF));
                        ptx::tc_fence_after_sync();
                    }
                    const uint32_t tmem_d = tmem_base + buf * C;
                    const uint32_t w_mat = w_u32 + (s < k ? 0 : Cfg::W_HALF_BYTES);  // W_a | W_d
                    for (int kc = 0; kc < Cfg::KCH; ++kc) {
                        ptx::mbar_wait(bar_full + 8 * stage, phase, WD_MMA_FULL | (uint32_t)(s * Cfg::KCH + kc));
                        ptx::tc_fence_after_sync();
                        const uint64_t adesc = ptx::umma_desc_sw128(stages_u32 + stage * A_STAGE_BYTES);
                        const uint64_t bdesc = ptx::umma_desc_sw128(w_mat + kc * Cfg::W_CHUNK_BYTES);
                        if (ptx::elect_one()) {
#ifndef K2_DBG_NO_MMA
#pragma unroll
                            for (int ks = 0; ks < KCHUNK / 16; ++ks) {
                                // advance 16 bf16 = 32 bytes along K inside the 128-byte swizzled row: +2 (16-B units)
                                ptx::umma_bf16_pair(tmem_d, adesc + 2 * ks, bdesc + 2 * ks, idesc,
                                                    (uint32_t)(!(first && kc == 0 && ks == 0)));
                            }
#endif
                            ptx::umma_commit_pair(bar_empty + 8 * stage, 3);  // stage reusable in both CTAs
                            if (last && kc == Cfg::KCH - 1) ptx::umma_commit_pair(bar_acc_full + 8 * buf, 3);
                        }
                        __syncwarp();
                        if (++stage == Cfg::NSTAGE) { stage = 0; phase ^= 1; }
                    }
                    if (last) ++gc;
                }
            }
        }
    }

    // ---- teardown: nobody may leave while the peer can still read our smem / arrive on our barriers ----
    __syncwarp();  // reconverge the single-thread roles: barrier.cluster.*.aligned needs converged warps
    ptx::tc_fence_before_sync();
    ptx::cluster_sync();
    if (warp == WARP_ALLOC) {
        ptx::tc_fence_after_sync();
        ptx::tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS);
    }
}

}  // namespace geotx

// ---- C ABI --------------------------------------------------------------------------------------------------
using namespace geotx;

extern "C" int geotx_watchdog_status(uint32_t* host_out4) {
    GEOTX_REQUIRE(host_out4, "geotx_watchdog_status: null pointer");
    int rc = check_cuda(cudaMemcpyFromSymbol(host_out4, ptx::g_watchdog, 16), "cudaMemcpyFromSymbol(g_watchdog)");
    if (rc) return rc;
    if (host_out4[0]) {
        const uint32_t zero[4] = {0, 0, 0, 0};
        rc = check_cuda(cudaMemcpyToSymbol(ptx::g_watchdog, zero, 16), "cudaMemcpyToSymbol(g_watchdog)");
    }
    return rc;
}

extern "C" int64_t geotx_gse_packed_weight_bytes(int C) {
    if (C <= 0 || C % 64 != 0) return 0;
    return (int64_t)4 * C * C + (int64_t)4 * C;
}

extern "C" int geotx_gse_pack_weights(const float* w_d, const float* b_d, const float* w_a, const float* b_a, int C,
                                      void* packed, void* stream) {
    GEOTX_REQUIRE(w_d && b_d && w_a && b_a && packed, "geotx_gse_pack_weights: null pointer");
    GEOTX_REQUIRE(C == 128 || C == 256, "geotx_gse_pack_weights: hidden_dim %d not supported by the tcgen05 path "
                  "(128 or 256); use geotx_gse_embed_fp32", C);
    GEOTX_REQUIRE(((uintptr_t)packed & 15) == 0, "geotx_gse_pack_weights: packed must be 16-byte aligned");
    int total = 2 * C * C;
    pack_weights_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w_d, b_d, w_a, b_a, C,
                                                                               (uint8_t*)packed);
    return launch_status("geotx_gse_pack_weights");
}

template <int C>
static int launch_k2(const K2Params& prm, cudaStream_t st) {
    using Cfg = K2Cfg<C>;
    const void* kernels[4] = {(const void*)gse_embed_tc_kernel<C, false, false>, (const void*)gse_embed_tc_kernel<C, true, false>,
                              (const void*)gse_embed_tc_kernel<C, false, true>, (const void*)gse_embed_tc_kernel<C, true, true>};
    // the opt-in to > 48 KB of dynamic shared memory belongs to the device's context: once per device, not per process
    static bool opted_in[64] = {};
    int dev = 0;
    int rc0 = check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
    if (rc0) return rc0;
    if (dev < 0 || dev >= 64 || !opted_in[dev]) {
        for (int i = 0; i < 4 && !rc0; ++i)
            rc0 = check_cuda(cudaFuncSetAttribute(kernels[i], cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES),
                             "cudaFuncSetAttribute(gse_embed_tc_kernel)");
        if (rc0) return rc0;
        if (dev >= 0 && dev < 64) opted_in[dev] = true;
    }
    const int num_sms = sm_count();
    // persistent: one CTA pair per SM pair (cluster of 2, same TPC)
    int clusters = num_sms / 2;
    if (prm.num_tile_pairs < clusters) clusters = prm.num_tile_pairs;
    const int which = (prm.reduce_mean ? 1 : 0) + (prm.scores ? 2 : 0);
    K2Params p = prm;
    void* args[1] = {&p};
    int rc = check_cuda(cudaLaunchKernel(kernels[which], dim3(2 * clusters), dim3(K2_THREADS), args, Cfg::SMEM_BYTES, st),
                        "cudaLaunchKernel(gse_embed_tc_kernel)");
    if (rc) return rc;
    return launch_status("geotx_gse_embed");
}

static int gse_embed_common(const char* who, const void* geometry, int B, int n, int k, int row_begin, int row_count,
                            const float* div_term, const void* packed_weights, int C, float sigma_d, float factor_a,
                            int reduce_mean, int out_dtype, void* out, const float* qt, int64_t qt_stride_b,
                            int64_t qt_stride_h, int64_t qt_stride_n, int H, float* scores_p, void* stream) {
    GEOTX_REQUIRE(geometry && div_term && packed_weights && out, "%s: null pointer", who);
    GEOTX_REQUIRE(B > 0 && n > 0 && k >= 1 && k <= GEOTX_MAX_ANGLE_K, "%s: bad sizes", who);
    GEOTX_REQUIRE(C == 128 || C == 256, "%s: hidden_dim %d not supported by the tcgen05 path (128 or "
                  "256); use geotx_gse_embed_fp32", who, C);
    GEOTX_REQUIRE(row_begin >= 0 && row_count > 0 && row_begin + row_count <= n, "%s: bad row slab", who);
    GEOTX_REQUIRE(out_dtype == GEOTX_F32 || out_dtype == GEOTX_BF16 || out_dtype == GEOTX_F16, "%s: bad out_dtype", who);
    GEOTX_REQUIRE(((uintptr_t)out & 31) == 0 && ((uintptr_t)packed_weights & 15) == 0,
                  "%s: out must be 32-byte and packed_weights 16-byte aligned", who);
    K2Params prm;
    prm.pts4 = (const float4*)geometry;
    prm.refv = prm.pts4 + (size_t)B * n;
    prm.div_term = div_term;
    prm.packed = (const uint8_t*)packed_weights;
    prm.out = out;
    prm.pairs_total = (long long)B * row_count * n;
    prm.n = n; prm.k = k; prm.row_begin = row_begin; prm.row_count = row_count;
    long long tiles = (prm.pairs_total + TILE_M - 1) / TILE_M;
    GEOTX_REQUIRE(tiles < 2147483647LL, "%s: too many pairs", who);
    prm.num_tile_pairs = (int)((tiles + 1) / 2);
    prm.sigma_d = sigma_d; prm.factor_a = factor_a;
    prm.reduce_mean = reduce_mean ? 1 : 0;
    prm.out_dtype = out_dtype;
    prm.qt = qt; prm.qt_sb = qt_stride_b; prm.qt_sh = qt_stride_h; prm.qt_sn = qt_stride_n;
    prm.scores = scores_p; prm.H = H;
    if (C == 256) return launch_k2<256>(prm, (cudaStream_t)stream);
    return launch_k2<128>(prm, (cudaStream_t)stream);
}

extern "C" int geotx_gse_embed(const void* geometry, int B, int n, int k, int row_begin, int row_count,
                               const float* div_term, const void* packed_weights, int C, float sigma_d,
                               float factor_a, int reduce_mean, int out_dtype, void* out, void* stream) {
    return gse_embed_common("geotx_gse_embed", geometry, B, n, k, row_begin, row_count, div_term, packed_weights, C,
                            sigma_d, factor_a, reduce_mean, out_dtype, out, nullptr, 0, 0, 0, 0, nullptr, stream);
}

extern "C" int geotx_gse_embed_scores(const void* geometry, int B, int n, int k, int row_begin, int row_count,
                                      const float* div_term, const void* packed_weights, int C, float sigma_d,
                                      float factor_a, int reduce_mean, int out_dtype, void* out, const float* qt,
                                      int64_t qt_stride_b, int64_t qt_stride_h, int64_t qt_stride_n, int H,
                                      float* scores_p, void* stream) {
    GEOTX_REQUIRE(qt && scores_p, "geotx_gse_embed_scores: null pointer");
    GEOTX_REQUIRE(H >= 1 && H <= 4, "geotx_gse_embed_scores: %d heads not supported by the fused scores (1..4); use "
                  "geotx_gse_embed + geotx_rpe_scores", H);
    GEOTX_REQUIRE(((uintptr_t)qt & 15) == 0 && qt_stride_b % 4 == 0 && qt_stride_h % 4 == 0 && qt_stride_n % 4 == 0,
                  "geotx_gse_embed_scores: qt must be 16-byte aligned with strides that are multiples of 4 floats");
    if (n % TILE_M != 0 || k > 3) {
        // tiles would straddle query rows (or the argument table needs the staging space): same result from the two
        // separate kernels, E read back once
        int rc = gse_embed_common("geotx_gse_embed_scores", geometry, B, n, k, row_begin, row_count, div_term,
                                  packed_weights, C, sigma_d, factor_a, reduce_mean, out_dtype, out, nullptr, 0, 0, 0, 0,
                                  nullptr, stream);
        if (rc) return rc;
        return geotx_rpe_scores(qt, qt_stride_b, qt_stride_h, qt_stride_n, out, out_dtype, B, H, n, n, C, row_begin,
                                row_count, scores_p, stream);
    }
    return gse_embed_common("geotx_gse_embed_scores", geometry, B, n, k, row_begin, row_count, div_term, packed_weights,
                            C, sigma_d, factor_a, reduce_mean, out_dtype, out, qt, qt_stride_b, qt_stride_h, qt_stride_n,
                            H, scores_p, stream);
}
