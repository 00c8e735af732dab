// K2c: GeometricStructureEmbedding as a LOW-RANK contraction (the default embedding kernel).
//
// geotransformer.py:57-72 computes E[i,j,:] = proj_d(sinus(d_ij)) + reduce_k proj_a(sinus(a_ijk)).  Per output channel
// c, proj_x(sinus(x))[c] is a function of ONE scalar (positional_embedding.py:28-31 followed by a Linear):
//     f_c(x) = sum_t W[c,2t] sin(x div_t) + W[c,2t+1] cos(x div_t),      0 < div_t <= 1,
// i.e. band-limited (highest frequency 1 rad per index unit) on a bounded interval [0, L] (angles: L = pi * factor_a
// = 12 index units for sigma_a = 15; distances: L = the cloud's largest distance / sigma_d).  Such a function has a
// Chebyshev expansion of length K ~ L/2 + 12 that is exact to ~1e-6 of sum_t |W[c,t]| (the coefficients decay like the
// Bessel functions J_m(L/2)): K = 16 covers L <= 13, 32 -> 37, 48 -> 64, 64 -> 93, 96 -> 150, 128 -> 210.  So
//     f_c(x) = sum_{m<K} G[m,c] T_m(2x/L - 1)          (G: K x C, built once per weight set in fp64)
// and the (n^2 (1+k)) x C x C contraction of the reference (K2, k2_gse_embed.cu: 524 kflop per pair) becomes a
// (n^2 (1+k)) x C x K one with K = 16 per angle slot and 32-48 for the distance: 41-49 kflop per pair.  Unlike a
// sampled table (K2t, k2_gse_embed_table.cu) the truncation error does not depend on the size of the weights (it is
// relative to them, and is measured in fp64 when the coefficients are built), and any angle_k and both reductions are
// covered ('mean' is linear: the slots' basis vectors are averaged before the contraction).
//
// The kernel (one persistent kernel, two outputs: E for K2c, scores_p = qt . E for K3r, where E never leaves the
// registers): a CTA of 12 warps walks tiles of 128 consecutive (i,j) pairs, two CTAs per SM, two or three tile buffers per CTA.
//   4 producer warps (setmaxnreg 48), one thread per pair: the pair's k+1 arguments (the arithmetic of
//            get_embedding_indices with fast reciprocal square roots and a minimax atan), the three-term recurrence
//            T_{m+1} = 2t T_m - T_{m-1} in fp32, the basis values as IEEE half into a shared-memory A tile
//            [128 pairs][16 (slots + distance tiles)] (row pitch = 16 (mod 32) bytes: conflict-free for ldmatrix);
//            in the scores mode also the query row's folded query (and u) as ready-made mma B fragments, copied
//            asynchronously first and converted in place last;
//   8 consumer warps (setmaxnreg 96): nothing but ldmatrix / mma.sync.m16n8k16 (half accumulators) / max.f16x2 /
//            the score contraction / stores.  The channel <-> fragment-column assignment is chosen so that a thread's 8
//            accumulator columns are 8 consecutive channels: a 16-byte store of E per row and thread (K2c), or the A
//            fragment of the score contraction as it stands (K3r).  Two layouts: the split instances of K3r give a
//            warp one m-tile and let it walk the channel blocks (consume_rows); everything else gives a warp one
//            channel block of all m-tiles with G in registers and partial scores in shared memory (consume).
//   hand-over through four mbarriers (full / empty x 2 buffers); the consumer warps never wait for each other.
// Nothing of size n x n x k x C exists; shared memory carries 160 B per pair instead of K2t's 5 KB of table reads and
// the accumulators are never read back from TMEM (K2's epilogue moves 4 KB per pair at TMEM's 64 B/clk).
// The distance range is not known on the host without a synchronisation: geotx_gse_knn leaves the largest squared
// distance of the cloud behind the geometry scratch, and the kernel picks the coefficient set (K_d) from it.
// Measurements, ablations and the history of the kernel: DESIGN.md 4.2, profiles/r02_k3r_ablations.txt.
#include <cuda_fp16.h>
#include "geotx_common.cuh"
#include "host_util.h"
#include "sm100_ptx.cuh"

namespace geotx {

#ifndef CHEB_UNROLL
#define CHEB_UNROLL 2  // m-tiles of a warp unrolled together (timing experiments: -DCHEB_UNROLL=1|2|4)
#endif
#ifndef CHEB_MIN_CTAS
#define CHEB_MIN_CTAS 2
#endif
constexpr int CHEB_CONSUMER_WARPS = 8, CHEB_PRODUCER_WARPS = 4;
constexpr int CHEB_THREADS = 32 * (CHEB_CONSUMER_WARPS + CHEB_PRODUCER_WARPS);
constexpr int CHEB_TILE = 32 * CHEB_PRODUCER_WARPS;  // pairs per tile: one producer thread per pair
constexpr int CHEB_MT = CHEB_TILE / 16;              // m-tiles (16 pairs) per tile
constexpr int CHEB_NUM_D_SETS = 6;
__host__ __device__ constexpr int cheb_kd_tiles(int s) { return s == 0 ? 1 : s == 1 ? 2 : s == 2 ? 3 : s == 3 ? 4 : s == 4 ? 6 : 8; }
// largest argument the set is built for (tail of the expansion of a unit sinusoid <= 1e-5, see the header)
__host__ __device__ constexpr float cheb_range(int s) {
    return s == 0 ? 13.f : s == 1 ? 37.f : s == 2 ? 64.f : s == 3 ? 93.f : s == 4 ? 150.f : 210.f;
}
constexpr int CHEB_BUILD_NODES = 256;          // Gauss-Chebyshev nodes of the coefficient quadrature (> 2 K_max)
constexpr int CHEB_STATS_PER_SET = 4;          // max |error|, max |f|, max |G|, unused

// halves of one coefficient set in the register-image layout [ktile][C/32][4 n-tiles][32 lanes][2 regs][2 halves]
__host__ __device__ constexpr size_t cheb_set_halves(int C, int ktiles) { return (size_t)ktiles * C * 16; }

__host__ inline size_t cheb_set_offset(int C, int ka_tiles, int set /* 0 = angle, 1.. = distance sets */) {
    size_t off = 0;
    if (set >= 1) off += cheb_set_halves(C, ka_tiles);
    for (int s = 1; s < set; ++s) off += cheb_set_halves(C, cheb_kd_tiles(s - 1));
    return off;  // in halves
}

// ---- coefficient build (once per weight set) ------------------------------------------------------------------------
// One block per output channel.  F_q = f_c(x_q) at the Gauss-Chebyshev nodes in fp64, G_m = (2/M) sum_q F_q cos(m theta_q)
// (discrete orthogonality: exact for the interpolant of degree M-1, of which the first K terms are kept), rounded to
// half and scattered into the mma B-fragment register image.  Then the half-rounded expansion is compared with f_c at
// M points between the nodes (both ends included): the largest deviation, the largest |f| and the largest |G| go to
// stats (atomicMax on the bit patterns of non-negative floats), which the host reads once per weight set.
__global__ void __launch_bounds__(CHEB_BUILD_NODES)
cheb_build_kernel(const float* __restrict__ div_term, const float* __restrict__ w, const float* __restrict__ bias0,
                  const float* __restrict__ bias1, int C, float L, int K, __half* __restrict__ frag,
                  unsigned int* __restrict__ stats, float* __restrict__ plain) {
    __shared__ double F[CHEB_BUILD_NODES];
    __shared__ double Gh[128];
    __shared__ float red[3][CHEB_BUILD_NODES / 32];
    const int c = blockIdx.x, q = threadIdx.x;
    constexpr int M = CHEB_BUILD_NODES;
    const double bias = (bias0 ? (double)bias0[c] : 0.0) + (bias1 ? (double)bias1[c] : 0.0);
    auto f = [&](double x) {
        double acc = bias;
        for (int t = 0; t < C / 2; ++t) {
            double sn, cs;
            sincos(x * (double)div_term[t], &sn, &cs);
            acc += (double)w[(size_t)c * C + 2 * t] * sn + (double)w[(size_t)c * C + 2 * t + 1] * cs;
        }
        return acc;
    };
    F[q] = f((cospi((q + 0.5) / M) + 1.0) * 0.5 * (double)L);
    __syncthreads();
    float gmax = 0.f;
    if (q < K) {
        double acc = 0.0;
        for (int j = 0; j < M; ++j) acc += F[j] * cospi((double)q * (j + 0.5) / M);
        acc *= (q == 0 ? 1.0 : 2.0) / M;
        const __half h = __double2half(acc);
        if (plain) plain[(size_t)q * C + c] = (float)acc;  // row-major [K][C] copy for host-side folds (distance split)
        Gh[q] = (double)__half2float(h);
        gmax = fabsf((float)acc);
        // fragment image: channel c = 32 cb + 8 qq + 2 t + e  <->  n-tile t, B column n = 2 qq + e;  row kk = q % 16 of
        // k-tile q / 16 lives in lane n * 4 + (kk % 8) / 2, register kk / 8, half kk & 1
        const int cb = c >> 5, cc = c & 31, qq = cc >> 3, t = (cc & 7) >> 1, e = cc & 1, n = 2 * qq + e;
        const int kt = q >> 4, kk = q & 15, lane = n * 4 + ((kk & 7) >> 1), reg = kk >> 3, hi = kk & 1;
        frag[((((size_t)kt * (C / 32) + cb) * 4 + t) * 32 + lane) * 4 + reg * 2 + hi] = h;
    }
    __syncthreads();
    // check at theta = q pi / (M - 1): q = 0 is x = L, q = M - 1 is x = 0
    const double th = (double)q / (M - 1);
    const double exact = f((cospi(th) + 1.0) * 0.5 * (double)L);
    double approx = 0.0;
    for (int m = 0; m < K; ++m) approx += Gh[m] * cospi(m * th);
    float err = (float)fabs(approx - exact), fmx = (float)fabs(exact);
    for (int off = 16; off > 0; off >>= 1) {
        err = fmaxf(err, __shfl_xor_sync(0xffffffffu, err, off));
        fmx = fmaxf(fmx, __shfl_xor_sync(0xffffffffu, fmx, off));
        gmax = fmaxf(gmax, __shfl_xor_sync(0xffffffffu, gmax, off));
    }
    if ((q & 31) == 0) { red[0][q >> 5] = err; red[1][q >> 5] = fmx; red[2][q >> 5] = gmax; }
    __syncthreads();
    if (q < 3) {
        float v = 0.f;
        for (int i = 0; i < M / 32; ++i) v = fmaxf(v, red[q][i]);
        atomicMax(stats + q, __float_as_uint(v));
    }
}

// ---- the embedding kernel -------------------------------------------------------------------------------------------
__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// d = a b (fresh accumulator)
__device__ __forceinline__ void mma_new(float (&d)[4], const uint32_t (&a)[4], const uint2& b) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y), "f"(0.f));
}
// d += a b
__device__ __forceinline__ void mma_acc(float (&d)[4], const uint32_t (&a)[4], const uint2& b) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}
// half accumulators (fast instances): d = a b and d += a b with D/C as two half2 registers (rows g and g + 8)
__device__ __forceinline__ void mma_h_new(uint32_t (&d)[2], const uint32_t (&a)[4], const uint2& b) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f16.f16.f16.f16 {%0,%1}, {%2,%3,%4,%5}, {%6,%7}, {%8,%8};"
                 : "=r"(d[0]), "=r"(d[1])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y), "r"(0u));
}
__device__ __forceinline__ void mma_h_acc(uint32_t (&d)[2], const uint32_t (&a)[4], const uint2& b) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f16.f16.f16.f16 {%0,%1}, {%2,%3,%4,%5}, {%6,%7}, {%0,%1};"
                 : "+r"(d[0]), "+r"(d[1])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}
__device__ __forceinline__ uint32_t hmax2_u(uint32_t x, uint32_t y) {
    uint32_t r;
    asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(y));
    return r;
}
__device__ __forceinline__ uint32_t hadd2_u(uint32_t x, uint32_t y) {
    uint32_t r;
    asm("add.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(y));
    return r;
}
__device__ __forceinline__ uint32_t pack_half2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ void st_na_v4(void* p, const uint4& v) {
    asm volatile("st.global.L1::no_allocate.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ void sts_v4(uint32_t addr, const uint4& v) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// named barriers between the producer and the consumer warps (all CHEB_THREADS threads take part: the signalling side
// arrives, the waiting side syncs)
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(uint32_t dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}
// Producer <-> consumer hand-over of a tile buffer.  Default: four mbarriers in shared memory (full[2]: 128 producer
// threads arrive, each consumer warp waits on its own; empty[2]: lane 0 of each of the 8 consumer warps arrives, the
// producers wait).  Nothing makes the consumer warps wait for EACH OTHER: with the named barriers of the first version
// (bar.sync over all 384 threads; kept under CHEB_NAMED_BARRIERS for A/B timing) every tile restarted the CTA's eight
// consumer warps in lockstep, and the tensor pipe idled through each ramp-down / ramp-up.
constexpr int CHEB_BAR_BYTES = 64;
#ifndef CHEB_SPLIT_BUFFERS
#define CHEB_SPLIT_BUFFERS 3
#endif
// tile buffers of a split instance with kd distance k-tiles: 3 where two CTAs still fit an SM, else 2 (set 3: 240-byte rows)
__host__ __device__ constexpr int cheb_split_buffers(int kd) { return kd >= 4 ? 2 : CHEB_SPLIT_BUFFERS; }
constexpr int CHEB_PSTRIDE = CHEB_TILE + 8;  // floats between the heads of a channel block's partial scores
__device__ __forceinline__ void cheb_bar_sync(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(CHEB_THREADS) : "memory"); }
__device__ __forceinline__ void cheb_bar_arrive(int id) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "n"(CHEB_THREADS) : "memory"); }

// 16 * TILES Chebyshev values of t (T_0 = 1, T_1 = t, T_{m+1} = 2 t T_m - T_{m-1}) as halves, to smem at `addr`
template <int TILES>
__device__ __forceinline__ void cheb_store(uint32_t addr, float t) {
    // (measured: the stride-2 form of the recurrence with one packed FFMA2 per half2 instead of two FFMAs is not
    // faster: n = 4096 1.693 ms against 1.662 ms)
    const float t2 = t + t;
    float a = 1.f, b = t;  // T_m, T_{m+1}
#pragma unroll
    for (int blk = 0; blk < 2 * TILES; ++blk) {
        uint32_t w[4];
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            w[h] = pack_half2(a, b);
            a = fmaf(t2, b, -a);
            b = fmaf(t2, a, -b);
        }
        sts_v4(addr + blk * 16, make_uint4(w[0], w[1], w[2], w[3]));
    }
}
// runtime tile count (generic instance)
__device__ __forceinline__ void cheb_store_rt(uint32_t addr, float t, int tiles) {
    const float t2 = t + t;
    float a = 1.f, b = t;
    for (int blk = 0; blk < 2 * tiles; ++blk) {
        uint32_t w[4];
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            w[h] = pack_half2(a, b);
            a = fmaf(t2, b, -a);
            b = fmaf(t2, a, -b);
        }
        sts_v4(addr + blk * 16, make_uint4(w[0], w[1], w[2], w[3]));
    }
}
// the same, accumulated in registers (for the 'mean' reduction): acc[m] += scale * T_m(t)
template <int TILES>
__device__ __forceinline__ void cheb_accumulate(float (&acc)[16 * TILES], float t, float scale) {
    const float t2 = t + t;
    float a = scale, b = t * scale;
#pragma unroll
    for (int m = 0; m < 16 * TILES; m += 2) {
        acc[m] += a;
        acc[m + 1] += b;
        a = fmaf(t2, b, -a);
        b = fmaf(t2, a, -b);
    }
}

// One embedding argument of a pair: the quantities of geotx_common.cuh (pair_argument_one / get_embedding_indices,
// geotransformer.py:38-53) without their IEEE slow paths -- the producers are issue-bound, and the basis values built
// from the argument are rounded to half (2.4e-4) anyway: sqrt as x rsqrt(x) (2 ulp), the division by sigma_d as a
// multiplication, atan2 as a branch-free minimax polynomial (7 odd terms on [0,1], |error| < 3.4e-7 rad = 1.3e-6 index
// units at sigma_a = 15).  atan2(0, 0) = 0 like the reference's (j = i).
__device__ __forceinline__ float cheb_sqrt(float x) { return x * rsqrtf(fmaxf(x, 1e-30f)); }

__device__ __forceinline__ float cheb_argument(const float4& pi, const float4& pj, const float4* __restrict__ refv_i, int k,
                                               int s, float inv_sigma_d, float factor_a) {
    if (s < k) {
        const float ax = pj.x - pi.x, ay = pj.y - pi.y, az = pj.z - pi.z;
        const float4 r = __ldg(refv_i + s);
        const float cx = r.y * az - r.z * ay, cy = r.z * ax - r.x * az, cz = r.x * ay - r.y * ax;
        const float sn = cheb_sqrt(cx * cx + cy * cy + cz * cz), cs = r.x * ax + r.y * ay + r.z * az;
        const float ac = fabsf(cs), mx = fmaxf(sn, ac), mn = fminf(sn, ac);
        const float a = mx > 0.f ? __fdividef(mn, mx) : 0.f, a2 = a * a;
        float t = fmaf(a2, 0.006811776198446751f, -0.033604178577661514f);
        t = fmaf(a2, t, 0.07962363958358765f);
        t = fmaf(a2, t, -0.1323334127664566f);
        t = fmaf(a2, t, 0.19807815551757812f);
        t = fmaf(a2, t, -0.3331736922264099f);
        t = fmaf(a2, t, 0.9999961256980896f);
        t *= a;
        t = sn > ac ? 1.5707963267948966f - t : t;
        t = cs < 0.f ? 3.141592653589793f - t : t;
        return t * factor_a;
    }
    return cheb_sqrt(sqdist_ref(pi, pj)) * inv_sigma_d;
}

struct ChebArgs {
    const float4* pts4;
    const float4* refv;
    const unsigned int* dmax2_bits;  // largest squared distance of the clouds (geotx_gse_knn)
    const __half* coef;              // angle set, then the distance sets
    void* out;                       // E (f16 / bf16), or scores_p (f32) in the scores mode
    int* overflow;
    long long pairs_total;           // E mode: pairs of the slab, flattened
    long long tiles;
    int n, k, slots, reduce_mean, row_begin, row_count, out_bf16, max_d_set;
    float sigma_d, inv_sigma_d, factor_a, inv_half_la;  // 2 / L_a
    unsigned int d_set_offset[CHEB_NUM_D_SETS];  // halves from coef to distance set s
    // scores mode (K3r): folded query qt[b,h,i,c] with element strides, H heads, key blocks per row
    const float* qt;
    long long qsb, qsh, qsn;
    int H, key_blocks;
    // distance split (scores mode, optional): u[b,h,i, u_offset[set] + m] = sum_c qt[b,h,i,c] G_d^{set}[m,c]
    const float* u;
    long long usb, ush, usn;
    int u_offset[CHEB_NUM_D_SETS];
};

// saturating hi / lo split of two floats into halves (the folded query enters the score contraction as hi + lo, like
// geotx_rpe_scores, so that only E carries a 16-bit rounding)
__device__ __forceinline__ void cheb_split(float x, float y, uint32_t& hi, uint32_t& lo) {
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(y), "f"(x));
    const float2 hf = __half22float2(*reinterpret_cast<const __half2*>(&hi));
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(y - hf.y), "f"(x - hf.x));
}

// SCORES = false: E of the slab is written (geotx_gse_embed_cheb); tiles are 128 consecutive flattened pairs.
// SCORES = true (K3r, geotx_rpe_scores_cheb): E never leaves the registers.  A tile is 128 consecutive keys j of ONE
// query row (b, i); the finished accumulators (rounded to half, exactly what the stored E would hold) become the A
// operand of a second contraction with that row's folded query (B operand: channels x heads, hi + lo halves), each
// warp producing the partial scores of its 32 channels; the 8 partials meet in shared memory and the tile's
// H x 128 scores are written as 512-byte rows of scores_p.
// KD, SL > 0: compile-time numbers of distance K tiles and angle slots (the fast instances: k = 3 with 'max', distance sets 1
// and 2, i.e. clouds up to 64 index units across -- every shipped configuration); G stays in registers, accumulators are
// halves.  KD = SL = 0: the generic instance (any k, 'mean', bf16 output, distance sets 3..5): runtime loops, fp32
// accumulators, G read through L1 for every m-tile.
//
// Warp-specialised CTA of 12 warps over tiles of 128 pairs (8 m-tiles), two tiles in flight:
//   warps 8-11  producers (48 registers): one thread per pair computes the pair's arguments, runs the recurrences and
//               writes the pair's row of the A tile; in the scores mode they also stage the tile row's folded query as
//               mma B fragments and, one tile later, add up the consumers' partial scores and write scores_p
//   warps 0-7   consumers (96 registers, setmaxnreg): warp w owns 32 channels (C = 256: channel block w of all 8 m-tiles;
//               C = 128: block w % 4 of four m-tiles); they do nothing but ldmatrix / mma / max / add / store, so that
//               the tensor pipe never waits for a produce phase
// Hand-over through named barriers (1, 2: tile buffer b is full; 3, 4: buffer b has been consumed), bar.arrive on the
// signalling side, bar.sync on the waiting side.
// Scores mode: a CTA's tiles blockIdx.x, blockIdx.x + gridDim.x, ... as (cloud b, query row i, key block kb), advanced
// by additions (the two integer divisions of tile -> (b, i, kb) cost ~50 instructions per tile and thread otherwise).
struct ChebTileWalk {
    unsigned b, i, kb, kb_step, row_step, key_blocks, row_count;
    __device__ __forceinline__ void start(const ChebArgs& p) {
        key_blocks = (unsigned)p.key_blocks;
        row_count = (unsigned)p.row_count;
        const unsigned row = blockIdx.x / key_blocks;
        kb = blockIdx.x - row * key_blocks;
        b = row / row_count;
        i = row - b * row_count;
        row_step = gridDim.x / key_blocks;
        kb_step = gridDim.x - row_step * key_blocks;
    }
    __device__ __forceinline__ void next() {
        kb += kb_step;
        i += row_step;
        if (kb >= key_blocks) { kb -= key_blocks; ++i; }
        while (i >= row_count) { i -= row_count; ++b; }
    }
};

template <int C, int KA, int KD, bool SCORES, int SL, bool SPLIT = false>
__device__ __forceinline__ void cheb_run(const ChebArgs& p, int d_set, unsigned char* smem) {
    constexpr bool FAST = KD > 0 && SL > 0;
    static_assert(!SPLIT || (SCORES && FAST), "the distance split exists in the fast scores instances only");
    constexpr int NCB = C / 32;                          // 32-channel blocks
    constexpr int MT_PER_WARP = CHEB_MT * NCB / CHEB_CONSUMER_WARPS;  // 8 (C = 256) or 4 (C = 128)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // tile buffers in flight: 2, or 3 for the split instances (their shared memory has the room since the partial scores
    // went: the producers run up to two tiles ahead; measured: producers waited for `empty` a third of their time and the
    // consumers for `full` 8 % of theirs with two buffers)
    constexpr int NB = SPLIT ? cheb_split_buffers(KD) : 2;
    const uint32_t bar_full = smem_u32(smem), bar_empty = bar_full + 8 * NB;  // + 8 * buf
    smem += CHEB_BAR_BYTES;
    const int slots = FAST ? SL : p.slots;
    const int kd_tiles = FAST ? KD : cheb_kd_tiles(d_set);
    const int ktot = slots * 16 * KA + 16 * kd_tiles;  // halves per A row
    const uint32_t pitch = ktot * 2 + 16;               // bytes: = 16 (mod 32), ldmatrix rows fall into distinct banks
    const uint32_t a_base = smem_u32(smem);
    const uint32_t buf_bytes = CHEB_TILE * pitch;
    // scores mode, behind the two A tiles.
    //   plain instances: partial[2][NCB][hp][136] floats (every channel block's share of the scores) and
    //     qimg[2][NCB][32 lanes][8 words]: the tile row's folded query as mma B fragments, hi | lo halves
    //   SPLIT instances (H <= 4; a consumer warp owns an m-tile and walks the channel blocks, see consume_rows):
    //     gimg[KA][NCB][2][32 lanes][4 words]: the angle coefficients as B fragments, two n-tiles per 16 bytes;
    //     qimg[2][NCB][32 lanes][4 words]: the folded query PACKED -- fragment columns 0-3 = the heads' hi halves,
    //     columns 4-7 = their lo halves, so that one HMMA contracts both; uimg[2][KD][32 lanes][4 words]: u[h][m]
    const int hp = p.H <= 4 ? 4 : 8;
    float* const partial = reinterpret_cast<float*>(smem + NB * buf_bytes);
    const int partial_floats = NCB * hp * CHEB_PSTRIDE;  // head stride 136 floats: the stores of heads 2q | 2q + 2 fall into different banks
    constexpr int GIMG_BYTES = KA * NCB * 2 * 32 * 16;
    const uint32_t gimg = a_base + NB * buf_bytes;
    const uint32_t qimg = SPLIT ? gimg + GIMG_BYTES : a_base + NB * buf_bytes + 2 * partial_floats * 4;
    constexpr int QIMG_BYTES = SPLIT ? NCB * 32 * 16 : NCB * 32 * 32;
    const uint32_t uimg = qimg + NB * QIMG_BYTES;
    constexpr int UIMG_BYTES = (KD > 0 ? KD : 1) * 32 * 16;
    const long long tiles = p.tiles;
    const long long per_b = (long long)p.row_count * p.n;

    // tile -> (cloud b, query row i, key j) of the tile's pair `local`.  32-bit arithmetic whenever the slab has fewer
    // than 2^32 pairs (64-bit divisions cost ~100 instructions each)
    const bool small = (unsigned long long)p.pairs_total < 0xffffffffull;
    auto locate = [&](long long tile, int local, long long& b, int& i, int& j) {
        if (SCORES) {
            const unsigned row = (unsigned)tile / (unsigned)p.key_blocks;
            const unsigned bb = row / (unsigned)p.row_count;
            b = bb;
            i = p.row_begin + (int)(row - bb * (unsigned)p.row_count);
            j = min((int)((unsigned)tile - row * (unsigned)p.key_blocks) * CHEB_TILE + local, p.n - 1);
        } else if (small) {
            const unsigned g = min((unsigned)tile * CHEB_TILE + (unsigned)local, (unsigned)p.pairs_total - 1u);
            const unsigned rowg = g / (unsigned)p.n;  // = b * row_count + local row
            const unsigned bb = rowg / (unsigned)p.row_count;
            b = bb;
            i = p.row_begin + (int)(rowg - bb * (unsigned)p.row_count);
            j = (int)(g - rowg * (unsigned)p.n);
        } else {
            const long long g = min(tile * CHEB_TILE + local, p.pairs_total - 1);
            b = g / per_b;
            const long long rem = g - b * per_b;
            i = p.row_begin + (int)(rem / p.n);
            j = (int)(rem % p.n);
        }
    };

    if (SCORES) {  // fragment lanes of heads >= H read zeros
        const int bytes = NB * QIMG_BYTES + (SPLIT ? NB * UIMG_BYTES : 0);
        for (int e = tid; e < bytes / 16; e += CHEB_THREADS) sts_v4(qimg + e * 16, make_uint4(0u, 0u, 0u, 0u));
    }
    if (SPLIT) {  // the angle coefficients, once per CTA: global image [kt][cb][t][lane] (8 bytes) -> [kt][cb][t / 2][lane] (16 bytes)
        const uint2* src = reinterpret_cast<const uint2*>(p.coef);
        for (int e = tid; e < KA * NCB * 4 * 32; e += CHEB_THREADS) {
            const int l = e & 31, t = (e >> 5) & 3, kc = e >> 7;  // kc = kt * NCB + cb
            const uint2 v = __ldg(src + e);
            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(gimg + ((kc * 2 + (t >> 1)) * 32 + l) * 16 + (t & 1) * 8), "r"(v.x), "r"(v.y) : "memory");
        }
    }
    if (tid == 0) {
        for (int b = 0; b < NB; ++b) {
            ptx::mbar_init(bar_full + 8 * b, CHEB_PRODUCER_WARPS * 32);
            ptx::mbar_init(bar_empty + 8 * b, CHEB_CONSUMER_WARPS);
        }
        ptx::fence_barrier_init();
    }
    __syncthreads();

    if (warp >= CHEB_CONSUMER_WARPS) {
        // ======================================== producers ========================================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 48;");
        const int ptid = tid - CHEB_CONSUMER_WARPS * 32;  // = the pair of the tile this thread builds
        const float inv_half_ld = 2.f / cheb_range(d_set);
        const float x_lim_d = cheb_range(d_set) * 1.0001f;
        bool clamped = false;

        ChebTileWalk walk;
        if (SCORES) walk.start(p);
        auto produce = [&](long long tile, int buf) {
            long long b;
            int i, j;
            if (SCORES) {
                b = walk.b;
                i = p.row_begin + (int)walk.i;
                j = min((int)walk.kb * CHEB_TILE + ptid, p.n - 1);
            } else {
                locate(tile, ptid, b, i, j);
            }
            const float4 pi = __ldg(p.pts4 + b * p.n + i), pj = __ldg(p.pts4 + b * p.n + j);
            const float4* refv_i = p.refv + (b * p.n + i) * p.k;
            const uint32_t row = a_base + buf * buf_bytes + ptid * pitch;
            constexpr int KD4 = KD > 0 ? KD * 4 : 1;
            if (SCORES) {
                // The tile row's folded query (and u) come from L2, a different row for every tile of this CTA: fetched
                // with plain loads at the point of use, that latency sat on the producers' critical path once per tile
                // (measured: 7 of their 17 clk per pair).  So the raw fp32 values are copied asynchronously, first thing,
                // into the very slots of the fragment images they will occupy (8 floats = the 32 bytes of hi | lo), and
                // the same thread converts them in place after the basis values are done.
                for (int it = ptid; it < p.H * (C / 8); it += CHEB_PRODUCER_WARPS * 32) {
                    const int h = it / (C / 8), grp = it - h * (C / 8);  // grp = 4 cb + q
                    const float* src = p.qt + b * p.qsb + (long long)h * p.qsh + (long long)i * p.qsn + grp * 8;
                    if (SPLIT) {  // hi slot = lane (g = h, q), lo slot = lane (g = h + 4, q): 16 bytes each
                        const uint32_t dst = qimg + buf * QIMG_BYTES + ((grp >> 2) * 32 + h * 4 + (grp & 3)) * 16;
                        cp_async16(dst, src);
                        cp_async16(dst + 256, src + 4);
                    } else {
                        const uint32_t dst = qimg + buf * QIMG_BYTES + ((grp >> 2) * 32 + h * 4 + (grp & 3)) * 32;
                        cp_async16(dst, src);
                        cp_async16(dst + 16, src + 4);
                    }
                }
                if (SPLIT && ptid < p.H * KD4) {
                    const int h = ptid / KD4, rem = ptid - h * KD4, kt = rem >> 2, qq = rem & 3;
                    const float* src = p.u + b * p.usb + (long long)h * p.ush + (long long)i * p.usn + p.u_offset[d_set] +
                                       16 * kt + 2 * qq;
                    const uint32_t dst = uimg + buf * UIMG_BYTES + (kt * 32 + h * 4 + qq) * 16;
                    cp_async8(dst, src);
                    cp_async8(dst + 8, src + 8);
                }
            }
            if (FAST) {  // k = SL slots, 'max': straight-line code, the recurrences of the slots interleave
#pragma unroll
                for (int s = 0; s < SL; ++s) {
                    const float x = cheb_argument(pi, pj, refv_i, SL, s, p.inv_sigma_d, p.factor_a);
                    cheb_store<KA>(row + s * (32 * KA), fminf(fmaxf(fmaf(x, p.inv_half_la, -1.f), -1.f), 1.f));
                }
            } else if (!p.reduce_mean) {
                for (int s = 0; s < slots; ++s) {
                    const float x = cheb_argument(pi, pj, refv_i, p.k, s, p.inv_sigma_d, p.factor_a);
                    cheb_store<KA>(row + s * (32 * KA), fminf(fmaxf(fmaf(x, p.inv_half_la, -1.f), -1.f), 1.f));
                }
            } else {  // geotransformer.py:68: the mean over the k neighbours is linear in the basis
                float acc[16 * KA];
#pragma unroll
                for (int m = 0; m < 16 * KA; ++m) acc[m] = 0.f;
                const float scale = 1.f / (float)p.k;
                for (int s = 0; s < p.k; ++s) {
                    const float x = cheb_argument(pi, pj, refv_i, p.k, s, p.inv_sigma_d, p.factor_a);
                    cheb_accumulate<KA>(acc, fminf(fmaxf(fmaf(x, p.inv_half_la, -1.f), -1.f), 1.f), scale);
                }
#pragma unroll
                for (int blk = 0; blk < 2 * KA; ++blk)
                    sts_v4(row + blk * 16, make_uint4(pack_half2(acc[8 * blk], acc[8 * blk + 1]), pack_half2(acc[8 * blk + 2], acc[8 * blk + 3]),
                                                      pack_half2(acc[8 * blk + 4], acc[8 * blk + 5]), pack_half2(acc[8 * blk + 6], acc[8 * blk + 7])));
            }
            {
                const float x = cheb_argument(pi, pj, refv_i, p.k, p.k, p.inv_sigma_d, p.factor_a);
                clamped |= !(x <= x_lim_d);
                const float t = fminf(fmaxf(fmaf(x, inv_half_ld, -1.f), -1.f), 1.f);
                if (FAST) cheb_store<FAST ? KD : 1>(row + slots * (32 * KA), t);
                else cheb_store_rt(row + slots * (32 * KA), t, kd_tiles);
            }
            if (SCORES) {
                // the folded query of the tile's row as B fragments of the score contraction: fragment column n = head;
                // lane (g = head, q) of channel block cb holds the 8 consecutive channels 32 cb + 8 q .. + 7 (the channel
                // <-> k-slot assignment of the accumulators that become the A operand), as two half2 pairs per 16-channel
                // k-tile, hi then lo.  Item = (head, 8-channel group); lanes of heads >= H stay zero.  The raw values
                // were copied into the item's own slot above, by this thread.
                asm volatile("cp.async.wait_all;" ::: "memory");
                for (int it = ptid; it < p.H * (C / 8); it += CHEB_PRODUCER_WARPS * 32) {
                    const int h = it / (C / 8), grp = it - h * (C / 8);
                    const uint32_t dst = qimg + buf * QIMG_BYTES + ((grp >> 2) * 32 + h * 4 + (grp & 3)) * (SPLIT ? 16 : 32);
                    const uint32_t dst_lo = dst + (SPLIT ? 256 : 16);
                    const uint4 ua = lds_v4(dst), va = lds_v4(dst_lo);
                    uint4 hi, lo;
                    cheb_split(__uint_as_float(ua.x), __uint_as_float(ua.y), hi.x, lo.x);
                    cheb_split(__uint_as_float(ua.z), __uint_as_float(ua.w), hi.y, lo.y);
                    cheb_split(__uint_as_float(va.x), __uint_as_float(va.y), hi.z, lo.z);
                    cheb_split(__uint_as_float(va.z), __uint_as_float(va.w), hi.w, lo.w);
                    sts_v4(dst, hi);
                    sts_v4(dst_lo, lo);
                }
                if (SPLIT && ptid < p.H * KD4) {
                    // u[h][m] = sum_c qt[h][c] G_d[m][c] (it arrives with the query: the caller's projection produces it)
                    // turns the distance term's share of the scores into a K_d-term product with the basis values that
                    // are in the A tile anyway.  Item = (head, k-tile, q): B rows 2q, 2q+1 | 2q+8, 2q+9 of the k-tile.
                    const int h = ptid / KD4, rem = ptid - h * KD4, kt = rem >> 2, qq = rem & 3;
                    const uint32_t dst = uimg + buf * UIMG_BYTES + (kt * 32 + h * 4 + qq) * 16;
                    const uint4 raw = lds_v4(dst);  // rows 2q, 2q+1 | rows 2q+8, 2q+9
                    uint4 w;  // hi(rows 2q..) | hi(rows 2q+8..) | lo | lo
                    cheb_split(__uint_as_float(raw.x), __uint_as_float(raw.y), w.x, w.z);
                    cheb_split(__uint_as_float(raw.z), __uint_as_float(raw.w), w.y, w.w);
                    sts_v4(dst, w);
                }
            }
        };
        // the channel blocks' partial scores of a consumed tile -> scores_p[b, h, i, j0 .. j0 + CHEB_TILE - 1].  A task is
        // (head, 4 consecutive keys): NCB 16-byte reads, one 16-byte store (the first version read and added one float per
        // instruction and spent 37 % of the producers' instructions here)
        const bool out_vec = (p.n & 3) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0;
        auto reduce = [&](long long tile, int buf) {
            long long b;
            int i, j0;
            locate(tile, 0, b, i, j0);
            const float* src = partial + buf * partial_floats;
            float* const out = reinterpret_cast<float*>(p.out) + (((size_t)b * p.H) * p.n + i) * p.n + j0;
            for (int task = ptid; task < p.H * (CHEB_TILE / 4); task += CHEB_PRODUCER_WARPS * 32) {
                const int h = task / (CHEB_TILE / 4), jl = (task - h * (CHEB_TILE / 4)) * 4;
                float4 v = *reinterpret_cast<const float4*>(src + h * CHEB_PSTRIDE + jl);
#pragma unroll
                for (int w = 1; w < NCB; ++w) {
                    const float4 t = *reinterpret_cast<const float4*>(src + (w * hp + h) * CHEB_PSTRIDE + jl);
                    v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
                }
                float* dst = out + (size_t)h * p.n * p.n + jl;
                if (out_vec) {  // n % 4 == 0: a group of 4 keys is inside the row or outside it as a whole
                    if (j0 + jl < p.n) *reinterpret_cast<float4*>(dst) = v;
                } else {
                    if (j0 + jl < p.n) dst[0] = v.x;
                    if (j0 + jl + 1 < p.n) dst[1] = v.y;
                    if (j0 + jl + 2 < p.n) dst[2] = v.z;
                    if (j0 + jl + 3 < p.n) dst[3] = v.w;
                }
            }
        };

        int it = 0;
        long long tile = blockIdx.x, older = 0, old = 0;  // the tiles of the two previous iterations
        for (; tile < tiles; tile += gridDim.x, ++it) {
            const int buf = it % NB;
            if (it >= NB) {
                // the consumers are done with tile it - NB: A[buf] is free, partial[buf] complete
#ifdef CHEB_NAMED_BARRIERS
                cheb_bar_sync(3 + buf);
#else
                ptx::mbar_wait(bar_empty + 8 * buf, ((it / NB) - 1) & 1, 0x100);
#endif
#if !defined(CHEB_ABL_NO_REDUCE) && !defined(CHEB_ABL_NO_PRODUCE)  // timing ablations (tools/build_variants.sh), never in the shipped library
                if (SCORES && !SPLIT) reduce(older, buf);  // SPLIT: the consumers write scores_p themselves
#endif
            }
#ifndef CHEB_ABL_NO_PRODUCE
            produce(tile, buf);
#endif
            if (SCORES) walk.next();
#ifdef CHEB_NAMED_BARRIERS
            __threadfence_block();
            cheb_bar_arrive(1 + buf);
#else
            ptx::mbar_arrive(bar_full + 8 * buf);  // release: this thread's rows and fragment images are visible to the waiters
#endif
            older = old;
            old = tile;
        }
        // drain: the last two tiles' scores (plain instances: NB = 2; the split instances have nothing left to do)
        if (!SPLIT && it >= 2) {
#ifdef CHEB_NAMED_BARRIERS
            cheb_bar_sync(3 + (it & 1));
#else
            ptx::mbar_wait(bar_empty + 8 * (it & 1), ((it - 2) >> 1) & 1, 0x101);
#endif
            if (SCORES && !SPLIT) reduce(older, it & 1);
        }
        if (!SPLIT && it >= 1) {
#ifdef CHEB_NAMED_BARRIERS
            cheb_bar_sync(3 + ((it + 1) & 1));
#else
            ptx::mbar_wait(bar_empty + 8 * ((it + 1) & 1), ((it - 1) >> 1) & 1, 0x102);
#endif
            if (SCORES && !SPLIT) reduce(old, (it + 1) & 1);
        }
        if (clamped) *p.overflow = 1;
    } else {
        // ======================================== consumers ========================================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 96;");
        const int cb = warp % NCB, mt0 = (warp / NCB) * MT_PER_WARP;
        // B operand: this warp's 32 channels of G
        const uint2* const g_a = reinterpret_cast<const uint2*>(p.coef) + cb * 128 + lane;  // + (kt * NCB * 4 + t) * 32
        const uint2* const g_d = reinterpret_cast<const uint2*>(p.coef + p.d_set_offset[d_set]) + cb * 128 + lane;
        uint2 b_a[KA][4], b_d[FAST ? KD : 1][4];
        if (FAST && !SPLIT) {
#pragma unroll
            for (int kt = 0; kt < KA; ++kt)
#pragma unroll
                for (int t = 0; t < 4; ++t) b_a[kt][t] = __ldg(g_a + (kt * NCB * 4 + t) * 32);
#pragma unroll
            for (int kt = 0; kt < KD; ++kt)
#pragma unroll
                for (int t = 0; t < 4; ++t) b_d[kt][t] = __ldg(g_d + (kt * NCB * 4 + t) * 32);
        }
        // ldmatrix: lane l addresses row (l & 15) of the m-tile, columns 8 (l >> 4) ..; mma: g = lane / 4 is the fragment
        // row, q = lane % 4 the column pair
        const uint32_t ld_off = (lane & 15) * pitch + (lane >> 4) * 16;
        const int g = lane >> 2, q = lane & 3;
        unsigned char* const out_bytes = reinterpret_cast<unsigned char*>(p.out);

        auto consume = [&](long long tile, int buf) {
            const uint32_t tile_addr = a_base + buf * buf_bytes + ld_off;
            const long long pair0 = tile * CHEB_TILE;
            uint4 qh, ql;  // the tile row's folded query, this warp's channel block (hi | lo)
            if (SCORES) {
                const uint32_t src = qimg + buf * QIMG_BYTES + (cb * 32 + lane) * 32;
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(qh.x), "=r"(qh.y), "=r"(qh.z), "=r"(qh.w) : "r"(src));
                asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(ql.x), "=r"(ql.y), "=r"(ql.z), "=r"(ql.w) : "r"(src + 16));
            }
#ifndef CHEB_NO_PIPELINE  // timing ablation switch (tools/build_variants.sh)
            constexpr bool PIPELINED = FAST && SCORES && !SPLIT;
#else
            constexpr bool PIPELINED = false;
#endif
            if constexpr (PIPELINED) {
                // Software-pipelined by one m-tile, fully unrolled: the 12 expansion HMMAs of m-tile t+1 are issued BEFORE
                // the score contraction of m-tile t, and the partial scores of t are stored one step later still.  Run
                // back to back (expansions -> max -> two dependent score HMMAs -> add -> store -> next ldmatrix), every
                // m-tile left the warp ~140 clocks without an HMMA to issue for 128 clocks of HMMA work, and four such
                // warps kept the pipe 61 % busy with the producers switched off (tools/build_variants.sh ablation).
                constexpr int AD = SPLIT ? 0 : KD;  // without the split the distance term is expanded into E like the angles
                uint32_t a[SL * KA + AD][4], c[SL][4][2], cd[4][2];
                float sc[4], sc1[4], scp[4], sc1p[4];
                auto load_a = [&](int mt) {
                    const uint32_t addr = tile_addr + mt * 16 * pitch;
#pragma unroll
                    for (int x = 0; x < SL * KA + AD; ++x) ldsm4(a[x], addr + x * 32);
                };
                auto expand = [&]() {
#pragma unroll
                    for (int sl = 0; sl < SL; ++sl)
#pragma unroll
                        for (int kt = 0; kt < KA; ++kt)
#pragma unroll
                            for (int t = 0; t < 4; ++t) {
                                if (kt == 0) mma_h_new(c[sl][t], a[sl * KA + kt], b_a[kt][t]);
                                else mma_h_acc(c[sl][t], a[sl * KA + kt], b_a[kt][t]);
                            }
#pragma unroll
                    for (int kt = 0; kt < AD; ++kt)
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            if (kt == 0) mma_h_new(cd[t], a[SL * KA + kt], b_d[kt][t]);
                            else mma_h_acc(cd[t], a[SL * KA + kt], b_d[kt][t]);
                        }
                };
                float sink = 0.f;  // timing ablations only
                auto store_partial = [&](int mt, const float (&s0)[4], const float (&s1)[4]) {
#ifdef CHEB_ABL_NO_PSTORE
                    sink += (s0[0] + s1[0]) + (s0[1] + s1[1]) + (s0[2] + s1[2]) + (s0[3] + s1[3]);
                    return;
#endif
                    if (2 * q < hp) {  // thread (g, q): keys 16 mt + g and + 8, heads 2 q and 2 q + 1
                        float* dst = partial + (buf * NCB + cb) * hp * CHEB_PSTRIDE + (2 * q) * CHEB_PSTRIDE + mt * 16 + g;
                        dst[0] = s0[0] + s1[0];
                        dst[CHEB_PSTRIDE] = s0[1] + s1[1];
                        dst[8] = s0[2] + s1[2];
                        dst[CHEB_PSTRIDE + 8] = s0[3] + s1[3];
                    }
                };
                load_a(mt0);
                expand();
#pragma unroll
                for (int mi = 0; mi < MT_PER_WARP; ++mi) {
                    if (mi + 1 < MT_PER_WARP) load_a(mt0 + mi + 1);  // the HMMAs of m-tile mi have read `a` at issue
                    uint32_t e0[4], e1[4];  // max over the slots (geotransformer.py:66) = the score contraction's A operand
#pragma unroll
                    for (int t = 0; t < 4; ++t)
#pragma unroll
                        for (int r = 0; r < 2; ++r) {
                            uint32_t v = c[0][t][r];
#pragma unroll
                            for (int sl = 1; sl < SL; ++sl) v = hmax2_u(v, c[sl][t][r]);
                            if (AD > 0) v = hadd2_u(v, cd[t][r]);  // geotransformer.py:70
                            (t < 2 ? e0 : e1)[(t & 1) * 2 + r] = v;
                        }
                    if (mi + 1 < MT_PER_WARP) expand();
                    if (mi > 0) store_partial(mt0 + mi - 1, scp, sc1p);
#ifdef CHEB_ABL_NO_SCORE_MMA
#pragma unroll
                    for (int r = 0; r < 4; ++r) { sc[r] = __uint_as_float(e0[r]); sc1[r] = __uint_as_float(e1[r]); }
#else
                    mma_new(sc, e0, make_uint2(qh.x, qh.y));
                    mma_new(sc1, e1, make_uint2(qh.z, qh.w));
                    mma_acc(sc, e0, make_uint2(ql.x, ql.y));
                    mma_acc(sc1, e1, make_uint2(ql.z, ql.w));
#endif
#pragma unroll
                    for (int r = 0; r < 4; ++r) { scp[r] = sc[r]; sc1p[r] = sc1[r]; }
                }
                store_partial(mt0 + MT_PER_WARP - 1, scp, sc1p);
                if (sink == 12345.678f) partial[lane] = sink;
            } else {
            constexpr int kUnroll = FAST ? CHEB_UNROLL : 1;
#pragma unroll kUnroll
            for (int mi = 0; mi < MT_PER_WARP; ++mi) {
                const int mt = mt0 + mi;
                uint32_t addr = tile_addr + mt * 16 * pitch;
                uint32_t bh[4][2];  // E of the m-tile as half2: [n-tile][row g | row g + 8]
                if (FAST) {
                    // HALF accumulators: max over the slots (geotransformer.py:66) and the sum with the distance term
                    // (:70) as packed half2 operations; the registers are already in the layout of the store and of the
                    // score contraction's A operand
                    uint32_t a[4], cur[4][2], accd[4][2];
#pragma unroll
                    for (int sl = 0; sl < SL; ++sl) {
#pragma unroll
                        for (int kt = 0; kt < KA; ++kt) {
                            ldsm4(a, addr + (sl * KA + kt) * 32);
#pragma unroll
                            for (int t = 0; t < 4; ++t) {
                                if (kt == 0) mma_h_new(cur[t], a, b_a[kt][t]);
                                else mma_h_acc(cur[t], a, b_a[kt][t]);
                            }
                        }
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            bh[t][0] = sl == 0 ? cur[t][0] : hmax2_u(bh[t][0], cur[t][0]);
                            bh[t][1] = sl == 0 ? cur[t][1] : hmax2_u(bh[t][1], cur[t][1]);
                        }
                    }
                    if (!SPLIT) {
#pragma unroll
                        for (int kt = 0; kt < KD; ++kt) {
                            ldsm4(a, addr + (SL * KA + kt) * 32);
#pragma unroll
                            for (int t = 0; t < 4; ++t) {
                                if (kt == 0) mma_h_new(accd[t], a, b_d[kt][t]);
                                else mma_h_acc(accd[t], a, b_d[kt][t]);
                            }
                        }
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            bh[t][0] = hadd2_u(bh[t][0], accd[t][0]);
                            bh[t][1] = hadd2_u(bh[t][1], accd[t][1]);
                        }
                    }  // SPLIT: bh is the angle term alone; the distance term joins the SCORES below
                } else {
                    // generic instance: fp32 accumulators, runtime loops, G read through L1
                    uint32_t a[4];
                    float best[4][4];
                    for (int s = 0; s < slots; ++s) {
                        float cur[4][4];
#pragma unroll
                        for (int kt = 0; kt < KA; ++kt) {
                            ldsm4(a, addr);
                            addr += 32;
#pragma unroll
                            for (int t = 0; t < 4; ++t) {
                                const uint2 bb = __ldg(g_a + (kt * NCB * 4 + t) * 32);
                                if (kt == 0) mma_new(cur[t], a, bb);
                                else mma_acc(cur[t], a, bb);
                            }
                        }
#pragma unroll
                        for (int t = 0; t < 4; ++t)
#pragma unroll
                            for (int r = 0; r < 4; ++r) best[t][r] = s == 0 ? cur[t][r] : fmaxf(best[t][r], cur[t][r]);
                    }
                    float accd[4][4];
                    for (int kt = 0; kt < kd_tiles; ++kt) {
                        ldsm4(a, addr);
                        addr += 32;
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            const uint2 bb = __ldg(g_d + (kt * NCB * 4 + t) * 32);
                            if (kt == 0) mma_new(accd[t], a, bb);
                            else mma_acc(accd[t], a, bb);
                        }
                    }
                    const bool bf16 = !SCORES && p.out_bf16;
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const float v0 = best[t][0] + accd[t][0], v1 = best[t][1] + accd[t][1];
                        const float v2 = best[t][2] + accd[t][2], v3 = best[t][3] + accd[t][3];
                        bh[t][0] = bf16 ? pack_bf16x2(v0, v1) : pack_half2(v0, v1);
                        bh[t][1] = bf16 ? pack_bf16x2(v2, v3) : pack_half2(v2, v3);
                    }
                }
                if (SCORES) {
                    // E (half, exactly what a stored E would hold) x folded query (hi + lo): rows = the m-tile's 16 keys,
                    // columns = heads; two independent chains (channels 0-15 | 16-31 of the block)
                    float sc[4], sc1[4];
                    const uint32_t e0[4] = {bh[0][0], bh[0][1], bh[1][0], bh[1][1]};
                    const uint32_t e1[4] = {bh[2][0], bh[2][1], bh[3][0], bh[3][1]};
                    mma_new(sc, e0, make_uint2(qh.x, qh.y));
                    mma_new(sc1, e1, make_uint2(qh.z, qh.w));
                    mma_acc(sc, e0, make_uint2(ql.x, ql.y));
                    mma_acc(sc1, e1, make_uint2(ql.z, ql.w));
                    if (2 * q < hp) {  // thread (g, q): keys 16 mt + g and + 8, heads 2 q and 2 q + 1
                        float* dst = partial + (buf * NCB + cb) * hp * CHEB_PSTRIDE + (2 * q) * CHEB_PSTRIDE + mt * 16 + g;
                        dst[0] = sc[0] + sc1[0];
                        dst[CHEB_PSTRIDE] = sc[1] + sc1[1];
                        dst[8] = sc[2] + sc1[2];
                        dst[CHEB_PSTRIDE + 8] = sc[3] + sc1[3];
                    }
                } else {
                    // thread (g, q): rows g and g + 8 of the m-tile, channels 32 cb + 8 q .. + 7 (n-tile t holds 2 t, 2 t + 1)
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const long long pair = pair0 + mt * 16 + g + 8 * h;
                        if (pair < p.pairs_total)
                            st_na_v4(out_bytes + ((size_t)pair * C + cb * 32 + q * 8) * 2, make_uint4(bh[0][h], bh[1][h], bh[2][h], bh[3][h]));
                    }
                }
            }
            }  // !PIPELINED
        };


        // SPLIT instances: warp w owns m-tile w of the tile (16 keys) and walks the channel blocks.  Its A fragments
        // (the basis values) are loaded ONCE per tile instead of once per channel-block warp, the scores accumulate in
        // registers across the channel blocks (two chains), so there are no partial scores, no reduction pass and no
        // shared-memory round trip: the warp writes its 16 x H scores straight to scores_p.  Per channel block it reads
        // 48 bytes per lane of B fragments (G and the packed query) from shared memory.  Per pair: 0.3 ldmatrix + 1.5
        // LDS.128 instead of 2.5 ldmatrix + 2 STS + the reduction, and 7.25 HMMA instead of 8.25 (hi and lo halves of
        // the query ride in one HMMA's eight columns).
        ChebTileWalk cwalk;
        if (SPLIT) cwalk.start(p);
        auto consume_rows = [&](long long tile, int buf) {
            const uint32_t tile_addr = a_base + buf * buf_bytes + ld_off + warp * 16 * pitch;
            constexpr int SLR = SL > 0 ? SL : 1;  // (the generic instance, SL = 0, never calls this)
            uint32_t a[SLR * KA][4], c[SLR][4][2];
#pragma unroll
            for (int x = 0; x < SL * KA; ++x) ldsm4(a[x], tile_addr + x * 32);
            // the distance term u . t(d) first: its dependent chain runs in the shadow of the expansions
            float sc[4] = {0.f, 0.f, 0.f, 0.f}, sc1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int kt = 0; kt < KD; ++kt) {
                uint32_t ad[4];
                ldsm4(ad, tile_addr + (SL * KA + kt) * 32);
                const uint4 ub = lds_v4(uimg + buf * UIMG_BYTES + (kt * 32 + lane) * 16);
                mma_acc(sc, ad, make_uint2(ub.x, ub.y));
                mma_acc(sc1, ad, make_uint2(ub.z, ub.w));
            }
            uint4 gb[2][KA][2], qc[2];  // B fragments of channel block cb live in slot cb & 1
            auto load_b = [&](int cb) {
#pragma unroll
                for (int kt = 0; kt < KA; ++kt)
#pragma unroll
                    for (int tp = 0; tp < 2; ++tp) gb[cb & 1][kt][tp] = lds_v4(gimg + (((kt * NCB + cb) * 2 + tp) * 32 + lane) * 16);
                qc[cb & 1] = lds_v4(qimg + buf * QIMG_BYTES + (cb * 32 + lane) * 16);
            };
            auto expand = [&](int cb) {
#pragma unroll
                for (int sl = 0; sl < SL; ++sl)
#pragma unroll
                    for (int kt = 0; kt < KA; ++kt)
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            const uint4& w = gb[cb & 1][kt][t >> 1];
                            const uint2 bb = (t & 1) ? make_uint2(w.z, w.w) : make_uint2(w.x, w.y);
                            if (kt == 0) mma_h_new(c[sl][t], a[sl * KA + kt], bb);
                            else mma_h_acc(c[sl][t], a[sl * KA + kt], bb);
                        }
            };
            load_b(0);
            expand(0);
#pragma unroll
            for (int cb = 0; cb < NCB; ++cb) {
                if (cb + 1 < NCB) load_b(cb + 1);
                uint32_t e0[4], e1[4];  // max over the slots (geotransformer.py:66) = the score contraction's A operand
#pragma unroll
                for (int t = 0; t < 4; ++t)
#pragma unroll
                    for (int r = 0; r < 2; ++r) {
                        uint32_t v = c[0][t][r];
#pragma unroll
                        for (int sl = 1; sl < SL; ++sl) v = hmax2_u(v, c[sl][t][r]);
                        (t < 2 ? e0 : e1)[(t & 1) * 2 + r] = v;
                    }
                if (cb + 1 < NCB) expand(cb + 1);
                mma_acc(sc, e0, make_uint2(qc[cb & 1].x, qc[cb & 1].y));
                mma_acc(sc1, e1, make_uint2(qc[cb & 1].z, qc[cb & 1].w));
            }
            // fragment columns 0-3 hold the hi halves' share (and the distance term), columns 4-7 the lo halves': lanes
            // q and q ^ 2 add up; lanes q < 2 then hold heads 2q, 2q + 1 of keys 16 w + g and + 8
            float v[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                v[r] = sc[r] + sc1[r];
                v[r] += __shfl_xor_sync(0xffffffffu, v[r], 2);
            }
            const long long b = cwalk.b;
            const int i = p.row_begin + (int)cwalk.i;
            const int j = (int)cwalk.kb * CHEB_TILE + warp * 16 + g;
            if (q < 2) {
                float* out = reinterpret_cast<float*>(p.out) + (((size_t)b * p.H + 2 * q) * p.n + i) * p.n + j;
                const size_t hs = (size_t)p.n * p.n;
                if (2 * q < p.H) {
                    if (j < p.n) out[0] = v[0];
                    if (j + 8 < p.n) out[8] = v[2];
                }
                if (2 * q + 1 < p.H) {
                    if (j < p.n) out[hs] = v[1];
                    if (j + 8 < p.n) out[hs + 8] = v[3];
                }
            }
        };

        int it = 0;
        for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int buf = it % NB;
            if (SPLIT && it > 0) cwalk.next();
            // the producers have filled A[buf] (and the query image)
#ifdef CHEB_NAMED_BARRIERS
            cheb_bar_sync(1 + buf);
#else
            ptx::mbar_wait(bar_full + 8 * buf, (it / NB) & 1, 0x200);
#endif
#ifndef CHEB_ABL_NO_CONSUME
            if constexpr (SPLIT) consume_rows(tile, buf);
            else consume(tile, buf);
#endif
#ifdef CHEB_NAMED_BARRIERS
            __threadfence_block();
            cheb_bar_arrive(3 + buf);
#else
            __syncwarp();  // every lane's reads of A[buf] and partial-score stores are done ...
            if (lane == 0) ptx::mbar_arrive(bar_empty + 8 * buf);  // ... before the warp's one arrival (release)
#endif
        }
    }
}

template <int C, int KA, bool SCORES>
__global__ void __launch_bounds__(CHEB_THREADS, CHEB_MIN_CTAS) gse_embed_cheb_kernel(const ChebArgs p) {
    extern __shared__ __align__(16) unsigned char cheb_smem[];
    // the coefficient set of the distance term: the shortest expansion that covers the clouds' largest distance
    const float x_max = __fdiv_rn(__fsqrt_rn(__uint_as_float(__ldg(p.dmax2_bits))), p.sigma_d);
    int d_set = 0;
    while (d_set < p.max_d_set && x_max > cheb_range(d_set)) ++d_set;
    // fast instances: 3 angle slots (k = 3, 'max'), distance sets 1 (also serving set 0: its 32 coefficients cover the
    // shorter range too) and 2; everything else runs the generic instance
    const bool half_out = SCORES || !p.out_bf16;  // the fast instances accumulate in half
    // split instances (the distance term in coefficient space, a consumer warp per m-tile): distance sets <= 3; the host
    // sized the shared memory for them under the same condition (cheb_launch)
    const bool split = SCORES && p.u != nullptr && p.H <= 4 && p.slots == 3 && p.max_d_set >= 1;
    if (split && d_set <= 1) cheb_run<C, KA, 2, SCORES, 3, SCORES>(p, 1, cheb_smem);
    else if (split && d_set == 2) cheb_run<C, KA, 3, SCORES, 3, SCORES>(p, 2, cheb_smem);
    else if (split && d_set == 3) cheb_run<C, KA, 4, SCORES, 3, SCORES>(p, 3, cheb_smem);
    else if (half_out && p.slots == 3 && d_set <= 1 && p.max_d_set >= 1) cheb_run<C, KA, 2, SCORES, 3>(p, 1, cheb_smem);
    else if (half_out && p.slots == 3 && d_set == 2) cheb_run<C, KA, 3, SCORES, 3>(p, 2, cheb_smem);
    else cheb_run<C, KA, 0, SCORES, 0>(p, d_set, cheb_smem);
}

}  // namespace geotx

using namespace geotx;

static bool cheb_shape_ok(int C, int ka_tiles) { return (C == 256 || C == 128) && (ka_tiles == 1 || ka_tiles == 2); }

extern "C" int geotx_gse_cheb_ka_tiles(float angle_index_max) {
    if (!(angle_index_max > 0.f)) return 0;
    if (angle_index_max <= cheb_range(0)) return 1;
    if (angle_index_max <= cheb_range(1)) return 2;
    return 0;
}

extern "C" float geotx_gse_cheb_d_range(int d_set) {
    return d_set >= 0 && d_set < CHEB_NUM_D_SETS ? cheb_range(d_set) : 0.f;
}

extern "C" int64_t geotx_gse_cheb_bytes(int C, int ka_tiles) {
    if (!cheb_shape_ok(C, ka_tiles)) return 0;
    return (int64_t)cheb_set_offset(C, ka_tiles, 1 + CHEB_NUM_D_SETS) * 2;
}

extern "C" int64_t geotx_gse_cheb_plain_rows(int ka_tiles) {
    if (ka_tiles != 1 && ka_tiles != 2) return 0;
    int64_t rows = 16 * ka_tiles;
    for (int s = 0; s < CHEB_NUM_D_SETS; ++s) rows += 16 * cheb_kd_tiles(s);
    return rows;
}

extern "C" int geotx_gse_build_cheb(const float* div_term, const float* w_d, const float* b_d, const float* w_a,
                                    const float* b_a, int C, float angle_index_max, int ka_tiles, void* coef,
                                    float* stats, float* plain, void* stream) {
    GEOTX_REQUIRE(div_term && w_d && b_d && w_a && b_a && coef && stats, "geotx_gse_build_cheb: null pointer");
    GEOTX_REQUIRE(cheb_shape_ok(C, ka_tiles), "geotx_gse_build_cheb: C must be 128 or 256, ka_tiles 1 or 2");
    GEOTX_REQUIRE(angle_index_max > 0.f && angle_index_max <= cheb_range(ka_tiles - 1),
                  "geotx_gse_build_cheb: angle range %g is not covered by %d coefficients", angle_index_max, 16 * ka_tiles);
    cudaStream_t st = (cudaStream_t)stream;
    int rc = check_cuda(cudaMemsetAsync(stats, 0, sizeof(float) * CHEB_STATS_PER_SET * (1 + CHEB_NUM_D_SETS), st),
                        "geotx_gse_build_cheb: memset");
    if (rc) return rc;
    __half* base = (__half*)coef;
    // the angle term carries no bias: max over the slots commutes with adding b_a, which rides with b_d in G_d[0]
    cheb_build_kernel<<<C, CHEB_BUILD_NODES, 0, st>>>(div_term, w_a, nullptr, nullptr, C, angle_index_max, 16 * ka_tiles,
                                                     base, (unsigned int*)stats, plain);
    size_t plain_row = 16 * ka_tiles;
    for (int s = 0; s < CHEB_NUM_D_SETS; ++s) {
        cheb_build_kernel<<<C, CHEB_BUILD_NODES, 0, st>>>(div_term, w_d, b_d, b_a, C, cheb_range(s), 16 * cheb_kd_tiles(s),
                                                         base + cheb_set_offset(C, ka_tiles, 1 + s),
                                                         (unsigned int*)stats + CHEB_STATS_PER_SET * (1 + s),
                                                         plain ? plain + plain_row * C : nullptr);
        plain_row += 16 * cheb_kd_tiles(s);
    }
    return launch_status("geotx_gse_build_cheb");
}

// shared launcher: E mode (scores == nullptr) or scores mode
static int cheb_launch(const char* what, const void* geometry, int B, int n, int k, int row_begin, int row_count,
                       const void* coef, int C, float angle_index_max, int ka_tiles, int max_d_set, float sigma_d,
                       float factor_a, int reduce_mean, int out_dtype, void* out, const float* qt, int64_t qsb,
                       int64_t qsh, int64_t qsn, int H, const float* u, int64_t usb, int64_t ush, int64_t usn,
                       int32_t* overflow, void* stream) {
    const bool scores = qt != nullptr;
    ChebArgs p;
    p.pts4 = (const float4*)geometry;
    p.refv = p.pts4 + (size_t)B * n;
    p.dmax2_bits = (const unsigned int*)(p.refv + (size_t)B * n * k);
    p.coef = (const __half*)coef;
    p.out = out;
    p.overflow = overflow;
    p.pairs_total = (long long)B * row_count * n;
    p.n = n;
    p.k = k;
    p.slots = reduce_mean ? 1 : k;
    p.reduce_mean = reduce_mean;
    p.row_begin = row_begin;
    p.row_count = row_count;
    p.out_bf16 = out_dtype == GEOTX_BF16;
    p.max_d_set = max_d_set;
    p.sigma_d = sigma_d;
    p.inv_sigma_d = 1.f / sigma_d;
    p.factor_a = factor_a;
    p.inv_half_la = 2.f / angle_index_max;
    for (int s = 0; s < CHEB_NUM_D_SETS; ++s) p.d_set_offset[s] = (unsigned int)cheb_set_offset(C, ka_tiles, 1 + s);
    p.qt = qt;
    p.qsb = qsb;
    p.qsh = qsh;
    p.qsn = qsn;
    p.H = H;
    p.u = u;
    p.usb = usb;
    p.ush = ush;
    p.usn = usn;
    for (int s = 0; s < CHEB_NUM_D_SETS; ++s)  // u holds sets 1, 2 (, 3) back to back; set 0 uses set 1's coefficients
        p.u_offset[s] = s == 2 ? 16 * cheb_kd_tiles(1) : s == 3 ? 16 * (cheb_kd_tiles(1) + cheb_kd_tiles(2)) : 0;
    p.key_blocks = (n + CHEB_TILE - 1) / CHEB_TILE;
    p.tiles = scores ? (long long)B * row_count * p.key_blocks : (p.pairs_total + CHEB_TILE - 1) / CHEB_TILE;
    GEOTX_REQUIRE(p.tiles < 2147483647LL, "%s: too many tiles", what);
    const int ktot_max = p.slots * 16 * ka_tiles + 16 * cheb_kd_tiles(max_d_set);
    size_t smem = CHEB_BAR_BYTES + (size_t)2 * CHEB_TILE * (ktot_max * 2 + 16);
    if (scores) smem += (size_t)2 * (C / 32) * (H <= 4 ? 4 : 8) * CHEB_PSTRIDE * sizeof(float) + (size_t)2 * (C / 32) * 32 * 32;
    if (scores && u && H <= 4 && p.slots == 3 && max_d_set >= 1) {
        // the split instances' own layout (no partial scores): A tiles for the widest set they handle, the angle
        // coefficients, the packed query image, u.  Sets 4 and 5 fall back to the generic instance and its layout.
        const int split_set = max_d_set < 3 ? max_d_set : 3;
        size_t split = 0;  // the largest layout among the instances this call may dispatch to (sets 1 .. split_set)
        for (int s = 1; s <= split_set; ++s) {
            const size_t kd = cheb_kd_tiles(s), nb = cheb_split_buffers((int)kd);
            const size_t need = CHEB_BAR_BYTES + nb * CHEB_TILE * ((3 * 16 * ka_tiles + 16 * kd) * 2 + 16) +
                                (size_t)ka_tiles * (C / 32) * 2 * 32 * 16 + nb * (C / 32) * 32 * 16 + nb * kd * 32 * 16;
            if (need > split) split = need;
        }
        smem = max_d_set <= 3 ? split : (split > smem ? split : smem);
    }
    GEOTX_REQUIRE(smem <= 220 * 1024, "%s: %zu B of shared memory (k=%d, max_d_set=%d) do not fit", what, smem, k, max_d_set);
    int dev = 0, sms = 0;
    int rc = check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
    if (rc) return rc;
    rc = check_cuda(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev), "cudaDeviceGetAttribute");
    if (rc) return rc;
    auto launch = [&](auto kernel) -> int {
        int r = check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), what);
        if (r) return r;
        int per_sm = 0;  // persistent grid: exactly the CTAs that are resident at once
        r = check_cuda(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, CHEB_THREADS, smem), what);
        if (r) return r;
        const int grid = (int)std::min<long long>(p.tiles, (long long)sms * std::max(per_sm, 1));
        kernel<<<grid, CHEB_THREADS, smem, (cudaStream_t)stream>>>(p);
        return 0;
    };
    if (scores) {
        if (C == 256) rc = ka_tiles == 1 ? launch(gse_embed_cheb_kernel<256, 1, true>) : launch(gse_embed_cheb_kernel<256, 2, true>);
        else rc = ka_tiles == 1 ? launch(gse_embed_cheb_kernel<128, 1, true>) : launch(gse_embed_cheb_kernel<128, 2, true>);
    } else {
        if (C == 256) rc = ka_tiles == 1 ? launch(gse_embed_cheb_kernel<256, 1, false>) : launch(gse_embed_cheb_kernel<256, 2, false>);
        else rc = ka_tiles == 1 ? launch(gse_embed_cheb_kernel<128, 1, false>) : launch(gse_embed_cheb_kernel<128, 2, false>);
    }
    if (rc) return rc;
    return launch_status(what);
}

extern "C" int geotx_gse_embed_cheb(const void* geometry, int B, int n, int k, int row_begin, int row_count,
                                    const void* coef, int C, float angle_index_max, int ka_tiles, int max_d_set,
                                    float sigma_d, float factor_a, int reduce_mean, int out_dtype, void* out,
                                    int32_t* overflow, void* stream) {
    GEOTX_REQUIRE(geometry && coef && out && overflow, "geotx_gse_embed_cheb: null pointer");
    GEOTX_REQUIRE(B > 0 && n > 0 && k >= 1 && k <= GEOTX_MAX_ANGLE_K, "geotx_gse_embed_cheb: bad sizes (B=%d n=%d k=%d)", B, n, k);
    GEOTX_REQUIRE(cheb_shape_ok(C, ka_tiles), "geotx_gse_embed_cheb: C must be 128 or 256, ka_tiles 1 or 2");
    GEOTX_REQUIRE(row_begin >= 0 && row_count > 0 && row_begin + row_count <= n, "geotx_gse_embed_cheb: bad row slab");
    GEOTX_REQUIRE(out_dtype == GEOTX_F16 || out_dtype == GEOTX_BF16, "geotx_gse_embed_cheb: out_dtype must be f16 or bf16");
    GEOTX_REQUIRE(max_d_set >= 0 && max_d_set < CHEB_NUM_D_SETS, "geotx_gse_embed_cheb: max_d_set outside [0,%d)", CHEB_NUM_D_SETS);
    GEOTX_REQUIRE(angle_index_max > 0.f && sigma_d > 0.f, "geotx_gse_embed_cheb: bad scales");
    return cheb_launch("geotx_gse_embed_cheb", geometry, B, n, k, row_begin, row_count, coef, C, angle_index_max, ka_tiles,
                       max_d_set, sigma_d, factor_a, reduce_mean, out_dtype, out, nullptr, 0, 0, 0, 0, nullptr, 0, 0, 0, overflow,
                       stream);
}

extern "C" int geotx_rpe_scores_cheb(const void* geometry, int B, int n, int k, int row_begin, int row_count,
                                     const void* coef, int C, float angle_index_max, int ka_tiles, int max_d_set,
                                     float sigma_d, float factor_a, int reduce_mean, const float* qt,
                                     int64_t qt_stride_b, int64_t qt_stride_h, int64_t qt_stride_n, int H,
                                     const float* u, int64_t u_stride_b, int64_t u_stride_h, int64_t u_stride_n,
                                     float* scores_p, int32_t* overflow, void* stream) {
    GEOTX_REQUIRE(geometry && coef && qt && scores_p && overflow, "geotx_rpe_scores_cheb: null pointer");
    GEOTX_REQUIRE(B > 0 && n > 0 && k >= 1 && k <= GEOTX_MAX_ANGLE_K, "geotx_rpe_scores_cheb: bad sizes (B=%d n=%d k=%d)", B, n, k);
    GEOTX_REQUIRE(cheb_shape_ok(C, ka_tiles), "geotx_rpe_scores_cheb: C must be 128 or 256, ka_tiles 1 or 2");
    GEOTX_REQUIRE(row_begin >= 0 && row_count > 0 && row_begin + row_count <= n, "geotx_rpe_scores_cheb: bad row slab");
    GEOTX_REQUIRE(max_d_set >= 0 && max_d_set < CHEB_NUM_D_SETS, "geotx_rpe_scores_cheb: max_d_set outside [0,%d)", CHEB_NUM_D_SETS);
    GEOTX_REQUIRE(angle_index_max > 0.f && sigma_d > 0.f, "geotx_rpe_scores_cheb: bad scales");
    GEOTX_REQUIRE(H >= 1 && H <= 8, "geotx_rpe_scores_cheb: H=%d outside [1,8]", H);
    GEOTX_REQUIRE(((uintptr_t)qt & 15) == 0 && qt_stride_b % 4 == 0 && qt_stride_h % 4 == 0 && qt_stride_n % 4 == 0,
                  "geotx_rpe_scores_cheb: qt must be 16-byte aligned with strides that are multiples of 4 elements");
    GEOTX_REQUIRE(!u || (((uintptr_t)u & 7) == 0 && u_stride_b % 2 == 0 && u_stride_h % 2 == 0 && u_stride_n % 2 == 0),
                  "geotx_rpe_scores_cheb: u must be 8-byte aligned with even strides");
    return cheb_launch("geotx_rpe_scores_cheb", geometry, B, n, k, row_begin, row_count, coef, C, angle_index_max, ka_tiles,
                       max_d_set, sigma_d, factor_a, reduce_mean, GEOTX_F16, scores_p, qt, qt_stride_b, qt_stride_h,
                       qt_stride_n, H, u, u_stride_b, u_stride_h, u_stride_n, overflow, stream);
}
