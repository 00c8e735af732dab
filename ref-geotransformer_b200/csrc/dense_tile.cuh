// Shared by K4f (k4_post_attention.cu) and K13 (k13_linear.cu): a 32-token tile's GEMMs against nn.Linear weights that
// stream through a ring of TMA bulk copies.
//
// Packed weight format (geotx_pack_linear_weight): W [N][K] fp32 -> halves, K chunks of 32, chunk-major:
//     packed[(c * N + r) * 32 + ((u ^ ((r >> 1) & 3)) * 8) + e] = half(W[r][32 c + 8 u + e]),  u = 0..3, e = 0..7
// i.e. chunk c holds 64 dense bytes per output row r, rows in order, so any row range [n0, n0 + NB) of a chunk is one
// contiguous block of NB * 64 bytes (one bulk copy), and the XOR on the 16-byte unit index makes the eight rows an
// ldmatrix phase touches fall into eight different bank groups without padding bytes.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>
#include "sm100_ptx.cuh"

namespace geotx {

constexpr int DT_BM = 32;          // tokens per CTA
constexpr int DT_KC = 32;          // K per chunk (halves)
constexpr int DT_ROW_BYTES = DT_KC * 2;
constexpr int DT_WARPS = 8;        // compute warps: each owns 1/8 of the CTA's output columns
constexpr int DT_THREADS = (DT_WARPS + 1) * 32;  // + the loader warp
#ifndef DT_STAGES_N
#define DT_STAGES_N 4  // ring depth (timing experiments: -DDT_STAGES_N=2|3)
#endif
constexpr int DT_STAGES = DT_STAGES_N;

__device__ __forceinline__ void dt_ldsm4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void dt_ldsm2(uint32_t (&r)[2], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
__device__ __forceinline__ void dt_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t dt_pack(float lo, float hi) {  // saturating: a stray large activation stays finite
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
// compute warps only (the loader warp never joins): named barrier 1
__device__ __forceinline__ void dt_sync_compute() { asm volatile("bar.sync 1, %0;" ::"n"(DT_WARPS * 32) : "memory"); }

// Ring bookkeeping shared by the loader and the compute warps: chunk number g of the CTA's flat schedule lives in stage
// g % DT_STAGES; its `full` barrier completes phase g / DT_STAGES, so does its `empty` barrier once all warps are done.
struct DtRing {
    uint32_t bar_full, bar_empty, stages, stage_bytes;
    __device__ __forceinline__ void init(int tid) const {
        if (tid == 0) {
            for (int s = 0; s < DT_STAGES; ++s) {
                ptx::mbar_init(bar_full + 8 * s, 1);
                ptx::mbar_init(bar_empty + 8 * s, DT_WARPS);
            }
            ptx::fence_barrier_init();
        }
    }
    // loader (one thread): chunk g <- `bytes` contiguous bytes at src
    __device__ __forceinline__ void load(int g, const void* src, uint32_t bytes) const {
        const int s = g % DT_STAGES;
        if (g >= DT_STAGES) ptx::mbar_wait(bar_empty + 8 * s, ((g / DT_STAGES) - 1) & 1, 0x400u | (uint32_t)g);
        ptx::mbar_arrive_expect_tx(bar_full + 8 * s, bytes);
        ptx::bulk_g2s(stages + s * stage_bytes, src, bytes, bar_full + 8 * s);
    }
    __device__ __forceinline__ uint32_t wait(int g) const {  // compute warps: -> shared address of chunk g
        const int s = g % DT_STAGES;
        ptx::mbar_wait(bar_full + 8 * s, (g / DT_STAGES) & 1, 0x500u | (uint32_t)g);
        return stages + s * stage_bytes;
    }
    __device__ __forceinline__ void release(int g, int lane) const {
        // The stage was read through the generic proxy (ldmatrix) and will be overwritten through the async proxy (the
        // loader's next bulk copy): the mbarrier hand-over orders generic-proxy accesses only, so every reader fences
        // its reads against the async proxy first.  Without this fence CTAs of the second wave, starting next to a
        // running CTA on the same SM, sporadically contracted against partly overwritten weight rows
        // (tools/dense_stress.py: 14 of 300 calls at N = 1792, every call at N = 2112; none with the fence).
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(bar_empty + 8 * (g % DT_STAGES));
    }
};

// One GEMM of a token tile: acc[mt][nt] (2 m-tiles x NT n-tiles: the 8 NT columns of this warp that start at row
// `n_warp` of the staged chunk) += A (shared memory, half, row pitch a_pitch bytes, K from column 0) x W^T, over
// `entries` ring entries whose flat numbers start at g0.  A ring entry holds SUB consecutive K chunks of the packed
// weight (they are contiguous in memory: one bulk copy), `sub_bytes` apart: one wait / fence / release per entry.
template <int NT, int SUB = 1>
__device__ __forceinline__ void dt_gemm(float (&acc)[2][NT][4], const DtRing& ring, int g0, int entries, uint32_t a_addr,
                                        int a_pitch, int n_warp, int lane, uint32_t sub_bytes = 0) {
    const int brow = n_warp + (lane & 7) + ((lane >> 4) & 1) * 8;  // + 16 np: this lane's row of an n-tile pair
    const int bunit = (lane >> 3) & 1;                             // + 2 ks: its 16-byte unit
    for (int c0 = 0; c0 < entries; ++c0) {
      const uint32_t w_entry = ring.wait(g0 + c0);
#pragma unroll
      for (int sub = 0; sub < SUB; ++sub) {
        const int c = c0 * SUB + sub;  // K chunk
        const uint32_t w_addr = w_entry + sub * sub_bytes;
#pragma unroll
        for (int ks = 0; ks < DT_KC / 16; ++ks) {
            uint32_t a[2][4];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
                dt_ldsm4(a[mt], a_addr + (mt * 16 + (lane & 15)) * a_pitch + (c * DT_KC + ks * 16 + (lane >> 4) * 8) * 2);
#pragma unroll
            for (int np = 0; np < NT / 2; ++np) {  // two n-tiles per ldmatrix.x4
                const int r = brow + np * 16;
                uint32_t b[4];
                dt_ldsm4(b, w_addr + r * DT_ROW_BYTES + (((ks * 2 + bunit) ^ ((r >> 1) & 3)) << 4));
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) {
                    dt_mma(acc[mt][2 * np], a[mt], b[0], b[1]);
                    dt_mma(acc[mt][2 * np + 1], a[mt], b[2], b[3]);
                }
            }
            if (NT & 1) {  // a last single n-tile: lanes 0-15 address its rows (unit = lane / 8)
                const int r = n_warp + (NT - 1) * 8 + (lane & 7);
                uint32_t b[2];
                dt_ldsm2(b, w_addr + r * DT_ROW_BYTES + (((ks * 2 + ((lane >> 3) & 1)) ^ ((r >> 1) & 3)) << 4));
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) dt_mma(acc[mt][NT - 1], a[mt], b[0], b[1]);
            }
        }
      }
      ring.release(g0 + c0, lane);
    }
}

}  // namespace geotx
