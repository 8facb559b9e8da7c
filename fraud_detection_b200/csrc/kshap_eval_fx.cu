// K2f - factored Kernel SHAP evaluator for a SHARED coalition set (sm_100a, tcgen05 + TMEM + TMA).
//
// With one coalition set for every explained row, the background part of the logit does not depend on the
// row:  logit[r][m][b] = G[m][b] + c + s[r][m],  G = ((1-Z) diag(w)) bg^T,  s[r][m] = sum_k z_mk w_k x_rk.
// So  2^(-log2e logit) = E[m][b] * u[r][m]  with  E = 2^(-log2e (G + c) - o_m)  built ONCE per plan and
// u = 2^(-log2e s + o_m)  one exponential per (row, coalition) instead of one per (row, coalition, background):
//     ey[r][m] = mean_b 1 / (1 + E[m][b] u[r][m])
// (o_m centres the exponents of coalition m so E and u stay inside fp32 range.)
//
// TWO small GEMMs per work unit (128 explained rows x 128 coalitions) run on the tensor cores:
//  (1) the per-row exponents  D[r][m] = Q[r][:] . Abar[m][:]  with
//        Q    (128 rows x 32, smem, A operand) = 3-term f16/bf16 split of [log2e w_k x_rk | -log2e sum_k w_k x_rk | 0]
//        Abar (128 coalitions x 32, smem, B operand) = [1 - z_mk | 1 | 0]   (the same tile image K2 uses as A)
//        D    (TMEM, fp32, lane = explained row, column = coalition) = -log2e s[r][m]
//  (2) the first half of the least-squares solve,  G[r][j] = sum_m v[r][m] x[m][j]  (X'W y of K3), with
//        v    (128 rows x 128 coalitions) = kw_m (link(ey[r][m]) - link(fnull)), written by the compute threads
//             STRAIGHT INTO TENSOR MEMORY as a 3-term bf16 split (exact to 24 bits) - the A operand is read from TMEM
//        x^T  (32 x 128, smem, B operand) = z_mj - z_m29 in {-1, 0, 1}: exact in bf16
//        G    (TMEM, fp32, 32 columns), read back once per unit and summed over the tiles of a row block in registers
//             (24 MMA steps per accumulator, so the tensor core's truncating accumulation stays below fp32 rounding).
//     Only 128 B per explained row (and per CTA that touched its row block: at most 2 for any real batch) go to HBM.
// Each compute thread owns one explained row (TMEM lane) and 32 coalitions; E[m][b] is the same for the
// whole warp, so it is read from shared memory as a broadcast LDS.128 and the inner loop is
//   d = fma(E, a, b);  four d share one reciprocal       (3.5 issue slots, 0.25 SFU ops per sigmoid)
// against 6.25 issue slots and 1.25 SFU ops in the general evaluator.  E streams through a double-buffered
// TMA ring in background chunks of <= 128 columns.
//
// CTA = 20 warps, persistent, one per SM.  The units are numbered row-block-major (u = row_block * tiles + tile)
// and every CTA takes one CONTIGUOUS range of them (equal to +-1 unit), so Q is built once per row block, the X'W y
// sums stay in registers across the tiles of a row block, and no wave is left half empty:
//   warp 0 (one lane)  TMA: per unit the coalition tile image + x^T image + (o_m, kw_m), and the E chunks
//   warp 1 (one lane)  tcgen05.mma issue: 6 per unit for D, 24 per unit for G; commits
//   warps 2-3          build the Q block of a row block (x rows -> scaled, split, canonical K-major layout)
//   warps 4-19         compute: TMEM lane quarter = warp % 4, coalition group = (warp - 4) / 4

#include <math.h>

#include <algorithm>
#include <new>
#include <utility>
#include <vector>

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "kshap_internal.cuh"
#include "ptx_sm100.cuh"

namespace {

constexpr int kFxThreads = 640;
constexpr int kFxComputeWarps = 16;
constexpr int kTileBytes = 128 * 32 * 2;         // 8192: one 128 x 32 operand image (also the 32 x 128 x^T image)
constexpr int kXStageBytes = FDB_FX_TILE_BYTES;  // x^T image + o_m[128] + kw_m[128]
constexpr int kQBytes = 3 * kTileBytes;          // three split terms
constexpr int kEStageBytes = 128 * 128 * 4;      // 65536: 128 coalitions x <= 128 background columns
constexpr uint32_t kLbo = 128, kSbo = 512;
constexpr uint32_t kSboX = 2048;                 // x^T image: 8-row groups of 16 K-chunks
constexpr uint32_t kTmemA2 = 256;                // TMEM columns: D stages 0..255, v splits 256..447 (3 x 64), G 448..479
constexpr uint32_t kTmemG = 448;

struct FxParams {
    const float* x;
    const uint16_t* zimg;      // coalition tile images (plan->d_aimg, set 0)
    const unsigned char* xt;   // per tile: x^T image (8192 B) + o_m[128] + kw_m[128]   (plan->d_fxtile)
    const float* E;            // [tile][chunk][128][cbs]
    const float* wq;
    float* ey;                 // optional [n_rows][m_pad]
    float* gpart;              // [n_rows][slots][32] X'W y per (row, CTA that touched the row block)
    int n_rows, tiles, m_pad, n_bg, fmt;
    int n_chunks;              // background chunks of 128 columns (last one narrower)
    int slots;
    long long n_units;
    float inv_bg, lnull;
};

struct FxBars {
    uint64_t z_full[2], z_empty[2], q_full[2], q_empty[2], t_full[2], t_empty[2], e_full[2], e_empty[2];
    uint64_t a2_full, g_full;
    uint32_t tmem_base;
};

__device__ __forceinline__ uint16_t fx_to_h16(float v, int fmt) {
    return fmt == 0 ? __half_as_ushort(__float2half_rn(v)) : __bfloat16_as_ushort(__float2bfloat16_rn(v));
}
__device__ __forceinline__ float fx_from_h16(uint16_t h, int fmt) {
    return fmt == 0 ? __half2float(__ushort_as_half(h)) : __bfloat162float(__ushort_as_bfloat16(h));
}

// Sums over one background chunk for one row and FOUR coalitions at once (independent chains hide the LDS and
// SFU latency and share the loop overhead).  e0 points at E[m0][chunk columns] in smem, rows are cbs floats
// apart; every lane reads the same addresses -> broadcast LDS.128.
// MODE 0 / 1 (clamped): d = min(E u + 1, clamp), u = 2^t;  0 = four sigmoids per reciprocal, clamp 2^30 (a sigmoid
//   moves by < 1e-9);  1 = logit link: two per reciprocal and clamp 2^62 so tiny sigmoids keep their relative accuracy.
// MODE 2 (balanced, no clamp): sigma = 1 / (1 + E u) = b / (E a + b) with (a, b) = (u, 1) for u <= 1 and (1, 1/u)
//   otherwise, so E <= d = E a + b <= E + 1 whatever the row: the product of the four d of a column quad is bounded by
//   the E values alone.  The columns are ordered so that a quad mixes the four quarters of the logit-sorted
//   background (fdb_kshap_build_factored) and kshap_build_E measures the bound; when every quad stays within
//   2^+-114 there is no min() in the loop, tiny sigmoids keep their relative accuracy (both links) and a sigmoid
//   costs 3.5 issue slots.  acc gathers sum 1/d; the caller multiplies by b once per (row, coalition).
// MODE 3 (balanced, logit link): as MODE 2 plus the complements  1 - sigma = g / d,  g = E a  (d = g + b), summed with
//   the same reciprocal:  sum_quad g_i / d_i = ((g0 d1 + g1 d0) p23 + (g2 d3 + g3 d2) p01) / (p01 p23).  accc gathers
//   them; with both sums the stored expectation is exact to float32 in both tails (fdb_kshap_encode_ey).
#ifndef FDB_FX_QUAD_UNROLL
#define FDB_FX_QUAD_UNROLL 2      // quads of one coalition in flight (x 4 interleaved coalitions), multi-chunk backgrounds
#endif
#ifndef FDB_FX_QUAD_UNROLL_SINGLE
#define FDB_FX_QUAD_UNROLL_SINGLE 3   // ... single-chunk backgrounds: no per-coalition partial sums in registers, room for three
#endif                                //     (measured at config 3, ms per 100k rows: 1 -> 2.97, 2 -> 2.91, 3 -> 2.77, 4 -> 2.83)
#ifndef FDB_FX_PREFETCH_MULTI
#define FDB_FX_PREFETCH_MULTI 0   // multi-chunk backgrounds: fetch the next group's exponents from TMEM ahead of the sums (costs 4 registers)
#endif
#define FDB_FX_PRAGMA(x) _Pragma(#x)
#define FDB_FX_UNROLL(n) FDB_FX_PRAGMA(unroll n)
enum { kFxClampedQuads = 0, kFxClampedPairs = 1, kFxBalanced = 2, kFxBalancedLogit = 3 };

template <int MODE>
__device__ __forceinline__ void fx_scales(const float (&t)[4], float (&a)[4], float (&b)[4]) {
    constexpr bool kBal = MODE == kFxBalanced || MODE == kFxBalancedLogit;
#pragma unroll
    for (int mm = 0; mm < 4; ++mm) {
        a[mm] = ptx::ex2_approx(kBal ? fminf(t[mm], 0.f) : t[mm]);
        b[mm] = kBal ? ptx::ex2_approx(fminf(-t[mm], 0.f)) : 1.0f;
    }
}

// sums of this chunk are gathered apart and added to the running totals once: with ~1000 background rows a
// single float32 chain drifts by 1e-5 on saturated coalitions (every term 1, every add rounded the same way)
template <int MODE, int UNROLL>
__device__ __forceinline__ void chunk_sum4(const float* __restrict__ e0, int cbs, int nquads, int ntail,
                                           const float (&a)[4], const float (&b)[4], float (&sum)[4]) {
    constexpr float kClamp = MODE == kFxClampedPairs ? 4.611686018427388e18f : 1073741824.f;
#pragma unroll
    for (int mm = 0; mm < 4; ++mm) sum[mm] = 0.f;
#pragma unroll UNROLL
    for (int q = 0; q < nquads; ++q) {
#pragma unroll
        for (int mm = 0; mm < 4; ++mm) {
            const float4 e = *reinterpret_cast<const float4*>(e0 + mm * cbs + 4 * q);
            float d0 = fmaf(e.x, a[mm], b[mm]), d1 = fmaf(e.y, a[mm], b[mm]);
            float d2 = fmaf(e.z, a[mm], b[mm]), d3 = fmaf(e.w, a[mm], b[mm]);
            if (MODE != kFxBalanced) {
                d0 = fminf(d0, kClamp); d1 = fminf(d1, kClamp); d2 = fminf(d2, kClamp); d3 = fminf(d3, kClamp);
            }
            const float n01 = d0 + d1, p01 = d0 * d1, n23 = d2 + d3, p23 = d2 * d3;
            if (MODE == kFxClampedPairs) {
                sum[mm] = fmaf(n01, ptx::rcp_approx(p01), sum[mm]);
                sum[mm] = fmaf(n23, ptx::rcp_approx(p23), sum[mm]);
            } else {
                sum[mm] = fmaf(fmaf(n01, p23, n23 * p01), ptx::rcp_approx(p01 * p23), sum[mm]);
            }
        }
    }
    for (int i = 0; i < ntail; ++i)
#pragma unroll
        for (int mm = 0; mm < 4; ++mm) {
            const float d = fmaf(e0[mm * cbs + 4 * nquads + i], a[mm], b[mm]);
            sum[mm] += ptx::rcp_approx(MODE == kFxBalanced ? d : fminf(d, kClamp));
        }
}

// MODE 3 body: sigmoid and complement sums for four coalitions over one background chunk (see the mode list above)
__device__ __forceinline__ void chunk_sum4_logit(const float* __restrict__ e0, int cbs, int nquads, int ntail,
                                                 const float (&a)[4], const float (&b)[4], float (&sum)[4], float (&sumc)[4]) {
#pragma unroll
    for (int mm = 0; mm < 4; ++mm) { sum[mm] = 0.f; sumc[mm] = 0.f; }
#pragma unroll 2
    for (int q = 0; q < nquads; ++q) {
#pragma unroll
        for (int mm = 0; mm < 4; ++mm) {
            const float4 e = *reinterpret_cast<const float4*>(e0 + mm * cbs + 4 * q);
            const float g0 = e.x * a[mm], g1 = e.y * a[mm], g2 = e.z * a[mm], g3 = e.w * a[mm];
            const float d0 = g0 + b[mm], d1 = g1 + b[mm], d2 = g2 + b[mm], d3 = g3 + b[mm];
            const float n01 = d0 + d1, p01 = d0 * d1, n23 = d2 + d3, p23 = d2 * d3;
            const float r = ptx::rcp_approx(p01 * p23);
            sum[mm] = fmaf(fmaf(n01, p23, n23 * p01), r, sum[mm]);
            const float c01 = fmaf(g0, d1, g1 * d0), c23 = fmaf(g2, d3, g3 * d2);
            sumc[mm] = fmaf(fmaf(c01, p23, c23 * p01), r, sumc[mm]);
        }
    }
    for (int i = 0; i < ntail; ++i)
#pragma unroll
        for (int mm = 0; mm < 4; ++mm) {
            const float g = e0[mm * cbs + 4 * nquads + i] * a[mm];
            const float r = ptx::rcp_approx(g + b[mm]);
            sum[mm] += r;
            sumc[mm] = fmaf(g, r, sumc[mm]);
        }
}

#ifndef FDB_FX_SERVICE_SLEEP_NS
#define FDB_FX_SERVICE_SLEEP_NS 800     // a unit lasts ~30 us and every hand-off is double-buffered
#endif
// waits of the single-lane service warps back off so their polling does not take issue slots from the compute warps
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
    while (!ptx::mbar_try_wait(bar, parity)) __nanosleep(FDB_FX_SERVICE_SLEEP_NS);
}

__device__ __forceinline__ int fx_cbw(int n_bg, int c) { const int r = n_bg - 128 * c; return r < 128 ? r : 128; }

// the CTA that owns unit u when n_units units are dealt out as contiguous ranges [c n_units / grid, (c + 1) n_units / grid)
__host__ __device__ __forceinline__ long long fx_owner(long long u, long long n_units, long long grid) {
    return ((u + 1) * grid - 1) / n_units;
}

// Four finished coalitions of one row: expectation -> stored form (ey output), link, v = kw (link(ey) - link(fnull)),
// 3-term bf16 split of v written into the A operand of the G MMA (TMEM, two coalitions per 32-bit column).
template <int MODE>
__device__ __forceinline__ void fx_finish4(const float (&S)[4], const float (&C)[4], const float (&b)[4], const float4 kw4,
                                           float inv_bg, float lnull, uint32_t a2_addr, float (&stored)[4]) {
    constexpr bool kBal = MODE == kFxBalanced || MODE == kFxBalancedLogit;
    const float kw[4] = {kw4.x, kw4.y, kw4.z, kw4.w};
    float v[4];
#pragma unroll
    for (int mm = 0; mm < 4; ++mm) {
        const float s = kBal ? S[mm] * b[mm] : S[mm];         // sigma sums were gathered as sum 1/d: scale by b = min(1, 2^-t)
        float y;
        if (MODE == kFxBalancedLogit) {                        // both sums known: keep the smaller one (sign tells which)
            stored[mm] = fdb_kshap_encode_ey(s, C[mm], inv_bg);
            y = fdb_kshap_link(stored[mm], 1);
        } else {
            stored[mm] = s * inv_bg;
            y = MODE == kFxClampedPairs ? fdb_kshap_link(stored[mm], 1) : stored[mm];
        }
        v[mm] = kw[mm] * (y - lnull);
    }
#pragma unroll
    for (int sp = 0; sp < 3; ++sp) {
        const uint32_t p01 = ptx::pack_bf16x2(v[1], v[0]), p23 = ptx::pack_bf16x2(v[3], v[2]);
        ptx::tmem_st2(a2_addr + sp * 64, p01, p23);
        if (sp < 2) {
            v[0] -= __uint_as_float(p01 << 16); v[1] -= __uint_as_float(p01 & 0xffff0000u);
            v[2] -= __uint_as_float(p23 << 16); v[3] -= __uint_as_float(p23 & 0xffff0000u);
        }
    }
}

template <int MODE, bool SINGLE>
__global__ void __launch_bounds__(kFxThreads, 1) kshap_eval_fx_kernel(const FxParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* sE = smem;                                   // [2][65536]
    unsigned char* sZ = sE + 2 * kEStageBytes;                  // [2][8192]   coalition tile image (B operand of the D MMA)
    unsigned char* sX = sZ + 2 * kTileBytes;                    // [2][9216]   x^T image (B operand of the G MMA) + o_m + kw_m
    unsigned char* sQ = sX + 2 * kXStageBytes;                  // [2][24576]  row block (A operand of the D MMA)
    FxBars* bars = reinterpret_cast<FxBars*>(sQ + 2 * kQBytes);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const long long n_units = p.n_units;
    const long long u0 = (long long)blockIdx.x * n_units / gridDim.x;
    const long long u1 = ((long long)blockIdx.x + 1) * n_units / gridDim.x;
    const int T = p.tiles;
    const int cbs_last = (fx_cbw(p.n_bg, p.n_chunks - 1) + 3) & ~3;
    const size_t tile_floats = (size_t)128 * (128 * (p.n_chunks - 1) + cbs_last);

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(&bars->z_full[i], 1);
            ptx::mbar_init(&bars->z_empty[i], 1);
            ptx::mbar_init(&bars->q_full[i], 2);
            ptx::mbar_init(&bars->q_empty[i], 1);
            ptx::mbar_init(&bars->t_full[i], 1);
            ptx::mbar_init(&bars->t_empty[i], kFxComputeWarps);
            ptx::mbar_init(&bars->e_full[i], 1);
            ptx::mbar_init(&bars->e_empty[i], kFxComputeWarps);
        }
        ptx::mbar_init(&bars->a2_full, kFxComputeWarps);
        ptx::mbar_init(&bars->g_full, 1);
        ptx::mbar_fence_init();
    }
    if (warp == 1) ptx::tmem_alloc(&bars->tmem_base, 512);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ================================================================ TMA: tile images + E chunks
        if (lane == 0) {
            auto issue_tile = [&](int kk, long long unit) {
                const int s = kk & 1;
                const int tile = (int)(unit % T);
                mbar_wait_relaxed(&bars->z_empty[s], ((kk >> 1) & 1) ^ 1);
                ptx::mbar_arrive_expect_tx(&bars->z_full[s], kTileBytes + kXStageBytes);
                ptx::bulk_g2s(sZ + s * kTileBytes, reinterpret_cast<const unsigned char*>(p.zimg) + (size_t)tile * kTileBytes,
                              kTileBytes, &bars->z_full[s]);
                ptx::bulk_g2s(sX + s * kXStageBytes, p.xt + (size_t)tile * kXStageBytes, kXStageBytes, &bars->z_full[s]);
            };
            int k = 0;
            uint32_t j = 0;
            if (u0 < u1) issue_tile(0, u0);
            for (long long u = u0; u < u1; ++u, ++k) {
                // E chunks of this unit first (their stages free up one unit earlier than the tile stage of unit k + 1)
                const float* et = p.E + (size_t)(u % T) * tile_floats;
                for (int c = 0; c < p.n_chunks; ++c, ++j) {
                    const int es = j & 1;
                    const uint32_t bytes = 128u * (uint32_t)((fx_cbw(p.n_bg, c) + 3) & ~3) * 4u;
                    mbar_wait_relaxed(&bars->e_empty[es], ((j >> 1) & 1) ^ 1);
                    ptx::mbar_arrive_expect_tx(&bars->e_full[es], bytes);
                    ptx::bulk_g2s(sE + es * kEStageBytes, et + (size_t)c * 128 * 128, bytes, &bars->e_full[es]);
                }
                if (u + 1 < u1) issue_tile(k + 1, u + 1);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ================================================================ MMA issue
        if (lane == 0 && u0 < u1) {
            const uint32_t idesc_d = ptx::make_idesc_f16(p.fmt, 128, 128);
            const uint32_t idesc_g = ptx::make_idesc_f16(1, 128, 32);        // v splits and x^T are bf16
            const long long rb0 = u0 / T;
            auto issue_d = [&](int kk, long long unit) {
                const int s = kk & 1;
                const uint32_t ph = (kk >> 1) & 1;
                const long long rb = unit / T;
                const int qi = (int)(rb - rb0), qs = qi & 1;
                mbar_wait_relaxed(&bars->z_full[s], ph);
                if (unit == u0 || unit % T == 0) mbar_wait_relaxed(&bars->q_full[qs], (qi >> 1) & 1);
                mbar_wait_relaxed(&bars->t_empty[s], ph ^ 1);
                ptx::tc_fence_after();
                const uint32_t d = tmem_base + s * 128;
                const uint32_t qa = ptx::smem_u32(sQ + qs * kQBytes), za = ptx::smem_u32(sZ + s * kTileBytes);
#pragma unroll
                for (int sp = 0; sp < 3; ++sp)
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks)
                        ptx::mma_f16(d, ptx::make_desc_kmajor_noswizzle(qa + sp * kTileBytes + ks * 256, kLbo, kSbo),
                                     ptx::make_desc_kmajor_noswizzle(za + ks * 256, kLbo, kSbo), idesc_d, (sp | ks) != 0);
                ptx::mma_commit(&bars->t_full[s]);
                if (unit + 1 == u1 || (unit + 1) % T == 0) ptx::mma_commit(&bars->q_empty[qs]);   // last tile of this row block here
            };
            issue_d(0, u0);
            int k = 0;
            for (long long u = u0; u < u1; ++u, ++k) {
                if (u + 1 < u1) issue_d(k + 1, u + 1);
                // G[r][j] (+)= v[r][m] x[m][j]: A = the three bf16 terms of v in TMEM, B = x^T image of this tile
                mbar_wait_relaxed(&bars->a2_full, k & 1);
                ptx::tc_fence_after();
                const uint32_t xa = ptx::smem_u32(sX + (k & 1) * kXStageBytes);
#pragma unroll
                for (int sp = 0; sp < 3; ++sp)
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk)
                        ptx::mma_f16_ts(tmem_base + kTmemG, tmem_base + kTmemA2 + sp * 64 + kk * 8,
                                        ptx::make_desc_kmajor_noswizzle(xa + kk * 256, kLbo, kSboX), idesc_g, (sp | kk) != 0);
                ptx::mma_commit(&bars->g_full);
                ptx::mma_commit(&bars->z_empty[k & 1]);
            }
        }
        __syncwarp();
    } else if (warp < 4) {
        // ================================================================ Q block builders (64 rows each), once per row block
        const int w = warp - 2;
        const float wk = lane < FDB_F ? 1.4426950408889634f * p.wq[lane] : 0.f;
        if (u0 < u1) {
            const long long rb0 = u0 / T, rb1 = (u1 - 1) / T;
            int qi = 0;
            for (long long rb = rb0; rb <= rb1; ++rb, ++qi) {
                const int s = qi & 1;
                const int64_t row0 = rb * 128 + w * 64;
                mbar_wait_relaxed(&bars->q_empty[s], ((qi >> 1) & 1) ^ 1);
                unsigned char* qb = sQ + s * kQBytes + (lane >> 3) * 128 + (lane & 7) * 2;
#pragma unroll 4
                for (int i = 0; i < 64; ++i) {
                    const int64_t gr = row0 + i;
                    float v = (lane < FDB_F && gr < p.n_rows) ? wk * p.x[(size_t)gr * FDB_F + lane] : 0.f;
                    v = fminf(fmaxf(v, -30000.f), 30000.f);
                    float tot = v;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
                    if (lane == 30) v = fminf(fmaxf(-tot, -60000.f), 60000.f);
                    const int r = w * 64 + i;
                    unsigned char* dst = qb + (r >> 3) * 512 + (r & 7) * 16;
                    float rem = v;
#pragma unroll
                    for (int jj = 0; jj < 3; ++jj) {
                        const uint16_t h = fx_to_h16(rem, p.fmt);
                        *reinterpret_cast<uint16_t*>(dst + jj * kTileBytes) = h;
                        rem -= fx_from_h16(h, p.fmt);
                    }
                }
                ptx::fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&bars->q_full[s]);
            }
        }
    } else {
        // ================================================================ compute: lane = explained row
        const int qd = warp & 3;
        const int cg = (warp - 4) >> 2;
        const uint32_t lane_base = (uint32_t)(qd * 32) << 16;
        const uint32_t g_addr = tmem_base + lane_base + kTmemG + cg * 8;
        float gacc[8];                                  // this thread's share of X'W y: row = lane, features 8 cg .. 8 cg + 7
#pragma unroll
        for (int i = 0; i < 8; ++i) gacc[i] = 0.f;
        bool pending = false;                           // G of the previous unit has not been folded into gacc yet
        long long prev_rb = -1;
        int k = 0;
        uint32_t j = 0;

        // fold G of unit kk (finished by the MMA warp) into gacc
        auto take_g = [&](int kk) {
            ptx::mbar_wait(&bars->g_full, kk & 1);
            ptx::tc_fence_after();
            uint32_t gr[8];
            ptx::tmem_ld8(g_addr, gr);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 8; ++i) gacc[i] += __uint_as_float(gr[i]);
        };
        // one 128-byte record per (row, CTA that worked on its row block); K3 folds the records of a row in slot order
        auto flush_g = [&](long long rb) {
            const long long row = rb * 128 + qd * 32 + lane;
            if (row < p.n_rows) {
                const int slot = (int)((long long)blockIdx.x - fx_owner(rb * T, n_units, gridDim.x));
                float4* dst = reinterpret_cast<float4*>(p.gpart + ((size_t)row * p.slots + slot) * 32 + cg * 8);
                dst[0] = make_float4(gacc[0], gacc[1], gacc[2], gacc[3]);
                dst[1] = make_float4(gacc[4], gacc[5], gacc[6], gacc[7]);
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) gacc[i] = 0.f;
        };

        for (long long u = u0; u < u1; ++u, ++k) {
            const int s = k & 1;
            const int tile = (int)(u % T);
            const long long rb = u / T;
            const long long row = rb * 128 + qd * 32 + lane;
            ptx::mbar_wait(&bars->z_full[s], (k >> 1) & 1);          // o_m / kw_m of this tile (written by the TMA)
            ptx::mbar_wait(&bars->t_full[s], (k >> 1) & 1);
            ptx::tc_fence_after();
            const uint32_t ta = tmem_base + lane_base + s * 128 + cg * 32;
            const uint32_t a2 = tmem_base + lane_base + kTmemA2 + cg * 16;
            const float4* op = reinterpret_cast<const float4*>(sX + s * kXStageBytes + kTileBytes) + cg * 8;   // o_m
            const float4* kwp = op + 32;                                                                        // kw_m
            float4* eyd = p.ey != nullptr && row < p.n_rows
                              ? reinterpret_cast<float4*>(p.ey + (size_t)row * p.m_pad + tile * 128 + cg * 32) : nullptr;

            // before the first write into the v operand of this unit: the G MMA of the previous unit must be done with
            // it (g_full); its result joins gacc, and a finished row block is written out
            auto before_first_store = [&]() {
                if (pending) {
                    take_g(k - 1);
                    pending = false;
                    if (rb != prev_rb) flush_g(prev_rb);
                }
            };
            auto load_t = [&](int g4, uint32_t (&rr)[4]) { ptx::tmem_ld4(ta + 4 * g4, rr); };
            auto exps = [&](const uint32_t (&rr)[4], int g4, float (&t4)[4]) {
                const float4 o = op[g4];
                t4[0] = fminf(fmaxf(__uint_as_float(rr[0]) + o.x, -126.f), 126.f);
                t4[1] = fminf(fmaxf(__uint_as_float(rr[1]) + o.y, -126.f), 126.f);
                t4[2] = fminf(fmaxf(__uint_as_float(rr[2]) + o.z, -126.f), 126.f);
                t4[3] = fminf(fmaxf(__uint_as_float(rr[3]) + o.w, -126.f), 126.f);
            };

            if (SINGLE) {
                // the whole background fits one E chunk: every group of four coalitions is finished in one go, so no
                // per-coalition partial sums live in registers; the exponents of the next group are fetched from TMEM
                // while this group is summed
                const int es = j & 1;
                const int cbw = p.n_bg, cbs = (cbw + 3) & ~3;
                ptx::mbar_wait(&bars->e_full[es], (j >> 1) & 1);
                const float* eb = reinterpret_cast<const float*>(sE + es * kEStageBytes) + (size_t)cg * 32 * cbs;
                uint32_t rr[4];
                load_t(0, rr);
#pragma unroll
                for (int g4 = 0; g4 < 8; ++g4) {
                    float t4[4], a[4], b[4], S[4], C[4] = {0.f, 0.f, 0.f, 0.f}, st[4];
                    ptx::tmem_ld_wait();
                    exps(rr, g4, t4);
                    if (g4 < 7) load_t(g4 + 1, rr);
                    fx_scales<MODE>(t4, a, b);
                    if (MODE == kFxBalancedLogit) chunk_sum4_logit(eb + (size_t)g4 * 4 * cbs, cbs, cbw >> 2, cbw & 3, a, b, S, C);
                    else chunk_sum4<MODE, FDB_FX_QUAD_UNROLL_SINGLE>(eb + (size_t)g4 * 4 * cbs, cbs, cbw >> 2, cbw & 3, a, b, S);
                    if (g4 == 0) before_first_store();
                    fx_finish4<MODE>(S, C, b, kwp[g4], p.inv_bg, p.lnull, a2 + 2 * g4, st);
                    if (eyd != nullptr) eyd[g4] = make_float4(st[0], st[1], st[2], st[3]);
                }
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&bars->e_empty[es]);
                ++j;
            } else {
                float part[32];
                float partc[MODE == kFxBalancedLogit ? 32 : 1];      // complement sums (logit link only)
#pragma unroll
                for (int mi = 0; mi < 32; ++mi) part[mi] = 0.f;
#pragma unroll
                for (int mi = 0; mi < (MODE == kFxBalancedLogit ? 32 : 1); ++mi) partc[mi] = 0.f;
                for (int c = 0; c < p.n_chunks; ++c, ++j) {
                    const int es = j & 1;
                    const int cbw = fx_cbw(p.n_bg, c);
                    const int cbs = (cbw + 3) & ~3;
                    ptx::mbar_wait(&bars->e_full[es], (j >> 1) & 1);
                    const float* eb = reinterpret_cast<const float*>(sE + es * kEStageBytes) + (size_t)cg * 32 * cbs;
                    uint32_t rr[4];
                    if (FDB_FX_PREFETCH_MULTI) load_t(0, rr);
#pragma unroll
                    for (int g4 = 0; g4 < 8; ++g4) {
                        float t4[4], a[4], b[4], S[4], C[4];
                        if (!FDB_FX_PREFETCH_MULTI) load_t(g4, rr);
                        ptx::tmem_ld_wait();
                        exps(rr, g4, t4);
                        if (FDB_FX_PREFETCH_MULTI && g4 < 7) load_t(g4 + 1, rr);
                        fx_scales<MODE>(t4, a, b);
                        if (MODE == kFxBalancedLogit) {
                            chunk_sum4_logit(eb + (size_t)g4 * 4 * cbs, cbs, cbw >> 2, cbw & 3, a, b, S, C);
#pragma unroll
                            for (int mm = 0; mm < 4; ++mm) partc[MODE == kFxBalancedLogit ? 4 * g4 + mm : 0] += C[mm];
                        } else {
                            chunk_sum4<MODE, FDB_FX_QUAD_UNROLL>(eb + (size_t)g4 * 4 * cbs, cbs, cbw >> 2, cbw & 3, a, b, S);
                        }
#pragma unroll
                        for (int mm = 0; mm < 4; ++mm) part[4 * g4 + mm] += S[mm];
                    }
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(&bars->e_empty[es]);
                }
                uint32_t rr[4];
                load_t(0, rr);
#pragma unroll
                for (int g4 = 0; g4 < 8; ++g4) {
                    float t4[4], a[4], b[4], S[4], C[4], st[4];
                    ptx::tmem_ld_wait();
                    exps(rr, g4, t4);
                    if (g4 < 7) load_t(g4 + 1, rr);
#pragma unroll
                    for (int mm = 0; mm < 4; ++mm) {
                        b[mm] = ptx::ex2_approx(fminf(-t4[mm], 0.f));
                        S[mm] = part[4 * g4 + mm];
                        C[mm] = MODE == kFxBalancedLogit ? partc[MODE == kFxBalancedLogit ? 4 * g4 + mm : 0] : 0.f;
                    }
                    (void)a;
                    if (g4 == 0) before_first_store();
                    fx_finish4<MODE>(S, C, b, kwp[g4], p.inv_bg, p.lnull, a2 + 2 * g4, st);
                    if (eyd != nullptr) eyd[g4] = make_float4(st[0], st[1], st[2], st[3]);
                }
            }
            // hand the exponents' TMEM stage back to the MMA warp and the finished v operand over to it
            ptx::tmem_st_wait();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                ptx::mbar_arrive(&bars->t_empty[s]);
                ptx::mbar_arrive(&bars->a2_full);
            }
            pending = true;
            prev_rb = rb;
        }
        if (pending) {
            take_g(k - 1);
            flush_g(prev_rb);
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, 512);
    }
}

// x^T image of every coalition tile (B operand of the G MMA: 32 feature rows x 128 coalitions, bf16, canonical K-major
// layout in 8-row groups of 16 K-chunks) followed by the tile's exponent centres o_m and kernel weights kw_m.
__global__ void kshap_build_fx_tiles(const uint32_t* __restrict__ bits, const float* __restrict__ kw, const float* __restrict__ eoff,
                                     int m_pad, unsigned char* __restrict__ out) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= m_pad) return;
    const uint32_t b = bits[m];
    const float w = kw[m];
    const int z29 = (b >> 29) & 1u;
    unsigned char* tile = out + (size_t)(m >> 7) * FDB_FX_TILE_BYTES;
    const int ml = m & 127;
    for (int jf = 0; jf < 32; ++jf) {
        // columns 0..28: x_mj = z_mj - z_m29 (the WLS regressors); column 29: z_m29 itself, so that the fused GEMM also leaves
        // T_29 = sum kw z_m29 e_m behind and num_features(k) can rebuild every T_j = sum kw z_mj e_m from the records
        int x = jf < FDB_F - 1 ? (int)((b >> jf) & 1u) - z29 : (jf == FDB_F - 1 ? z29 : 0);
        if (w == 0.f) x = 0;                                        // padding coalitions
        const uint16_t h = x > 0 ? 0x3F80 : (x < 0 ? 0xBF80 : 0);   // +1 / -1 / 0 in bf16
        *reinterpret_cast<uint16_t*>(tile + (jf >> 3) * 2048 + (ml >> 3) * 128 + (jf & 7) * 16 + (ml & 7) * 2) = h;
    }
    float* meta = reinterpret_cast<float*>(tile + 8192);
    meta[ml] = eoff[m];
    meta[128 + ml] = w;
}

// One block per coalition: exponents e_b = -log2e (c + sum_k (1 - z_mk) w_k bg_bk) in f64, centred per
// coalition, stored as E = 2^(e_b - o_m) in the [tile][chunk][128][cbs] layout the evaluator streams.
__global__ void __launch_bounds__(128) kshap_build_E(const uint32_t* __restrict__ bits, const double* __restrict__ Pd,
                                                     int n_bg, int n_chunks, float* __restrict__ E,
                                                     float* __restrict__ eoff, float* __restrict__ half_spread,
                                                     float* __restrict__ quad_logsum) {
    __shared__ double s_min[128], s_max[128];
    const int m = blockIdx.x, tid = threadIdx.x;
    const uint32_t b = bits[m];
    double ev[8];
    double mn = 1e300, mx = -1e300;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int col = tid + i * 128;
        ev[i] = 0.0;
        if (col < n_bg) {
            const double* pr = Pd + (size_t)col * 32;
            double g = pr[30];
            for (int k = 0; k < FDB_F; ++k) g += ((b >> k) & 1u) ? 0.0 : pr[k];
            ev[i] = -1.4426950408889634074 * g;
            mn = fmin(mn, ev[i]);
            mx = fmax(mx, ev[i]);
        }
    }
    s_min[tid] = mn;
    s_max[tid] = mx;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if (tid < o) { s_min[tid] = fmin(s_min[tid], s_min[tid + o]); s_max[tid] = fmax(s_max[tid], s_max[tid + o]); }
        __syncthreads();
    }
    const double centre = 0.5 * (s_min[0] + s_max[0]);
    if (tid == 0) {
        eoff[m] = (float)centre;
        half_spread[m] = (float)(0.5 * (s_max[0] - s_min[0]));
    }
    // The clamp-free inner loop multiplies the four d = E a + b (a, b <= 1, one of them 1) of a column quad:
    // prod <= 2^(4 + sum max(lev, 0)) and prod >= 2^(-sum max(-lev, 0)) with lev = log2 E.  Record the worst quad.
    const int full_cols = n_bg & ~3;
    double qs = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int col = tid + i * 128;
        const double lev = col < full_cols ? ev[i] - centre : 0.0;
        double up = fmax(lev, 0.0), dn = fmax(-lev, 0.0);
        up += __shfl_xor_sync(0xffffffffu, up, 1); dn += __shfl_xor_sync(0xffffffffu, dn, 1);
        up += __shfl_xor_sync(0xffffffffu, up, 2); dn += __shfl_xor_sync(0xffffffffu, dn, 2);
        qs = fmax(qs, fmax(up, dn));
    }
    __syncthreads();
    s_max[tid] = qs;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if (tid < o) s_max[tid] = fmax(s_max[tid], s_max[tid + o]);
        __syncthreads();
    }
    if (tid == 0) quad_logsum[m] = (float)s_max[0];
    const int cbs_last = (fx_cbw(n_bg, n_chunks - 1) + 3) & ~3;
    const size_t tile_floats = (size_t)128 * (128 * (n_chunks - 1) + cbs_last);
    float* et = E + (size_t)(m >> 7) * tile_floats;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int col = tid + i * 128;
        if (col < n_bg) {
            const int c = col >> 7;
            const int cbs = (fx_cbw(n_bg, c) + 3) & ~3;
            const double ex = fmin(fmax(ev[i] - centre, -120.0), 120.0);
            et[(size_t)c * 128 * 128 + (size_t)(m & 127) * cbs + (col & 127)] = (float)exp2(ex);
        }
    }
}

}  // namespace

// Build the factored operands of a single-set plan (called from fdb_kshap_plan_create).
int fdb_kshap_build_factored(fdb_kshap_plan* pl, const double* bg_host) {
    const int n_chunks = (pl->n_bg + 127) / 128;
    const int cbs_last = ((pl->n_bg - 128 * (n_chunks - 1)) + 3) & ~3;
    const size_t tile_floats = (size_t)128 * (128 * (n_chunks - 1) + cbs_last);
    const size_t total = tile_floats * (pl->m_pad / 128);
    double* h = new (std::nothrow) double[(size_t)pl->n_bg * 32];
    FDB_REQUIRE(h != nullptr, FDB_ERR_NOMEM, "out of host memory");
    // Column order of E (the mean over the background does not care): rows sorted by their logit are dealt out so
    // that every quad of 4 consecutive columns takes one row from each quarter of the sorted order.  A quad then
    // never holds four large (or four small) exponents at once, which keeps the product of its four denominators
    // inside fp32 range without clamps (kshap_build_E measures the actual bound).  The n_bg % 4 last rows form the tail.
    std::vector<std::pair<double, int>> order(pl->n_bg);
    for (int b = 0; b < pl->n_bg; ++b) {
        double z = pl->intercept;
        for (int k = 0; k < FDB_F; ++k) z += pl->coef[k] * bg_host[(size_t)b * FDB_F + k];
        order[b] = {z, b};
    }
    std::sort(order.begin(), order.end());
    const int n_quads = pl->n_bg / 4;
    for (int i = 0; i < pl->n_bg; ++i) {
        const int col = i < 4 * n_quads ? 4 * (i % n_quads) + i / n_quads : i;
        const int b = order[i].second;
        for (int k = 0; k < 32; ++k) h[(size_t)col * 32 + k] = 0.0;
        for (int k = 0; k < FDB_F; ++k) h[(size_t)col * 32 + k] = pl->coef[k] * bg_host[(size_t)b * FDB_F + k];
        h[(size_t)col * 32 + 30] = pl->intercept;
    }
    double* dP = nullptr;
    cudaError_t e = cudaMalloc((void**)&dP, (size_t)pl->n_bg * 32 * 8);
    if (e == cudaSuccess) e = cudaMemcpy(dP, h, (size_t)pl->n_bg * 32 * 8, cudaMemcpyHostToDevice);
    delete[] h;
    if (e == cudaSuccess) e = cudaMalloc((void**)&pl->d_E, total * 4);
    if (e == cudaSuccess) e = cudaMemset(pl->d_E, 0, total * 4);
    if (e == cudaSuccess) e = cudaMalloc((void**)&pl->d_eoff, (size_t)pl->m_pad * 4);
    float* d_spread = nullptr;
    if (e == cudaSuccess) e = cudaMalloc((void**)&d_spread, (size_t)pl->m_pad * 2 * 4);
    if (e == cudaSuccess) {
        kshap_build_E<<<pl->m_pad, 128>>>(pl->d_bits, dP, pl->n_bg, n_chunks, pl->d_E, pl->d_eoff, d_spread, d_spread + pl->m_pad);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    // The factorisation needs every coalition's background exponents within +-100 of their centre (fp32 range of
    // E and u).  Backgrounds that spread the logit further (outliers of ~+-140 in logit) keep the general evaluator.
    float worst = 0.f, worst_quad = 0.f;
    if (e == cudaSuccess) {
        float* hs = new (std::nothrow) float[(size_t)pl->m_pad * 2];
        if (hs && (e = cudaMemcpy(hs, d_spread, (size_t)pl->m_pad * 2 * 4, cudaMemcpyDeviceToHost)) == cudaSuccess)
            for (int m = 0; m < pl->m_pad; ++m) {
                worst = fmaxf(worst, hs[m]);
                worst_quad = fmaxf(worst_quad, hs[pl->m_pad + m]);
            }
        delete[] hs;
    }
    pl->fx_quad_logsum = worst_quad;
    cudaFree(d_spread);
    cudaFree(dP);
    FDB_CUDA(e);
    pl->fx_half_spread = worst;
    if (worst > 100.f) {
        cudaFree(pl->d_E);
        pl->d_E = nullptr;
        return FDB_OK;
    }
    // per-tile operands of the fused X'W y GEMM: x^T image, exponent centres, kernel weights
    if (!pl->d_fxtile) FDB_CUDA(cudaMalloc((void**)&pl->d_fxtile, (size_t)(pl->m_pad / 128) * FDB_FX_TILE_BYTES));
    kshap_build_fx_tiles<<<(pl->m_pad + 127) / 128, 128>>>(pl->d_bits, pl->d_kw, pl->d_eoff, pl->m_pad, pl->d_fxtile);
    FDB_CUDA(cudaGetLastError());
    FDB_CUDA(cudaDeviceSynchronize());
    return FDB_OK;
}

// Launch geometry of the factored evaluator for n_rows rows: units, CTAs and X'W y records per row (K3 needs the same
// three numbers to find the records of a row).
void fdb_kshap_fx_geometry(const fdb_kshap_plan* pl, int n_rows, long long* n_units, int* grid, int* slots) {
    const int tiles = pl->m_pad / 128;
    const long long units = (long long)((n_rows + 127) / 128) * tiles;
    long long g = units < pl->ctx->sm_count ? units : pl->ctx->sm_count;
    if (g < 1) g = 1;
    *n_units = units;
    *grid = (int)g;
    // a row block (tiles consecutive units) meets at most 2 CTAs once every CTA's range holds >= tiles units
    *slots = (units / g >= tiles) ? 2 : tiles;
}

int fdb_kshap_launch_eval_fx(fdb_kshap_plan* pl, cudaStream_t st, const float* x, int n_rows, float* ey, float* gpart) {
    FDB_REQUIRE(pl->d_E != nullptr && pl->d_fxtile != nullptr, FDB_ERR_STATE,
                "the factored evaluator is not available for this plan (needs one coalition set whose background logits "
                "stay within +-100 log2 units per coalition)");
    FxParams p;
    p.x = x; p.zimg = pl->d_aimg; p.xt = pl->d_fxtile; p.E = pl->d_E; p.wq = pl->d_wq; p.ey = ey; p.gpart = gpart;
    p.n_rows = n_rows; p.tiles = pl->m_pad / 128; p.m_pad = pl->m_pad; p.n_bg = pl->n_bg; p.fmt = pl->fmt;
    p.n_chunks = (pl->n_bg + 127) / 128;
    p.inv_bg = 1.0f / (float)pl->n_bg;
    p.lnull = pl->link == 1 ? (float)log(pl->fnull / (1.0 - pl->fnull)) : (float)pl->fnull;
    int grid = 1;
    fdb_kshap_fx_geometry(pl, n_rows, &p.n_units, &grid, &p.slots);
    const size_t smem = 2 * kEStageBytes + 2 * kTileBytes + 2 * kXStageBytes + 2 * kQBytes + sizeof(FxBars) + 64;
    static bool configured[FDB_MAX_DEVICES] = {};          // function attributes are per device
    if (!configured[pl->ctx->device]) {
#define FDB_FX_ATTR(MODE)                                                                                                       \
        FDB_CUDA(cudaFuncSetAttribute(kshap_eval_fx_kernel<MODE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        FDB_CUDA(cudaFuncSetAttribute(kshap_eval_fx_kernel<MODE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem))
        FDB_FX_ATTR(kFxClampedQuads);
        FDB_FX_ATTR(kFxClampedPairs);
        FDB_FX_ATTR(kFxBalanced);
        FDB_FX_ATTR(kFxBalancedLogit);
#undef FDB_FX_ATTR
        configured[pl->ctx->device] = true;
    }
    // balanced (clamp-free) inner loop whenever the product of the four denominators of every column quad is
    // guaranteed to stay within 2^+-114 (kshap_build_E)
    const char* force = getenv("FDB_FX_CLAMPED");
    const bool balanced = pl->fx_quad_logsum <= 110.f && pl->fx_half_spread <= 100.f && !(force && atoi(force) != 0);
    const int mode = balanced ? (pl->link == 1 ? kFxBalancedLogit : kFxBalanced) : (pl->link == 1 ? kFxClampedPairs : kFxClampedQuads);
    const bool single = p.n_chunks == 1;
#define FDB_FX_LAUNCH(MODE)                                                                            \
    do {                                                                                               \
        if (single) kshap_eval_fx_kernel<MODE, true><<<grid, kFxThreads, smem, st>>>(p);              \
        else kshap_eval_fx_kernel<MODE, false><<<grid, kFxThreads, smem, st>>>(p);                    \
    } while (0)
    switch (mode) {
        case kFxClampedQuads: FDB_FX_LAUNCH(kFxClampedQuads); break;
        case kFxClampedPairs: FDB_FX_LAUNCH(kFxClampedPairs); break;
        case kFxBalanced: FDB_FX_LAUNCH(kFxBalanced); break;
        default: FDB_FX_LAUNCH(kFxBalancedLogit); break;
    }
#undef FDB_FX_LAUNCH
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}
