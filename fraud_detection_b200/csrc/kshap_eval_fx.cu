// K2f - factored Kernel SHAP evaluator for a SHARED coalition set (sm_100a, tcgen05 + TMEM + TMA).
//
// With one coalition set for every explained row, the background part of the logit does not depend on the
// row:  logit[r][m][b] = G[m][b] + c + s[r][m],  G = ((1-Z) diag(w)) bg^T,  s[r][m] = sum_k z_mk w_k x_rk.
// So  2^(-log2e logit) = E[m][b] * u[r][m]  with  E = 2^(-log2e (G + c) - o_m)  built ONCE per plan and
// u = 2^(-log2e s + o_m)  one exponential per (row, coalition) instead of one per (row, coalition, background):
//     ey[r][m] = mean_b 1 / (1 + E[m][b] u[r][m])
// (o_m centres the exponents of coalition m so E and u stay inside fp32 range.)
//
// The per-row term is a GEMM and runs on the tensor cores:  D[r][m] = Q[r][:] . Abar[m][:]  with
//   Q    (128 rows x 32, smem, A operand) = 3-term f16/bf16 split of [log2e w_k x_rk | -log2e sum_k w_k x_rk | 0]
//   Abar (128 coalitions x 32, smem, B operand) = [1 - z_mk | 1 | 0]   (the same tile image K2 uses as A)
//   D    (TMEM, fp32, lane = explained row, column = coalition) = -log2e s[r][m]
// Each compute thread owns one explained row (TMEM lane) and 32 coalitions; E[m][b] is the same for the
// whole warp, so it is read from shared memory as a broadcast LDS.128 and the inner loop is
//   d = min(fma(E, u, 1), 2^30);  four d share one reciprocal       (4.5 issue slots, 0.25 SFU ops per sigmoid)
// against 6.25 issue slots and 1.25 SFU ops in the general evaluator.  E streams through a double-buffered
// TMA ring in background chunks of <= 128 columns; partial sums stay in registers across chunks.
//
// CTA = 20 warps, persistent, one per SM; work unit = (block of 128 rows, tile of 128 coalitions):
//   warp 0 (one lane)  TMA: coalition tile image per unit, E chunks
//   warp 1 (one lane)  tcgen05.mma issue (6 per unit) + commits
//   warps 2-3          build the Q block (x rows -> scaled, split, canonical K-major layout)
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
constexpr int kTileBytes = 128 * 32 * 2;        // 8192: one 128 x 32 operand image
constexpr int kQBytes = 3 * kTileBytes;          // three split terms
constexpr int kEStageBytes = 128 * 128 * 4;      // 65536: 128 coalitions x <= 128 background columns
constexpr uint32_t kLbo = 128, kSbo = 512;

struct FxParams {
    const float* x;
    const uint16_t* zimg;      // coalition tile images (plan->d_aimg, set 0)
    const float* E;            // [tile][chunk][128][cbs]
    const float* eoff;         // [m_pad] exponent centre o_m
    const float* wq;
    const float* xw;           // [m_pad][32] kw_m * x_mj
    float* ey;                 // optional [n_rows][m_pad]
    float* gpart;              // [n_rows][m_pad/32][32] partial X'W y
    int n_rows, tiles, m_pad, n_bg, fmt;
    int n_chunks;              // background chunks of 128 columns (last one narrower)
    int link;
    float inv_bg, lnull;
};

struct FxBars {
    uint64_t zq_full[2], zq_empty[2], t_full[2], t_empty[2], e_full[2], e_empty[2];
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
// apart; every lane reads the same addresses -> broadcast LDS.128.  WIDE: logit link, two sigmoids per
// reciprocal and clamp 2^62 so that tiny sigmoids keep their relative accuracy.
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
#define FDB_FX_QUAD_UNROLL 2      // quads of one coalition in flight (x 4 interleaved coalitions)
#endif
#define FDB_FX_PRAGMA(x) _Pragma(#x)
#define FDB_FX_UNROLL(n) FDB_FX_PRAGMA(unroll n)
enum { kFxClampedQuads = 0, kFxClampedPairs = 1, kFxBalanced = 2, kFxBalancedLogit = 3 };

template <int MODE>
__device__ __forceinline__ void chunk_sum4(const float* __restrict__ e0, int cbs, int nquads, int ntail,
                                           const float* __restrict__ t, float* __restrict__ acc) {
    constexpr float kClamp = MODE == kFxClampedPairs ? 4.611686018427388e18f : 1073741824.f;
    // sums of this chunk are gathered apart and added to the running totals once: with ~1000 background rows a
    // single float32 chain drifts by 1e-5 on saturated coalitions (every term 1, every add rounded the same way)
    float a[4], b[4], sum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int mm = 0; mm < 4; ++mm) {
        a[mm] = ptx::ex2_approx(MODE == kFxBalanced ? fminf(t[mm], 0.f) : t[mm]);
        b[mm] = MODE == kFxBalanced ? ptx::ex2_approx(fminf(-t[mm], 0.f)) : 1.0f;
    }
FDB_FX_UNROLL(FDB_FX_QUAD_UNROLL)
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
#pragma unroll
    for (int mm = 0; mm < 4; ++mm) acc[mm] += sum[mm];
}

// MODE 3 body: sigmoid and complement sums for four coalitions over one background chunk (see the mode list above)
__device__ __forceinline__ void chunk_sum4_logit(const float* __restrict__ e0, int cbs, int nquads, int ntail,
                                                 const float* __restrict__ t, float* __restrict__ acc, float* __restrict__ accc) {
    float a[4], b[4], sum[4] = {0.f, 0.f, 0.f, 0.f}, sumc[4] = {0.f, 0.f, 0.f, 0.f};      // per-chunk sums, see chunk_sum4
#pragma unroll
    for (int mm = 0; mm < 4; ++mm) {
        a[mm] = ptx::ex2_approx(fminf(t[mm], 0.f));
        b[mm] = ptx::ex2_approx(fminf(-t[mm], 0.f));
    }
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
#pragma unroll
    for (int mm = 0; mm < 4; ++mm) {
        acc[mm] += sum[mm];
        accc[mm] += sumc[mm];
    }
}

#ifndef FDB_FX_SERVICE_SLEEP_NS
#define FDB_FX_SERVICE_SLEEP_NS 800     // a unit lasts ~35 us and every hand-off is double-buffered
#endif
// waits of the single-lane service warps back off so their polling does not take issue slots from the compute warps
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
    while (!ptx::mbar_try_wait(bar, parity)) __nanosleep(FDB_FX_SERVICE_SLEEP_NS);
}

__device__ __forceinline__ int fx_cbw(int n_bg, int c) { const int r = n_bg - 128 * c; return r < 128 ? r : 128; }

template <int MODE>
__global__ void __launch_bounds__(kFxThreads, 1) kshap_eval_fx_kernel(const FxParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* sE = smem;                                   // [2][65536]
    unsigned char* sZ = sE + 2 * kEStageBytes;                  // [2][8192]
    unsigned char* sQ = sZ + 2 * kTileBytes;                    // [2][24576]
    FxBars* bars = reinterpret_cast<FxBars*>(sQ + 2 * kQBytes);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int row_blocks = (p.n_rows + 127) >> 7;
    const int64_t n_units = (int64_t)row_blocks * p.tiles;
    const int cbs_last = (fx_cbw(p.n_bg, p.n_chunks - 1) + 3) & ~3;
    const size_t tile_floats = (size_t)128 * (128 * (p.n_chunks - 1) + cbs_last);

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(&bars->zq_full[i], 3);
            ptx::mbar_init(&bars->zq_empty[i], 1);
            ptx::mbar_init(&bars->t_full[i], 1);
            ptx::mbar_init(&bars->t_empty[i], kFxComputeWarps);
            ptx::mbar_init(&bars->e_full[i], 1);
            ptx::mbar_init(&bars->e_empty[i], kFxComputeWarps);
        }
        ptx::mbar_fence_init();
    }
    if (warp == 1) ptx::tmem_alloc(&bars->tmem_base, 256);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0) {
        // ================================================================ TMA: coalition tiles + E chunks
        if (lane == 0) {
            auto issue_z = [&](int kk, int64_t unit) {
                const int s = kk & 1;
                mbar_wait_relaxed(&bars->zq_empty[s], ((kk >> 1) & 1) ^ 1);
                ptx::mbar_arrive_expect_tx(&bars->zq_full[s], kTileBytes);
                ptx::bulk_g2s(sZ + s * kTileBytes, reinterpret_cast<const unsigned char*>(p.zimg) + (size_t)(unit % p.tiles) * kTileBytes,
                              kTileBytes, &bars->zq_full[s]);
            };
            int k = 0;
            uint32_t j = 0;
            if ((int64_t)blockIdx.x < n_units) issue_z(0, blockIdx.x);
            for (int64_t u = blockIdx.x; u < n_units; u += gridDim.x, ++k) {
                if (u + gridDim.x < n_units) issue_z(k + 1, u + gridDim.x);
                const float* et = p.E + (size_t)(u % p.tiles) * tile_floats;
                for (int c = 0; c < p.n_chunks; ++c, ++j) {
                    const int es = j & 1;
                    const uint32_t bytes = 128u * (uint32_t)((fx_cbw(p.n_bg, c) + 3) & ~3) * 4u;
                    mbar_wait_relaxed(&bars->e_empty[es], ((j >> 1) & 1) ^ 1);
                    ptx::mbar_arrive_expect_tx(&bars->e_full[es], bytes);
                    ptx::bulk_g2s(sE + es * kEStageBytes, et + (size_t)c * 128 * 128, bytes, &bars->e_full[es]);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ================================================================ MMA issue
        if (lane == 0) {
            const uint32_t idesc = ptx::make_idesc_f16(p.fmt, 128, 128);
            int k = 0;
            for (int64_t u = blockIdx.x; u < n_units; u += gridDim.x, ++k) {
                const int s = k & 1;
                const uint32_t ph = (k >> 1) & 1;
                mbar_wait_relaxed(&bars->zq_full[s], ph);
                mbar_wait_relaxed(&bars->t_empty[s], ph ^ 1);
                ptx::tc_fence_after();
                const uint32_t d = tmem_base + s * 128;
                const uint32_t qa = ptx::smem_u32(sQ + s * kQBytes), za = ptx::smem_u32(sZ + s * kTileBytes);
#pragma unroll
                for (int sp = 0; sp < 3; ++sp)
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks)
                        ptx::mma_f16(d, ptx::make_desc_kmajor_noswizzle(qa + sp * kTileBytes + ks * 256, kLbo, kSbo),
                                     ptx::make_desc_kmajor_noswizzle(za + ks * 256, kLbo, kSbo), idesc, (sp | ks) != 0);
                ptx::mma_commit(&bars->t_full[s]);
                ptx::mma_commit(&bars->zq_empty[s]);
            }
        }
        __syncwarp();
    } else if (warp < 4) {
        // ================================================================ Q block builders (64 rows each)
        const int w = warp - 2;
        const float wk = lane < FDB_F ? 1.4426950408889634f * p.wq[lane] : 0.f;
        int k = 0;
        for (int64_t u = blockIdx.x; u < n_units; u += gridDim.x, ++k) {
            const int s = k & 1;
            const int64_t row0 = (u / p.tiles) * 128 + w * 64;
            mbar_wait_relaxed(&bars->zq_empty[s], ((k >> 1) & 1) ^ 1);
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
            if (lane == 0) ptx::mbar_arrive(&bars->zq_full[s]);
        }
    } else {
        // ================================================================ compute: lane = explained row
        const int qd = warp & 3;
        const int cg = (warp - 4) >> 2;
        int k = 0;
        uint32_t j = 0;
        for (int64_t u = blockIdx.x; u < n_units; u += gridDim.x, ++k) {
            const int s = k & 1;
            const int tile = (int)(u % p.tiles);
            const int64_t row = (u / p.tiles) * 128 + qd * 32 + lane;
            ptx::mbar_wait(&bars->t_full[s], (k >> 1) & 1);
            ptx::tc_fence_after();
            const uint32_t ta = tmem_base + ((uint32_t)(qd * 32) << 16) + s * 128 + cg * 32;
            // The exponents t[r][m] = D + o_m stay in TMEM for the whole unit (the other accumulator stage keeps the
            // MMA warp busy meanwhile) and are fetched four coalitions at a time, so only the 32 running sums live
            // in registers across the background chunks.
            float part[32];
            float partc[MODE == kFxBalancedLogit ? 32 : 1];      // complement sums (logit link only)
#pragma unroll
            for (int mi = 0; mi < 32; ++mi) part[mi] = 0.f;
#pragma unroll
            for (int mi = 0; mi < (MODE == kFxBalancedLogit ? 32 : 1); ++mi) partc[mi] = 0.f;
            const float4* op = reinterpret_cast<const float4*>(p.eoff + tile * 128 + cg * 32);
            for (int c = 0; c < p.n_chunks; ++c, ++j) {
                const int es = j & 1;
                const int cbw = fx_cbw(p.n_bg, c);
                const int cbs = (cbw + 3) & ~3;
                ptx::mbar_wait(&bars->e_full[es], (j >> 1) & 1);
                const float* eb = reinterpret_cast<const float*>(sE + es * kEStageBytes) + (size_t)cg * 32 * cbs;
#pragma unroll
                for (int g4 = 0; g4 < 8; ++g4) {
                    uint32_t rr[4];
                    ptx::tmem_ld4(ta + 4 * g4, rr);
                    ptx::tmem_ld_wait();
                    const float4 o = __ldg(op + g4);
                    float t4[4];
                    t4[0] = fminf(fmaxf(__uint_as_float(rr[0]) + o.x, -126.f), 126.f);
                    t4[1] = fminf(fmaxf(__uint_as_float(rr[1]) + o.y, -126.f), 126.f);
                    t4[2] = fminf(fmaxf(__uint_as_float(rr[2]) + o.z, -126.f), 126.f);
                    t4[3] = fminf(fmaxf(__uint_as_float(rr[3]) + o.w, -126.f), 126.f);
                    if (MODE == kFxBalancedLogit)
                        chunk_sum4_logit(eb + (size_t)g4 * 4 * cbs, cbs, cbw >> 2, cbw & 3, t4, part + 4 * g4, partc + (MODE == kFxBalancedLogit ? 4 * g4 : 0));
                    else
                        chunk_sum4<MODE>(eb + (size_t)g4 * 4 * cbs, cbs, cbw >> 2, cbw & 3, t4, part + 4 * g4);
                }
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&bars->e_empty[es]);
            }
            if (MODE == kFxBalanced || MODE == kFxBalancedLogit) {   // sigma sums were gathered as sum 1/d: scale by b = min(1, 2^-t)
                uint32_t r0[16], r1[16];
                ptx::tmem_ld16(ta, r0);
                ptx::tmem_ld16(ta + 16, r1);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 o = __ldg(op + i);
                    const float oo[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int mi = i * 4 + q;
                        const float t = fminf(fmaxf(__uint_as_float(mi < 16 ? r0[mi] : r1[mi - 16]) + oo[q], -126.f), 126.f);
                        part[mi] *= ptx::ex2_approx(fminf(-t, 0.f));
                        if (MODE == kFxBalancedLogit)     // both sums known: keep the smaller one (sign tells which)
                            part[mi] = fdb_kshap_encode_ey(part[mi], partc[mi], p.inv_bg);
                    }
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&bars->t_empty[s]);
            if (row < p.n_rows) {
                // fused first half of the least-squares solve: this thread's share of X'W y,
                //   g[j] += (kw_m x_mj) * (link(ey[r][m]) - link(fnull))   over its 32 coalitions
                // (two passes of 16 features each: 16 sums + 32 y + a batch of loads in flight fit the register budget,
                //  so the warp-uniform xw loads are issued well ahead of the FMAs that consume them)
                if (MODE != kFxBalancedLogit) {
#pragma unroll
                    for (int mi = 0; mi < 32; ++mi) part[mi] *= p.inv_bg;
                }
                const float4* xwp = reinterpret_cast<const float4*>(p.xw + (size_t)(tile * 128 + cg * 32) * 32);
                float4* gdst = reinterpret_cast<float4*>(p.gpart + ((size_t)row * (p.m_pad / 32) + tile * 4 + cg) * 32);
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    float gq[16];
#pragma unroll
                    for (int jj = 0; jj < 16; ++jj) gq[jj] = 0.f;
#pragma unroll
                    for (int mi = 0; mi < 32; ++mi) {
                        // the link is a property of the instantiation (modes 1 and 3 serve the logit link): no dead logit code in the others
                        const float yv = fdb_kshap_link(part[mi], (MODE == kFxClampedPairs || MODE == kFxBalancedLogit) ? 1 : 0) - p.lnull;
                        const float4 x0 = __ldg(xwp + mi * 8 + half * 4), x1 = __ldg(xwp + mi * 8 + half * 4 + 1);
                        const float4 x2 = __ldg(xwp + mi * 8 + half * 4 + 2), x3 = __ldg(xwp + mi * 8 + half * 4 + 3);
                        gq[0] = fmaf(x0.x, yv, gq[0]); gq[1] = fmaf(x0.y, yv, gq[1]); gq[2] = fmaf(x0.z, yv, gq[2]); gq[3] = fmaf(x0.w, yv, gq[3]);
                        gq[4] = fmaf(x1.x, yv, gq[4]); gq[5] = fmaf(x1.y, yv, gq[5]); gq[6] = fmaf(x1.z, yv, gq[6]); gq[7] = fmaf(x1.w, yv, gq[7]);
                        gq[8] = fmaf(x2.x, yv, gq[8]); gq[9] = fmaf(x2.y, yv, gq[9]); gq[10] = fmaf(x2.z, yv, gq[10]); gq[11] = fmaf(x2.w, yv, gq[11]);
                        gq[12] = fmaf(x3.x, yv, gq[12]); gq[13] = fmaf(x3.y, yv, gq[13]); gq[14] = fmaf(x3.z, yv, gq[14]); gq[15] = fmaf(x3.w, yv, gq[15]);
                    }
#pragma unroll
                    for (int v4 = 0; v4 < 4; ++v4)
                        gdst[half * 4 + v4] = make_float4(gq[4 * v4], gq[4 * v4 + 1], gq[4 * v4 + 2], gq[4 * v4 + 3]);
                }
                if (p.ey != nullptr) {
                    float4* dst = reinterpret_cast<float4*>(p.ey + (size_t)row * p.m_pad + tile * 128 + cg * 32);
#pragma unroll
                    for (int i = 0; i < 8; ++i) dst[i] = make_float4(part[4 * i], part[4 * i + 1], part[4 * i + 2], part[4 * i + 3]);
                }
            }
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, 256);
    }
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
    }
    return FDB_OK;
}

int fdb_kshap_launch_eval_fx(fdb_kshap_plan* pl, cudaStream_t st, const float* x, int n_rows, float* ey) {
    FDB_REQUIRE(pl->d_E != nullptr, FDB_ERR_STATE,
                "the factored evaluator is not available for this plan (needs one coalition set whose background logits "
                "stay within +-100 log2 units per coalition)");
    FxParams p;
    p.x = x; p.zimg = pl->d_aimg; p.E = pl->d_E; p.eoff = pl->d_eoff; p.wq = pl->d_wq; p.ey = ey;
    p.xw = pl->d_xw; p.gpart = pl->d_gpart;
    p.n_rows = n_rows; p.tiles = pl->m_pad / 128; p.m_pad = pl->m_pad; p.n_bg = pl->n_bg; p.fmt = pl->fmt;
    p.n_chunks = (pl->n_bg + 127) / 128;
    p.link = pl->link;
    p.inv_bg = 1.0f / (float)pl->n_bg;
    p.lnull = pl->link == 1 ? (float)log(pl->fnull / (1.0 - pl->fnull)) : (float)pl->fnull;
    const size_t smem = 2 * kEStageBytes + 2 * kTileBytes + 2 * kQBytes + sizeof(FxBars) + 64;
    static bool configured[FDB_MAX_DEVICES] = {};          // function attributes are per device
    if (!configured[pl->ctx->device]) {
        FDB_CUDA(cudaFuncSetAttribute(kshap_eval_fx_kernel<kFxClampedQuads>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        FDB_CUDA(cudaFuncSetAttribute(kshap_eval_fx_kernel<kFxClampedPairs>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        FDB_CUDA(cudaFuncSetAttribute(kshap_eval_fx_kernel<kFxBalanced>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        FDB_CUDA(cudaFuncSetAttribute(kshap_eval_fx_kernel<kFxBalancedLogit>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[pl->ctx->device] = true;
    }
    const int64_t n_units = (int64_t)((n_rows + 127) / 128) * p.tiles;
    int64_t grid = n_units < pl->ctx->sm_count ? n_units : pl->ctx->sm_count;
    if (grid < 1) grid = 1;
    // balanced (clamp-free) inner loop whenever the product of the four denominators of every column quad is
    // guaranteed to stay within 2^+-114 (kshap_build_E)
    const char* force = getenv("FDB_FX_CLAMPED");
    const bool balanced = pl->fx_quad_logsum <= 110.f && pl->fx_half_spread <= 100.f && !(force && atoi(force) != 0);
    if (balanced && pl->link == 1) kshap_eval_fx_kernel<kFxBalancedLogit><<<(int)grid, kFxThreads, smem, st>>>(p);
    else if (balanced) kshap_eval_fx_kernel<kFxBalanced><<<(int)grid, kFxThreads, smem, st>>>(p);
    else if (pl->link == 1) kshap_eval_fx_kernel<kFxClampedPairs><<<(int)grid, kFxThreads, smem, st>>>(p);
    else kshap_eval_fx_kernel<kFxClampedQuads><<<(int)grid, kFxThreads, smem, st>>>(p);
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}
