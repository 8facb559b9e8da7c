// K2 - Kernel SHAP coalition evaluator on the 5th-gen tensor cores (tcgen05 + TMEM), sm_100a.
//
//   ey[r][m] = mean_b sigmoid( sum_k (1-z_mk) w_k bg_bk  +  sum_k z_mk w_k x_rk  +  c )
//
// is computed per (explained row r, tile of 128 coalitions) as one small GEMM whose result
// never leaves the SM:
//   A  (128 x 32, smem)  = [1 - z_m0 .. 1 - z_m29 | 1 | 0]           exactly 0/1 in f16/bf16
//   B  (n_bg x 32, smem) = -log2e * [w_k bg_bk | c | 0]             split into 2 f16 (or 3 bf16) terms
//   Bq (8 x 32, smem)    = 3-term split of [log2e w_k x_rk | -log2e sum_k w_k x_rk | 0], the SAME row 8 times;
//                          its descriptor uses a row-group stride of 0, so the MMA sees that row broadcast
//                          to every background column
//   D  (TMEM, fp32)      = A B^T + A Bq^T   (all split terms accumulate into the same columns)
// so that D[m][b] = -log2e * logit[m][b] leaves the tensor core finished.  The epilogue reads TMEM with
// tcgen05.ld (one coalition per thread = one TMEM lane), evaluates sigmoid = 1 / (1 + ex2(D)) on the SFU -
// four sigmoids share one reciprocal - and keeps the running mean over b in a register; only ey[r][m]
// (4 bytes per 128 x n_bg logits) goes to HBM.  Because A is exactly representable and the
// B splits carry >= 22 mantissa bits, the products are fp32-accurate although the MMA runs at
// f16/bf16 rate.
//
// CTA = 24 warps, persistent, four independent pipelines g = 0..3 (each owns 128 TMEM columns), so every
// SM sub-partition always has four epilogue warps to hide SFU and TMEM latency:
//   warp g           producer : writes Bq for the item's row and expands the item's 128 coalition words (512 B from
//                               global memory) into the 8 KB A tile in shared memory - no 64-byte-per-coalition image in
//                               HBM; single-set plans keep ready-made tile images (K2f streams them) and bulk-copy them (TMA)
//   warp 4+g         MMA      : one lane issues tcgen05.mma / tcgen05.commit
//   warps 8+4g..11+4g epilogue: TMEM lane quarter (warp % 4), 32 coalitions per warp
// The B operand stays resident in shared memory for the whole kernel.  The background is walked in chunks of 64
// columns that alternate between the two halves of the pipeline's TMEM columns, so the MMA of chunk c + 1
// (or of the next item) runs while the epilogue reads chunk c; quads of the last group that lie entirely
// beyond the background are skipped.

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "kshap_internal.cuh"
#include "ptx_sm100.cuh"

namespace {

#ifndef FDB_K2_PIPES
#define FDB_K2_PIPES 4
#endif
#ifndef FDB_K2_BUF_COLS
#define FDB_K2_BUF_COLS 64
#endif
constexpr int kPipes = FDB_K2_PIPES;
constexpr int kThreadsTc = (2 * kPipes + 4 * kPipes) * 32;   // 768
constexpr int kATileBytes = 128 * 32 * 2;     // 8192
constexpr int kBqBytes = 3 * 512;             // 3 splits x (4 k-chunks x 8 identical rows x 16 B)
constexpr int kAStages = 2;
constexpr int kBufCols = FDB_K2_BUF_COLS;     // TMEM columns of one accumulator buffer (a background chunk)
constexpr int kPipeCols = 2 * kBufCols;       // per pipeline: two buffers, the MMA of background chunk c + 1 runs while chunk c is read
static_assert(kPipes * kPipeCols <= 512 && kBufCols % 16 == 0, "TMEM holds 512 columns");
constexpr uint32_t kLbo = 128;                // K-direction core-matrix stride (bytes)
constexpr uint32_t kSbo = 512;                // 8-row-group stride (bytes)

struct TcParams {
    const float* x;
    const int32_t* set_idx;
    const uint32_t* bits;                      // [S][m_pad] coalition words; the A tile is expanded from them in smem ...
    const uint16_t* aimg;                      // ... unless the plan holds ready-made tile images (single-set plans): TMA
    const uint16_t* bimg;
    const float* wq;
    float* ey;
    int64_t row_offset;
    int n_rows, n_sets, tiles, m_pad;
    int n_bg, nb_pad, nsplit, fmt;
    int n_chunks, chunk_cols;
    float inv_bg;
};

struct PipeBars {
    uint64_t a_full[kAStages];
    uint64_t a_empty[kAStages];
    uint64_t t_full[2];
    uint64_t t_empty[2];
};

struct TcBars {
    PipeBars pipe[kPipes];
    uint64_t b_full;
    uint32_t tmem_base;
};

__device__ __forceinline__ uint16_t to_h16(float v, int fmt) {
    return fmt == 0 ? __half_as_ushort(__float2half_rn(v)) : __bfloat16_as_ushort(__float2bfloat16_rn(v));
}
__device__ __forceinline__ float from_h16(uint16_t h, int fmt) {
    return fmt == 0 ? __half2float(__ushort_as_half(h)) : __bfloat162float(__ushort_as_bfloat16(h));
}


// 2^x on the FMA pipe (Cody-Waite split + degree-5 minimax, max rel. error 2.3e-7 ~ ex2.approx), x in [-120, 31].
// Used for a fixed share of the exponentials so the SFU (16 lanes/SM) and the FMA pipe (128 lanes/SM) finish together.
__device__ __forceinline__ float ex2_poly(float x) {
    const float magic = 12582912.f;                 // 1.5 * 2^23: rounds x to an integer held in the low mantissa bits
    const float r = x + magic;
    const float f = x - (r - magic);                // [-0.5, 0.5]
    float p = 0.001327647129073739f;
    p = fmaf(p, f, 0.009675540961325169f);
    p = fmaf(p, f, 0.05550713092088699f);
    p = fmaf(p, f, 0.24022120237350464f);
    p = fmaf(p, f, 0.6931469440460205f);
    p = fmaf(p, f, 1.0000001192092896f);
    return __int_as_float(__float_as_int(p) + (__float_as_int(r) << 23));
}

__device__ __forceinline__ void release_tmem(uint64_t* bar, int lane) {
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(bar);
}

// Share of the exponentials evaluated by ex2_poly instead of the SFU: bit i of the mask = element i of every
// 8.  Measured on B200: with four epilogue warps per sub-partition the kernel is SFU-bound, and every
// polynomial costs ~9 issue slots, so only a small share pays.
#ifndef FDB_K2_POLY_MASK
#define FDB_K2_POLY_MASK 0x00
#endif
constexpr uint32_t kPolyMask8 = FDB_K2_POLY_MASK;

#ifndef FDB_K2_SERVICE_SLEEP_NS
#define FDB_K2_SERVICE_SLEEP_NS 48
#endif
// producer / MMA lanes poll with a short back-off so they do not take issue slots from the epilogue warps
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
    while (!ptx::mbar_try_wait(bar, parity)) __nanosleep(FDB_K2_SERVICE_SLEEP_NS);
}

// sum_i sigmoid over 16 accumulator columns, t_i = D[m][b_i] = -log2e * logit.
// Four sigmoids share ONE reciprocal: s0+s1+s2+s3 = (n01 p23 + n23 p01) / (p01 p23) with d_i = 1 + 2^t_i,
// n_ab = d_a + d_b, p_ab = d_a d_b.  t is clamped to 30 so the product of four d stays below 2^121;
// the clamp changes a sigmoid by < 1e-9 absolute.  Masked columns (beyond the background) are pinned at the
// clamp, where the sigmoid is 1 / (1 + 2^30), and that known amount is subtracted again (no cancellation).
// (Identity link; the logit link uses sigmoid_and_complement_sums16 below.)
template <bool MASKED>
__device__ __forceinline__ float sigmoid_sum16(const uint32_t (&r)[16], int nvalid) {
    constexpr float kClamp = 30.f;
    constexpr float kFloor = 9.31322574e-10f;   // sigmoid at the clamp, 1 / (1 + 2^clamp)
    float sum = 0.f;
#pragma unroll
    for (int q = 0; q < 16; q += 4) {
        if (MASKED && q >= nvalid) break;              // quads that lie entirely beyond the background cost nothing
        float d[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float t = fminf(__uint_as_float(r[q + i]), kClamp);
            if (MASKED && q + i >= nvalid) t = kClamp;
            const float e = ((kPolyMask8 >> ((q + i) & 7)) & 1u) ? ex2_poly(fmaxf(t, -120.f)) : ptx::ex2_approx(t);
            d[i] = 1.0f + e;
        }
        const float n01 = d[0] + d[1], p01 = d[0] * d[1];
        const float n23 = d[2] + d[3], p23 = d[2] * d[3];
        const float num = fmaf(n01, p23, n23 * p01);
        sum = fmaf(num, ptx::rcp_approx(p01 * p23), sum);
    }
    return MASKED ? sum - (float)(((nvalid + 3) & ~3) - nvalid) * kFloor : sum;   // masked columns of the last quad sit at the clamp; take them out again
}

__device__ __forceinline__ float sigmoid_group(const uint32_t (&r)[16], int nvalid) {
    return nvalid >= 16 ? sigmoid_sum16<false>(r, 16) : sigmoid_sum16<true>(r, nvalid);
}

// Logit link: sum of the sigmoids AND of their complements over up to 16 accumulator columns, two columns per
// reciprocal:  s0 + s1 = (d0 + d1) / (d0 d1),  (1 - s0) + (1 - s1) = (e0 d1 + e1 d0) / (d0 d1),  e = 2^t, d = 1 + e.
// t is clamped at 62 (e0 d1 + e1 d0 < 2^126); a sigmoid keeps its relative accuracy down to 2^-62 and a
// complement down to 2^-126, so the stored expectation is exact to float32 in both tails (fdb_kshap_encode_ey).
// Columns beyond the background are simply left out (an odd last column takes a reciprocal of its own).
__device__ __forceinline__ void sigmoid_and_complement_sums16(const uint32_t (&r)[16], int nvalid, float& S, float& C) {
#pragma unroll
    for (int q = 0; q < 16; q += 2) {
        if (q >= nvalid) break;
        const float e0 = ptx::ex2_approx(fminf(__uint_as_float(r[q]), 62.f));
        const float e1 = ptx::ex2_approx(fminf(__uint_as_float(r[q + 1]), 62.f));
        const float d0 = 1.0f + e0, d1 = 1.0f + e1;
        if (q + 1 < nvalid) {
            const float rp = ptx::rcp_approx(d0 * d1);
            S = fmaf(d0 + d1, rp, S);
            C = fmaf(fmaf(e0, d1, e1 * d0), rp, C);
        } else {
            const float r0 = ptx::rcp_approx(d0);
            S += r0;
            C = fmaf(e0, r0, C);
        }
    }
}

// smem carve-up (dynamic): [B image][A stages x pipes][Bq stages x pipes][barriers]
// FDB_K2_MAXREG: cap on registers per thread.  At 64 the CTA leaves a quarter of the register file (and 130 KB of shared
// memory) free, so the coalition sampler of the NEXT batch (other stream, 51 registers, 16 KB per set) can run beside it.
#ifndef FDB_K2_MAXREG
#define FDB_K2_MAXREG 0
#endif
#if FDB_K2_MAXREG > 0
#define FDB_K2_REGCAP __maxnreg__(FDB_K2_MAXREG)
#else
#define FDB_K2_REGCAP __launch_bounds__(kThreadsTc, 1)
#endif
template <bool FROM_BITS, bool WIDE>
__global__ void FDB_K2_REGCAP kshap_eval_tc_kernel(const TcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const uint32_t b_bytes = (uint32_t)p.nsplit * p.nb_pad * 64;
    unsigned char* sB = smem;
    unsigned char* sA = sB + b_bytes;                                       // [pipes][kAStages][8192]
    unsigned char* sBq = sA + kPipes * kAStages * kATileBytes;              // [pipes][kAStages][1536]
    TcBars* bars = reinterpret_cast<TcBars*>(sBq + kPipes * kAStages * kBqBytes);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t n_items = (int64_t)p.n_rows * p.tiles;
    const int64_t stride = (int64_t)gridDim.x * kPipes;

    // ---- one-time set-up -----------------------------------------------------------------
    if (threadIdx.x == 0) {
        for (int g = 0; g < kPipes; ++g) {
            for (int s = 0; s < kAStages; ++s) {
                ptx::mbar_init(&bars->pipe[g].a_full[s], 1);
                ptx::mbar_init(&bars->pipe[g].a_empty[s], 1);
            }
            for (int b = 0; b < 2; ++b) {
                ptx::mbar_init(&bars->pipe[g].t_full[b], 1);
                ptx::mbar_init(&bars->pipe[g].t_empty[b], 4);
            }
        }
        ptx::mbar_init(&bars->b_full, 1);
        ptx::mbar_fence_init();
    }
    if (warp == kPipes) ptx::tmem_alloc(&bars->tmem_base, 512);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == 0 && lane == 0) {
        // resident B operand: one barrier phase, copies of <= 32 KB
        ptx::mbar_arrive_expect_tx(&bars->b_full, b_bytes);
        for (uint32_t off = 0; off < b_bytes; off += 32768) {
            const uint32_t n = (b_bytes - off) < 32768u ? (b_bytes - off) : 32768u;
            ptx::bulk_g2s(sB + off, reinterpret_cast<const unsigned char*>(p.bimg) + off, n, &bars->b_full);
        }
    }

    if (warp < kPipes) {
        // =================================================================== producer of pipeline g
        const int g = warp;
        PipeBars& pb = bars->pipe[g];
        const float wk = lane < FDB_F ? p.wq[lane] : 0.f;
        int k = 0;
        for (int64_t it = (int64_t)blockIdx.x * kPipes + g; it < n_items; it += stride, ++k) {
            const int s = k % kAStages;
            const uint32_t ph = (k / kAStages) & 1;
            const int row = (int)(it / p.tiles);
            const int tile = (int)(it % p.tiles);
            const int set = p.set_idx ? p.set_idx[row] : (p.n_sets == 1 ? 0 : (int)((p.row_offset + row) % p.n_sets));
            // Bq = 3-term split of [log2e w_k x_k (k<30) | -log2e sum_k w_k x_k | 0], each term stored as 8 equal rows
            float v = lane < FDB_F ? 1.4426950408889634f * wk * p.x[(size_t)row * FDB_F + lane] : 0.f;
            v = fminf(fmaxf(v, -30000.f), 30000.f);
            float tot = v;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            if (lane == 30) v = fminf(fmaxf(-tot, -60000.f), 60000.f);
            uint16_t h[3];
            float rem = v;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                h[j] = to_h16(rem, p.fmt);
                rem -= from_h16(h[j], p.fmt);
            }
            // the item's 128 coalition words (lane l: rows l, l + 32, l + 64, l + 96), fetched before the stage is waited for
            const uint32_t* mb = p.bits + ((size_t)set * p.m_pad + (size_t)tile * 128);
            uint32_t mw[4];
            if (FROM_BITS) {
#pragma unroll
                for (int j = 0; j < 4; ++j) mw[j] = __ldg(mb + j * 32 + lane);
            }
            mbar_wait_relaxed(&pb.a_empty[s], ph ^ 1);
            unsigned char* bq = sBq + (size_t)(g * kAStages + s) * kBqBytes + (lane >> 3) * 128 + (lane & 7) * 2;
#pragma unroll
            for (int j = 0; j < 3; ++j)
#pragma unroll
                for (int rr = 0; rr < 8; ++rr) *reinterpret_cast<uint16_t*>(bq + j * 512 + rr * 16) = h[j];
            // A tile (128 coalitions x 32 k, canonical K-major no-swizzle layout) straight from the coalition words:
            //   a[m][k] = 1 - z_mk (k < 30), a[m][30] = 1 (bias lane), a[m][31] = 0.  16 B per (row, k-chunk of 8);
            //   a quarter warp writes 128 contiguous bytes -> conflict-free STS.128.  64 B of operand per 4 B read.
            if (!FROM_BITS) {                 // ready-made tile image: one bulk copy (TMA, UBLKCP)
                ptx::fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    ptx::mbar_arrive_expect_tx(&pb.a_full[s], kATileBytes);
                    ptx::bulk_g2s(sA + (size_t)(g * kAStages + s) * kATileBytes,
                                  reinterpret_cast<const unsigned char*>(p.aimg) + ((size_t)set * p.tiles + tile) * kATileBytes,
                                  kATileBytes, &pb.a_full[s]);
                }
                continue;
            }
            unsigned char* at = sA + (size_t)(g * kAStages + s) * kATileBytes + (lane >> 3) * 512 + (lane & 7) * 16;
            const uint32_t one = p.fmt == 0 ? 0x3C00u : 0x3F80u;      // 1.0 in f16 / bf16
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t abar = (~mw[j] & 0x3FFFFFFFu) | (1u << 30);
#pragma unroll
                for (int kc = 0; kc < 4; ++kc) {
                    uint32_t w4[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint32_t two = (abar >> (kc * 8 + q * 2)) & 3u;          // two adjacent k
                        w4[q] = ((two & 1u) | ((two & 2u) << 15)) * one;
                    }
                    *reinterpret_cast<uint4*>(at + j * 2048 + kc * 128) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
                }
            }
            ptx::fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&pb.a_full[s]);
        }
    } else if (warp < 2 * kPipes) {
        // =================================================================== MMA issuer of pipeline g
        const int g = warp - kPipes;
        PipeBars& pb = bars->pipe[g];
        if (lane == 0) {
            ptx::mbar_wait(&bars->b_full, 0);
            const uint32_t a0 = ptx::smem_u32(sA + (size_t)g * kAStages * kATileBytes);
            const uint32_t q0 = ptx::smem_u32(sBq + (size_t)g * kAStages * kBqBytes);
            const uint32_t b0 = ptx::smem_u32(sB);
            const uint32_t d0 = tmem_base + g * kPipeCols;
            int k = 0;
            uint32_t job = 0;
            for (int64_t it = (int64_t)blockIdx.x * kPipes + g; it < n_items; it += stride, ++k) {
                const int s = k % kAStages;
                const uint32_t ph = (k / kAStages) & 1;
                mbar_wait_relaxed(&pb.a_full[s], ph);
                ptx::tc_fence_after();
                const uint32_t a_addr = a0 + s * kATileBytes;
                const uint32_t q_addr = q0 + s * kBqBytes;
                for (int c = 0; c < p.n_chunks; ++c, ++job) {
                    const uint32_t buf = job & 1, d = d0 + buf * kBufCols;
                    mbar_wait_relaxed(&pb.t_empty[buf], ((job >> 1) & 1) ^ 1);
                    ptx::tc_fence_after();
                    const int col0 = c * p.chunk_cols;
                    const int nc = (p.nb_pad - col0) < p.chunk_cols ? (p.nb_pad - col0) : p.chunk_cols;
                    const uint32_t idesc = ptx::make_idesc_f16(p.fmt, 128, nc);
                    bool acc = false;
                    for (int sp = 0; sp < p.nsplit; ++sp) {
#pragma unroll
                        for (int ks = 0; ks < 2; ++ks) {
                            const uint64_t ad = ptx::make_desc_kmajor_noswizzle(a_addr + ks * 256, kLbo, kSbo);
                            const uint64_t bd = ptx::make_desc_kmajor_noswizzle(
                                b0 + (uint32_t)sp * p.nb_pad * 64 + (uint32_t)col0 * 64 + ks * 256, kLbo, kSbo);
                            ptx::mma_f16(d, ad, bd, idesc, acc);
                            acc = true;
                        }
                    }
#pragma unroll
                    for (int sp = 0; sp < 3; ++sp) {
#pragma unroll
                        for (int ks = 0; ks < 2; ++ks) {
                            const uint64_t ad = ptx::make_desc_kmajor_noswizzle(a_addr + ks * 256, kLbo, kSbo);
                            // row-group stride 0: the 8 stored rows are reused for every group of 8 columns
                            const uint64_t qd = ptx::make_desc_kmajor_noswizzle(q_addr + sp * 512 + ks * 256, kLbo, 0);
                            ptx::mma_f16(d, ad, qd, idesc, true);
                        }
                    }
                    ptx::mma_commit(&pb.t_full[buf]);
                }
                ptx::mma_commit(&pb.a_empty[s]);
            }
        }
        __syncwarp();
    } else {
        // =================================================================== epilogue of pipeline g
        const int g = (warp - 2 * kPipes) >> 2;
        const int qd = warp & 3;                     // TMEM lane quarter this warp may read
        PipeBars& pb = bars->pipe[g];
        const uint32_t ta0 = tmem_base + ((uint32_t)(qd * 32) << 16) + g * kPipeCols;
        uint32_t job = 0;
        for (int64_t it = (int64_t)blockIdx.x * kPipes + g; it < n_items; it += stride) {
            const int row = (int)(it / p.tiles);
            const int tile = (int)(it % p.tiles);
            float acc = 0.f, comp = 0.f;
            for (int c = 0; c < p.n_chunks; ++c, ++job) {
                const uint32_t buf = job & 1, ta = ta0 + buf * kBufCols;
                ptx::mbar_wait(&pb.t_full[buf], (job >> 1) & 1);
                ptx::tc_fence_after();
                int valid = p.n_bg - c * p.chunk_cols;
                if (valid > p.chunk_cols) valid = p.chunk_cols;
                // double-buffered TMEM reads: the load of group g+1 is in flight while group g is evaluated; the
                // accumulator columns are handed back to the MMA warp as soon as the last one has been read
                const int ngroups = (valid + 15) >> 4;
                uint32_t ra[16], rb[16];
                ptx::tmem_ld16(ta, ra);
                for (int gi = 0; gi < ngroups; gi += 2) {
                    ptx::tmem_ld_wait();
                    if (gi + 1 < ngroups) ptx::tmem_ld16(ta + (gi + 1) * 16, rb);
                    else release_tmem(&pb.t_empty[buf], lane);
                    if (WIDE) sigmoid_and_complement_sums16(ra, valid - gi * 16, acc, comp);
                    else acc += sigmoid_group(ra, valid - gi * 16);
                    if (gi + 1 < ngroups) {
                        ptx::tmem_ld_wait();
                        if (gi + 2 < ngroups) ptx::tmem_ld16(ta + (gi + 2) * 16, ra);
                        else release_tmem(&pb.t_empty[buf], lane);
                        if (WIDE) sigmoid_and_complement_sums16(rb, valid - (gi + 1) * 16, acc, comp);
                        else acc += sigmoid_group(rb, valid - (gi + 1) * 16);
                    }
                }
            }
            p.ey[(size_t)row * p.m_pad + tile * 128 + qd * 32 + lane] = WIDE ? fdb_kshap_encode_ey(acc, comp, p.inv_bg) : acc * p.inv_bg;
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == kPipes) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace

// shared memory one CTA of the tensor-core evaluator needs for this plan (resident B image + A / Bq stages + barriers)
size_t fdb_kshap_tc_smem_bytes(const fdb_kshap_plan* pl) {
    return (size_t)pl->nsplit * pl->nb_pad * 64 + kPipes * kAStages * (kATileBytes + kBqBytes) + sizeof(TcBars) + 64;
}

int fdb_kshap_launch_eval_tc(fdb_kshap_plan* pl, cudaStream_t st, const float* x, int64_t row_offset, int n_rows,
                             const int32_t* set_idx, float* ey) {
    TcParams p;
    p.x = x; p.set_idx = set_idx; p.bits = pl->d_bits; p.aimg = pl->d_aimg; p.bimg = pl->d_bimg; p.wq = pl->d_wq; p.ey = ey;
    p.row_offset = row_offset; p.n_rows = n_rows; p.n_sets = pl->n_sets; p.tiles = pl->m_pad / 128; p.m_pad = pl->m_pad;
    p.n_bg = pl->n_bg; p.nb_pad = pl->nb_pad; p.nsplit = pl->nsplit; p.fmt = pl->fmt;
    p.n_chunks = (pl->nb_pad + kBufCols - 1) / kBufCols;
    p.chunk_cols = kBufCols;
    p.inv_bg = 1.0f / (float)pl->n_bg;
    const size_t smem = fdb_kshap_tc_smem_bytes(pl);
    FDB_REQUIRE(smem <= 227 * 1024, FDB_ERR_ARG, "background of %d rows with %d operand splits needs %zu bytes of shared memory",
                pl->n_bg, pl->nsplit, smem);
    static size_t configured[FDB_MAX_DEVICES] = {};        // function attributes are per device
    if (smem > configured[pl->ctx->device]) {
        FDB_CUDA(cudaFuncSetAttribute(kshap_eval_tc_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        FDB_CUDA(cudaFuncSetAttribute(kshap_eval_tc_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        FDB_CUDA(cudaFuncSetAttribute(kshap_eval_tc_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        FDB_CUDA(cudaFuncSetAttribute(kshap_eval_tc_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[pl->ctx->device] = smem;
    }
    const int64_t n_items = (int64_t)n_rows * p.tiles;
    int64_t grid = (n_items + kPipes - 1) / kPipes;
    if (grid > pl->ctx->sm_count) grid = pl->ctx->sm_count;
    if (grid < 1) grid = 1;
    const bool from_bits = pl->d_aimg == nullptr, wide = pl->link == 1;
    if (!from_bits && !wide) kshap_eval_tc_kernel<false, false><<<(int)grid, kThreadsTc, smem, st>>>(p);
    else if (from_bits && !wide) kshap_eval_tc_kernel<true, false><<<(int)grid, kThreadsTc, smem, st>>>(p);
    else if (!from_bits) kshap_eval_tc_kernel<false, true><<<(int)grid, kThreadsTc, smem, st>>>(p);
    else kshap_eval_tc_kernel<true, true><<<(int)grid, kThreadsTc, smem, st>>>(p);
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}
