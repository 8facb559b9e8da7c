// K2 - Kernel SHAP coalition evaluator on the 5th-gen tensor cores (tcgen05 + TMEM), sm_100a.
//
//   ey[r][m] = mean_b sigmoid( sum_k (1-z_mk) w_k bg_bk  +  sum_k z_mk w_k x_rk  +  c )
//
// is computed per (explained row r, tile of 128 coalitions) as one small GEMM whose result
// never leaves the SM:
//   A  (128 x 32, smem)  = [1 - z_m0 .. 1 - z_m29 | 1 | 0]           exactly 0/1 in f16/bf16
//   B  (n_bg x 32, smem) = -log2e * [w_k bg_bk | c | 0]             split into 2 f16 (or 3 bf16) terms
//   Bq (16 x 32, smem)   = rows 0..2: the 3-term split of [log2e w_k x_rk | -log2e sum_k w_k x_rk | 0]
//   D  (TMEM, fp32)      = A B^T  (all split terms accumulate into the same columns) and A Bq^T
// so that t[m][b] = D[m][b] + (Dq[m][0] + Dq[m][1] + Dq[m][2]) = -log2e * logit[m][b].  The epilogue
// reads TMEM with tcgen05.ld (one coalition per thread = one TMEM lane), evaluates
// sigmoid = rcp(1 + ex2(t)) on the SFU and keeps the running mean over b in a register; only
// ey[r][m] (4 bytes per 128 x n_bg logits) goes to HBM.  Because A is exactly representable and the
// B splits carry >= 22 mantissa bits, the products are fp32-accurate although the MMA runs at
// f16/bf16 rate.
//
// CTA = 12 warps, persistent, two independent pipelines g = 0,1 (each owns half of TMEM):
//   warp g      producer : writes Bq for the item's row, bulk-copies (TMA, UBLKCP) the A tile image
//   warp 2+g    MMA      : one lane issues tcgen05.mma / tcgen05.commit
//   warps 4+4g..7+4g  epilogue: TMEM lane quarter (warp % 4), 32 coalitions per warp
// The B operand stays resident in shared memory for the whole kernel.  Backgrounds wider than
// 240 rows are processed in column chunks that reuse the pipeline's TMEM buffer.

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "kshap_internal.cuh"
#include "ptx_sm100.cuh"

namespace {

constexpr int kThreadsTc = 384;
constexpr int kATileBytes = 128 * 32 * 2;     // 8192
constexpr int kBqBytes = 16 * 32 * 2;         // 1024
constexpr int kAStages = 2;
constexpr int kPipeCols = 256;                // TMEM columns per pipeline
constexpr uint32_t kLbo = 128;                // K-direction core-matrix stride (bytes)
constexpr uint32_t kSbo = 512;                // 8-row-group stride (bytes)

struct TcParams {
    const float* x;
    const int32_t* set_idx;
    const uint16_t* aimg;
    const uint16_t* bimg;
    const float* wq;
    float* ey;
    int64_t row_offset;
    int n_rows, n_sets, tiles, m_pad;
    int n_bg, nb_pad, nsplit, fmt;
    int n_chunks, chunk_cols, buf_cols, nbuf;
    float inv_bg;
};

struct PipeBars {
    uint64_t a_full[kAStages];
    uint64_t a_empty[kAStages];
    uint64_t t_full[2];
    uint64_t t_empty[2];
};

struct TcBars {
    PipeBars pipe[2];
    uint64_t b_full;
    uint32_t tmem_base;
};

__device__ __forceinline__ uint16_t to_h16(float v, int fmt) {
    return fmt == 0 ? __half_as_ushort(__float2half_rn(v)) : __bfloat16_as_ushort(__float2bfloat16_rn(v));
}
__device__ __forceinline__ float from_h16(uint16_t h, int fmt) {
    return fmt == 0 ? __half2float(__ushort_as_half(h)) : __bfloat162float(__ushort_as_bfloat16(h));
}


// smem carve-up (dynamic): [B image][A stages pipe0][A stages pipe1][Bq stages pipe0][Bq stages pipe1][barriers]
__global__ void __launch_bounds__(kThreadsTc, 1) kshap_eval_tc_kernel(const TcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const uint32_t b_bytes = (uint32_t)p.nsplit * p.nb_pad * 64;
    unsigned char* sB = smem;
    unsigned char* sA = sB + b_bytes;                                   // [2 pipes][kAStages][8192]
    unsigned char* sBq = sA + 2 * kAStages * kATileBytes;               // [2 pipes][kAStages][1024]
    TcBars* bars = reinterpret_cast<TcBars*>(sBq + 2 * kAStages * kBqBytes);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t n_items = (int64_t)p.n_rows * p.tiles;
    const int64_t stride = (int64_t)gridDim.x * 2;

    // ---- one-time set-up -----------------------------------------------------------------
    for (uint32_t i = threadIdx.x; i < 2 * kAStages * kBqBytes / 16; i += blockDim.x)
        reinterpret_cast<uint4*>(sBq)[i] = make_uint4(0, 0, 0, 0);     // rows 3..15 of every Bq stay zero
    if (threadIdx.x == 0) {
        for (int g = 0; g < 2; ++g) {
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
    if (warp == 2) ptx::tmem_alloc(&bars->tmem_base, 512);
    ptx::fence_proxy_async_smem();
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

    if (warp < 2) {
        // =================================================================== producer of pipeline g
        const int g = warp;
        PipeBars& pb = bars->pipe[g];
        const float wk = lane < FDB_F ? p.wq[lane] : 0.f;
        int k = 0;
        for (int64_t it = (int64_t)blockIdx.x * 2 + g; it < n_items; it += stride, ++k) {
            const int s = k % kAStages;
            const uint32_t ph = (k / kAStages) & 1;
            const int row = (int)(it / p.tiles);
            const int tile = (int)(it % p.tiles);
            const int set = p.set_idx ? p.set_idx[row] : (p.n_sets == 1 ? 0 : (int)((p.row_offset + row) % p.n_sets));
            ptx::mbar_wait(&pb.a_empty[s], ph ^ 1);
            // Bq rows 0..2 = 3-term split of [log2e w_k x_k (k<30) | -log2e sum_k w_k x_k | 0]
            float v = lane < FDB_F ? 1.4426950408889634f * wk * p.x[(size_t)row * FDB_F + lane] : 0.f;
            v = fminf(fmaxf(v, -30000.f), 30000.f);
            float tot = v;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            if (lane == 30) v = fminf(fmaxf(-tot, -60000.f), 60000.f);
            unsigned char* bq = sBq + (size_t)(g * kAStages + s) * kBqBytes + (lane >> 3) * 128 + (lane & 7) * 2;
            float rem = v;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const uint16_t h = to_h16(rem, p.fmt);
                *reinterpret_cast<uint16_t*>(bq + j * 16) = h;
                rem -= from_h16(h, p.fmt);
            }
            ptx::fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                ptx::mbar_arrive_expect_tx(&pb.a_full[s], kATileBytes);
                ptx::bulk_g2s(sA + (size_t)(g * kAStages + s) * kATileBytes,
                              reinterpret_cast<const unsigned char*>(p.aimg) + ((size_t)set * p.tiles + tile) * kATileBytes,
                              kATileBytes, &pb.a_full[s]);
            }
        }
    } else if (warp < 4) {
        // =================================================================== MMA issuer of pipeline g
        const int g = warp - 2;
        PipeBars& pb = bars->pipe[g];
        if (lane == 0) {
            ptx::mbar_wait(&bars->b_full, 0);
            const uint32_t a0 = ptx::smem_u32(sA + (size_t)g * kAStages * kATileBytes);
            const uint32_t q0 = ptx::smem_u32(sBq + (size_t)g * kAStages * kBqBytes);
            const uint32_t b0 = ptx::smem_u32(sB);
            const uint32_t idesc_q = ptx::make_idesc_f16(p.fmt, 128, 16);
            int k = 0;
            int64_t job = 0;
            for (int64_t it = (int64_t)blockIdx.x * 2 + g; it < n_items; it += stride, ++k) {
                const int s = k % kAStages;
                const uint32_t ph = (k / kAStages) & 1;
                ptx::mbar_wait(&pb.a_full[s], ph);
                ptx::tc_fence_after();
                const uint32_t a_addr = a0 + s * kATileBytes;
                for (int c = 0; c < p.n_chunks; ++c, ++job) {
                    const int buf = (int)(job % p.nbuf);
                    const uint32_t tph = (uint32_t)((job / p.nbuf) & 1);
                    ptx::mbar_wait(&pb.t_empty[buf], tph ^ 1);
                    ptx::tc_fence_after();
                    const uint32_t d = tmem_base + g * kPipeCols + buf * p.buf_cols;
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
                    if (c == 0) {
#pragma unroll
                        for (int ks = 0; ks < 2; ++ks) {
                            const uint64_t ad = ptx::make_desc_kmajor_noswizzle(a_addr + ks * 256, kLbo, kSbo);
                            const uint64_t qd = ptx::make_desc_kmajor_noswizzle(q0 + s * kBqBytes + ks * 256, kLbo, kSbo);
                            ptx::mma_f16(d + (p.buf_cols - 16), ad, qd, idesc_q, ks > 0);
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
        const int g = (warp - 4) >> 2;
        const int qd = warp & 3;                     // TMEM lane quarter this warp may read
        PipeBars& pb = bars->pipe[g];
        const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16) + g * kPipeCols;
        int64_t job = 0;
        for (int64_t it = (int64_t)blockIdx.x * 2 + g; it < n_items; it += stride) {
            const int row = (int)(it / p.tiles);
            const int tile = (int)(it % p.tiles);
            float acc = 0.f, sq = 0.f;
            for (int c = 0; c < p.n_chunks; ++c, ++job) {
                const int buf = (int)(job % p.nbuf);
                const uint32_t tph = (uint32_t)((job / p.nbuf) & 1);
                ptx::mbar_wait(&pb.t_full[buf], tph);
                ptx::tc_fence_after();
                const uint32_t ta = lane_base + buf * p.buf_cols;
                if (c == 0) {
                    uint32_t r4[4];
                    ptx::tmem_ld4(ta + (p.buf_cols - 16), r4);
                    ptx::tmem_ld_wait();
                    sq = __uint_as_float(r4[0]) + __uint_as_float(r4[1]) + __uint_as_float(r4[2]);
                }
                const int col0 = c * p.chunk_cols;
                int valid = p.n_bg - col0;
                if (valid > p.chunk_cols) valid = p.chunk_cols;
                for (int cc = 0; cc < valid; cc += 16) {
                    uint32_t r[16];
                    ptx::tmem_ld16(ta + cc, r);
                    ptx::tmem_ld_wait();
                    if (cc + 16 <= valid) {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            acc += ptx::rcp_approx(1.0f + ptx::ex2_approx(__uint_as_float(r[i]) + sq));
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (cc + i < valid) acc += ptx::rcp_approx(1.0f + ptx::ex2_approx(__uint_as_float(r[i]) + sq));
                    }
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&pb.t_empty[buf]);
            }
            p.ey[(size_t)row * p.m_pad + tile * 128 + qd * 32 + lane] = acc * p.inv_bg;
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace

int fdb_kshap_launch_eval_tc(fdb_kshap_plan* pl, cudaStream_t st, const float* x, int64_t row_offset, int n_rows,
                             const int32_t* set_idx, float* ey) {
    TcParams p;
    p.x = x; p.set_idx = set_idx; p.aimg = pl->d_aimg; p.bimg = pl->d_bimg; p.wq = pl->d_wq; p.ey = ey;
    p.row_offset = row_offset; p.n_rows = n_rows; p.n_sets = pl->n_sets; p.tiles = pl->m_pad / 128; p.m_pad = pl->m_pad;
    p.n_bg = pl->n_bg; p.nb_pad = pl->nb_pad; p.nsplit = pl->nsplit; p.fmt = pl->fmt;
    p.n_chunks = (pl->nb_pad + 239) / 240;
    p.chunk_cols = (((pl->nb_pad + p.n_chunks - 1) / p.n_chunks) + 15) & ~15;
    if (p.chunk_cols + 16 <= 128) { p.buf_cols = 128; p.nbuf = 2; } else { p.buf_cols = 256; p.nbuf = 1; }
    p.inv_bg = 1.0f / (float)pl->n_bg;
    const size_t smem = (size_t)pl->nsplit * pl->nb_pad * 64 + 2 * kAStages * (kATileBytes + kBqBytes) + sizeof(TcBars) + 64;
    FDB_REQUIRE(smem <= 227 * 1024, FDB_ERR_ARG, "background of %d rows with %d operand splits needs %zu bytes of shared memory",
                pl->n_bg, pl->nsplit, smem);
    static size_t configured = 0;
    if (smem > configured) {
        FDB_CUDA(cudaFuncSetAttribute(kshap_eval_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    const int64_t n_items = (int64_t)n_rows * p.tiles;
    int64_t grid = (n_items + 1) / 2;
    if (grid > pl->ctx->sm_count) grid = pl->ctx->sm_count;
    if (grid < 1) grid = 1;
    kshap_eval_tc_kernel<<<(int)grid, kThreadsTc, smem, st>>>(p);
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}
