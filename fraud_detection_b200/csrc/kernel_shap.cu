// Kernel SHAP on the B200: plan (per model/background/mask-set state), the SIMT reference
// evaluator, K3 (constrained weighted least squares) and the C-ABI entry points.
// The tensor-core evaluator (K2) lives in kshap_eval_tc.cu.
//
// Algorithm restated from shap.KernelExplainer (see oracle/kernel_shap.py for the step list):
//   ey[r][m]  = mean_b sigmoid( sum_k (1 - z_mk) w_k bg_bk + sum_k z_mk w_k x_rk + c )        (evaluate)
//   eyAdj     = link(ey) - link(fnull);  d = link(f(x)) - link(fnull)
//   X = Z[:, :29] - Z[:, 29];  y = eyAdj - Z[:, 29] d;  w = (X'WX)^-1 X'W y;  phi[29] = d - sum(w)   (solve)
// X'WX depends only on the mask set, so it is inverted once per set at plan creation
// (kshap_wls_prepare) and every explained row costs one X'W y accumulation and one 29x29
// mat-vec (kshap_wls_solve, one warp per row).

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <new>
#include <vector>

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "kshap_internal.cuh"

namespace {

constexpr int kNV = FDB_F - 1;   // 29 unknowns after eliminating the last feature

__device__ __forceinline__ float link_apply(float v, int link) {       // see fdb_kshap_link (stored-value encoding)
    if (link != 1) return v;
    if (v < 0.f) {
        const float c = fmaxf(-v, 1e-37f);
        return logf((1.0f - c) / c);
    }
    v = fminf(fmaxf(v, 1e-37f), 0.99999994f);
    return logf(v / (1.0f - v));
}

// ---------------------------------------------------------------------------------------
// bits -> A operand image (tcgen05 canonical K-major no-swizzle layout, 128 masks x 32 k per tile)
//   a[m][k] = 1 - z_mk (k < 30), a[m][30] = 1 (bias lane), a[m][31] = 0
__global__ void kshap_expand_masks(const uint32_t* __restrict__ bits, uint16_t* __restrict__ aimg, int64_t total,
                                   uint16_t one) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // global mask index (set * m_pad + m)
    if (idx >= total) return;
    const uint32_t b = bits[idx];
    const uint32_t abar = (~b & 0x3FFFFFFFu) | (1u << 30);
    const int64_t tile = idx >> 7;
    const int row = (int)(idx & 127);
    uint4* dst = reinterpret_cast<uint4*>(aimg + tile * (128 * 32)) + (row >> 3) * 32 + (row & 7);   // 16-byte units
#pragma unroll
    for (int kc = 0; kc < 4; ++kc) {
        uint32_t w[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int k = kc * 8 + p * 2;
            w[p] = (((abar >> k) & 1u) ? (uint32_t)one : 0u) | (((abar >> (k + 1)) & 1u) ? ((uint32_t)one << 16) : 0u);
        }
        dst[kc * 8] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// ---------------------------------------------------------------------------------------
// Per mask set: A = X'WX (29x29, f64), c2 = X'W z29, then A^-1 by Gauss-Jordan (SPD, no pivoting).
// One warp per set.  Lane j owns column j of [A | I] in registers (A is symmetric, so the row it accumulated is that
// column): step k needs the pivot row - element (k, j) is already in lane j - and the multipliers A[i][k], which
// are lane k's column and travel by shuffle.  No shared memory, no barriers.
__global__ void __launch_bounds__(32) kshap_wls_prepare(const uint32_t* __restrict__ bits, const float* __restrict__ kw,
                                                        int m_pad, float* __restrict__ ainv, float* __restrict__ c2out,
                                                        int* __restrict__ status, uint32_t varying, int* __restrict__ bad) {
    const int s = blockIdx.x;
    const int lane = threadIdx.x;
    const uint32_t* sb = bits + (size_t)s * m_pad;
    const float* sk = kw + (size_t)s * m_pad;
    // x_i x_j = zp_i zp_j with zp = z29 ? ~z : z (x_i = z_i - z_29 = z29 ? -zp_i : zp_i);  c2 = X'W z29 = -sum w zp z29
    double a[kNV];
    double c2;
    __shared__ float s_tile[2 * 32 * 33];
    auto word = [](uint32_t b) { return ((b >> 29) & 1u) ? ~b : b; };
    auto flag = [](uint32_t b) { return ((b >> 29) & 1u) != 0u; };
    // draw-count weights (sampled sets, shap's own sets) take the tensor-core Gram; anything else the exact generic path
    if (!fdb_kshap_weighted_gram_row_mma<kNV>(sb, sk, m_pad, lane, word, flag, s_tile, a, c2))
        fdb_kshap_weighted_gram_row<kNV>(sb, sk, m_pad, lane, word, flag, a, c2);
    c2out[(size_t)s * 32 + lane] = lane < kNV ? (float)(-c2) : 0.f;
    double inv[kNV];
#pragma unroll
    for (int i = 0; i < kNV; ++i) {
        inv[i] = (i == lane) ? 1.0 : 0.0;
        // a feature outside `varying` has an all-zero row and column (its bits follow bit 29): pin its unknown at 0
        if (i == lane && !((varying >> lane) & 1u)) a[i] = 1.0;
        if (lane >= kNV) a[i] = 0.0;
    }
    // A pivot that has lost 11 digits against its own diagonal entry means X'WX is singular to working precision
    // (too few or linearly dependent coalitions): report it instead of returning an inverse made of rounding noise.
    double diag0 = 0.0;
#pragma unroll
    for (int i = 0; i < kNV; ++i)
        if (i == lane) diag0 = a[i];
    bool singular = false;
#pragma unroll
    for (int k = 0; k < kNV; ++k) {
        const double pk = __shfl_sync(0xffffffffu, a[k], k);          // A[k][k]
        const double dk = __shfl_sync(0xffffffffu, diag0, k);
        if (!(pk > 1e-11 * dk) || !(pk > 1e-300)) { singular = true; break; }
        const double r = 1.0 / pk;
        const double ak = a[k] * r, ik = inv[k] * r;                    // pivot row, scaled: elements (k, lane)
#pragma unroll
        for (int i = 0; i < kNV; ++i) {
            const double f = __shfl_sync(0xffffffffu, a[i], k);       // A[i][k] before this step
            if (i != k) {
                a[i] = fma(-f, ak, a[i]);
                inv[i] = fma(-f, ik, inv[i]);
            }
        }
        a[k] = ak;
        inv[k] = ik;
    }
#pragma unroll
    for (int i = 0; i < kNV; ++i)
        ainv[((size_t)s * kNV + i) * 32 + lane] = (lane < kNV && !singular) ? (float)inv[i] : 0.f;
    if (lane == 0) {
        status[s] = singular ? 1 : 0;
        if (singular) atomicAdd(bad, 1);          // polled by fdb_kshap_plan_status; a singular set yields phi = 0 except the last feature
    }
}

// ---------------------------------------------------------------------------------------
// SIMT evaluator: one thread per (row, mask), plain fp32 FMAs and full-precision expf.
// The GPU-side cross-check of K2 and the evaluator for shapes K2 does not take.
__global__ void __launch_bounds__(128) kshap_eval_simt(const float* __restrict__ x, int64_t row_offset,
                                                       const int32_t* __restrict__ set_idx, int n_sets,
                                                       const uint32_t* __restrict__ bits, int m_pad,
                                                       const float* __restrict__ P, int n_bg,
                                                       const float* __restrict__ wq, float* __restrict__ ey, int link) {
    __shared__ float q[32];
    const int r = blockIdx.y;
    const int m = blockIdx.x * 128 + threadIdx.x;
    if (threadIdx.x < 32) q[threadIdx.x] = threadIdx.x < FDB_F ? wq[threadIdx.x] * x[(size_t)r * FDB_F + threadIdx.x] : 0.f;
    __syncthreads();
    const int s = set_idx ? set_idx[r] : (n_sets == 1 ? 0 : (int)((row_offset + r) % n_sets));
    const uint32_t b = bits[(size_t)s * m_pad + m];
    float sq = 0.f;
#pragma unroll
    for (int k = 0; k < FDB_F; ++k) sq += ((b >> k) & 1u) ? q[k] : 0.f;
    double acc = 0.0, comp = 0.0;          // (float64 running sums: a plain float32 sum over 1000 rows drifts by 1e-5)
    for (int j = 0; j < n_bg; ++j) {
        const float* pr = P + (size_t)j * 32;
        float g = pr[30];   // intercept
#pragma unroll
        for (int k = 0; k < FDB_F; ++k) g += ((b >> k) & 1u) ? 0.f : pr[k];
        acc += 1.0f / (1.0f + expf(-(g + sq)));
        if (link == 1) comp += 1.0f / (1.0f + expf(g + sq));
    }
    ey[(size_t)r * m_pad + m] = link == 1 ? fdb_kshap_encode_ey((float)(acc / n_bg), (float)(comp / n_bg), 1.0f) : (float)(acc / n_bg);
}

// ey_out: the caller's view of the coalition expectations (plain values, n_masks per row)
__global__ void kshap_copy_ey(const float* __restrict__ ey, int m_pad, float* __restrict__ out, int n_masks, int64_t total,
                              int link) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int64_t r = i / n_masks;
    const float v = ey[r * m_pad + (i - r * n_masks)];
    out[i] = link == 1 ? fdb_kshap_decode_ey(v) : v;
}

// ---------------------------------------------------------------------------------------
// K3: one warp per explained row.
__global__ void __launch_bounds__(128) kshap_wls_solve(
    const float* __restrict__ x, int64_t row_offset, int n_rows, const int32_t* __restrict__ set_idx, int n_sets,
    const uint32_t* __restrict__ bits, const float* __restrict__ kw, int m_pad, const float* __restrict__ ey,
    const float* __restrict__ ainv, const float* __restrict__ c2, const float* __restrict__ wcoef, float intercept,
    float fnull, int link, float* __restrict__ phi, double* __restrict__ abs_sum, int accumulate,
    int64_t total_rows, double* __restrict__ partials, unsigned int* __restrict__ ticket,
    const float* __restrict__ gpart, int gslots, long long fx_units, int fx_grid, int fx_tiles) {
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const float lnull = link_apply(fnull, link);
    const float wk = lane < FDB_F ? wcoef[lane] : 0.f;
    double abs_acc = 0.0;
    for (int r = blockIdx.x * 4 + warp; r < n_rows; r += gridDim.x * 4) {
        const int s = set_idx ? set_idx[r] : (n_sets == 1 ? 0 : (int)((row_offset + r) % n_sets));
        float z = lane < FDB_F ? wk * x[(size_t)r * FDB_F + lane] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) z += __shfl_xor_sync(0xffffffffu, z, o);
        // link(f(x)): under the logit link it is the row's logit itself (no round trip through sigmoid)
        const float lfx = link == 1 ? z + intercept : 1.0f / (1.0f + expf(-(z + intercept)));
        const float d = lfx - lnull;
        const uint32_t* sb = bits + (size_t)s * m_pad;
        const float* sk = kw + (size_t)s * m_pad;
        const float* er = ey + (size_t)r * m_pad;
        float g;
        if (gpart != nullptr) {
            // X'W y already accumulated by the factored evaluator (fused GEMM, kshap_eval_fx.cu): one record per CTA that
            // worked on this row's block of 128 rows.  Units are dealt to CTAs as contiguous ranges, so the CTAs of a
            // row block are consecutive; fold their records in CTA order (deterministic).
            const long long ub = (long long)(r >> 7) * fx_tiles;
            const int nseg = (int)((((ub + fx_tiles) * fx_grid - 1) / fx_units) - (((ub + 1) * fx_grid - 1) / fx_units)) + 1;
            const float* gp = gpart + (size_t)r * gslots * 32 + lane;
            g = 0.f;
            for (int q = 0; q < nseg; ++q) g += gp[(size_t)q * 32];
        } else {
            // X'W y: every lane owns one coalition per step and keeps all 29 partial sums in registers
            // (x_mj * v = zp_j ? vp : 0 after folding the sign of the eliminated feature into vp / zp) ...
            float gj[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) gj[j] = 0.f;
            for (int m0 = 0; m0 < m_pad; m0 += 32) {
                const uint32_t b = sb[m0 + lane];
                const float kwv = sk[m0 + lane];
                const float v = kwv != 0.f ? kwv * (link_apply(er[m0 + lane], link) - lnull) : 0.f;
                const bool z29 = (b >> 29) & 1u;
                const float vp = z29 ? -v : v;
                const uint32_t bp = z29 ? ~b : b;
#pragma unroll
                for (int j = 0; j < kNV; ++j) gj[j] += ((bp >> j) & 1u) ? vp : 0.f;
            }
            // ... then a butterfly transpose-reduce leaves the total of column j in lane j (31 shuffles)
#pragma unroll
            for (int half = 16; half >= 1; half >>= 1) {
                const bool up = (lane & half) != 0;
#pragma unroll
                for (int i = 0; i < half; ++i) {
                    const float send = up ? gj[i] : gj[i + half];
                    const float keep = up ? gj[i + half] : gj[i];
                    gj[i] = keep + __shfl_xor_sync(0xffffffffu, send, half);
                }
            }
            g = gj[0];
        }
        const float rhs = lane < kNV ? g - d * c2[(size_t)s * 32 + lane] : 0.f;
        const float* ai = ainv + (size_t)s * kNV * 32;
        float w = 0.f;
#pragma unroll
        for (int i = 0; i < kNV; ++i) w += ai[i * 32 + lane] * __shfl_sync(0xffffffffu, rhs, i);
        float sumw = lane < kNV ? w : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sumw += __shfl_xor_sync(0xffffffffu, sumw, o);
        float p = lane < kNV ? w : (lane == kNV ? d - sumw : 0.f);
        if (fabsf(p) < 1e-10f) p = 0.f;
        if (lane < FDB_F) phi[(size_t)r * FDB_F + lane] = p;
        abs_acc += fabs((double)p);
    }
    if (abs_sum == nullptr) return;
    __shared__ double s_acc[4][32];
    __shared__ bool s_last;
    s_acc[warp][lane] = abs_acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        partials[(size_t)blockIdx.x * 32 + lane] = s_acc[0][lane] + s_acc[1][lane] + s_acc[2][lane] + s_acc[3][lane];
        __threadfence();
    }
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    __syncthreads();
    if (s_last && threadIdx.x < 32) {
        __threadfence();
        double tot = 0.0;
        if (lane < FDB_F) {
            for (unsigned b = 0; b < gridDim.x; ++b) tot += __ldcg(partials + (size_t)b * 32 + lane);
            if (accumulate) tot += abs_sum[lane];
        } else if (lane == FDB_F) {
            tot = (double)total_rows;
        }
        abs_sum[lane] = tot;
        if (lane == 0) *ticket = 0u;
    }
}

// ---------------------------------------------------------------------------------------
uint16_t f32_to_bf16(float f) { return __bfloat16_as_ushort(__float2bfloat16_rn(f)); }
float bf16_to_f32(uint16_t h) { return __bfloat162float(__ushort_as_bfloat16(h)); }
uint16_t f32_to_f16(float f) { return __half_as_ushort(__float2half_rn(f)); }
float f16_to_f32(uint16_t h) { return __half2float(__ushort_as_half(h)); }

void free_plan(fdb_kshap_plan* p) {
    if (!p) return;
    cudaFree(p->d_bits); cudaFree(p->d_kw); cudaFree(p->d_P); cudaFree(p->d_wq); cudaFree(p->d_bimg); cudaFree(p->d_aimg);
    cudaFree(p->d_ainv); cudaFree(p->d_c2); cudaFree(p->d_status); cudaFree(p->d_ey);
    cudaFree(p->d_abs); cudaFree(p->d_E); cudaFree(p->d_eoff); cudaFree(p->d_fxtile); cudaFree(p->d_gpart); cudaFree(p->d_gz); cudaFree(p->d_bv); cudaFree(p->d_s0);
    fdb_kshap_free_lasso(p);
    cudaFree(p->d_partials); cudaFree(p->d_ticket); cudaFree(p->d_bad);
    fdb_kshap_sampler_free(p);
    for (int i = 0; i < 2; ++i) {
        cudaFree(p->d_x2[i]); cudaFree(p->d_phi2[i]); cudaFree(p->d_set2[i]);
        for (int e = 0; e < 3; ++e)
            if (p->ev[i][e]) cudaEventDestroy(p->ev[i][e]);
    }
    delete p;
}

}  // namespace

// Split v into `n` 16-bit pieces (f16 or bf16) whose sum reproduces v to ~n*11 / n*8 bits.
void fdb_kshap_split(double v, int fmt, int n, uint16_t* out) {
    double rem = v;
    for (int i = 0; i < n; ++i) {
        const uint16_t h = fmt == 0 ? f32_to_f16((float)rem) : f32_to_bf16((float)rem);
        out[i] = h;
        rem -= (double)(fmt == 0 ? f16_to_f32(h) : bf16_to_f32(h));
    }
}

// Everything derived from the coalition sets (d_bits / d_kw): operand tile images, per-set (X'WX)^-1 and, for a
// single set, the factored evaluator's operands.  Run at plan creation and again after a resample.
// Stream-ordered part of the refresh for plans with several sets: per-set (X'WX)^-1 (and the Z'WZ products when feature
// selection is on), enqueued on `st` with no host synchronisation.  Singular sets are counted in d_bad.
static int refresh_set_products_async(fdb_kshap_plan* p, cudaStream_t st) {
    // (X'WX)^-1 is NOT rebuilt here but by the first solve that needs it (ensure_inverses): a plan whose rows go through a
    // feature selection (l1_reg) never reads it, and its 32,768 Gauss-Jordan inversions are a quarter of a redraw
    p->ainv_valid = false;
    if (p->d_gz && p->l1_mode != FDB_KSHAP_L1_NONE) {          // Z'WZ products: every selection reads them
        int rc = fdb_kshap_refresh_lars(p, st);
        if (rc != FDB_OK) return rc;
        if (p->l1_mode > FDB_KSHAP_L1_NUM_FEATURES) return fdb_kshap_refresh_lasso(p, st);
        p->lasso_scaled = -1;                                  // centred Gram products (criteria / alpha) are stale now
    } else if (p->d_gz) {
        p->gz_valid = false;
        p->lasso_scaled = -1;
    }
    return FDB_OK;
}

// per-set (X'WX)^-1 for the current sets, stream-ordered (singular sets are counted in d_bad)
static int ensure_inverses(fdb_kshap_plan* p, cudaStream_t st) {
    if (p->ainv_valid) return FDB_OK;
    kshap_wls_prepare<<<p->n_sets, 32, 0, st>>>(p->d_bits, p->d_kw, p->m_pad, p->d_ainv, p->d_c2, p->d_status, p->varying, p->d_bad);
    FDB_CUDA(cudaGetLastError());
    p->ainv_valid = true;
    return FDB_OK;
}

static int refresh_set_products(fdb_kshap_plan* p) {
    const size_t SM = (size_t)p->n_sets * p->m_pad;
    const uint16_t one = p->fmt == 0 ? 0x3C00 : 0x3F80;
    if (p->d_aimg) {      // single-set plans only: the tile images the factored evaluator streams by TMA
        kshap_expand_masks<<<(unsigned)((SM + 255) / 256), 256>>>(p->d_bits, p->d_aimg, (int64_t)SM, one);
        FDB_CUDA(cudaGetLastError());
    }
    kshap_wls_prepare<<<p->n_sets, 32>>>(p->d_bits, p->d_kw, p->m_pad, p->d_ainv, p->d_c2, p->d_status, p->varying, p->d_bad);
    FDB_CUDA(cudaGetLastError());
    FDB_CUDA(cudaDeviceSynchronize());
    p->ainv_valid = true;
    FDB_CUDA(cudaMemset(p->d_bad, 0, sizeof(int)));
    std::vector<int> st(p->n_sets);
    FDB_CUDA(cudaMemcpy(st.data(), p->d_status, (size_t)p->n_sets * 4, cudaMemcpyDeviceToHost));
    for (int s = 0; s < p->n_sets; ++s)
        FDB_REQUIRE(st[s] == 0, FDB_ERR_ARG, "mask set %d gives a singular weighted least-squares system (X'WX); "
                    "use more samples or fewer features", s);
    if (p->n_sets == 1) {
        cudaFree(p->d_E); cudaFree(p->d_eoff);
        p->d_E = nullptr; p->d_eoff = nullptr;
        int rc = fdb_kshap_build_factored(p, p->bg_host.data());
        if (rc != FDB_OK) return rc;
    }
    if (p->d_gz) {
        cudaFree(p->d_gz); cudaFree(p->d_bv); cudaFree(p->d_s0);
        p->d_gz = nullptr; p->d_bv = nullptr; p->d_s0 = nullptr;
        int rc = fdb_kshap_prepare_lars(p);
        if (rc != FDB_OK) return rc;
        if (p->d_lg) {
            const int scale = p->lasso_scaled;
            fdb_kshap_free_lasso(p);
            if (scale >= 0) {                                  // -1: stale already, the next fdb_kshap_plan_set_l1 rebuilds them
                rc = fdb_kshap_prepare_lasso(p, scale);
                if (rc != FDB_OK) return rc;
            }
        }
    }
    return FDB_OK;
}

// bits == nullptr: draw the coalition sets on the device (n_masks = resolved nsamples, seed)
static int plan_create_impl(fdb_ctx* ctx, const double* coef, double intercept, int n_bg, const double* bg,
                            int n_sets, int n_masks, const uint32_t* bits, const double* kw, int link,
                            unsigned long long seed, uint32_t varying, fdb_kshap_plan** out) {
    FDB_REQUIRE(ctx && coef && bg && out, FDB_ERR_ARG, "NULL argument");
    *out = nullptr;
    FDB_REQUIRE((varying & ~0x3FFFFFFFu) == 0 && ((varying >> 29) & 1u) && __builtin_popcount(varying) >= 2, FDB_ERR_ARG,
                "varying_mask 0x%x must name at least 2 of the 30 features and include feature 29", varying);
    FDB_REQUIRE(n_bg >= 1 && n_bg <= FDB_KSHAP_MAX_BG, FDB_ERR_ARG, "background rows must be in [1, %d], got %d", FDB_KSHAP_MAX_BG, n_bg);
    FDB_REQUIRE(n_sets >= 1 && n_masks >= 1, FDB_ERR_ARG, "need at least one mask set and one mask");
    FDB_REQUIRE(link == 0 || link == 1, FDB_ERR_ARG, "link must be 0 (identity) or 1 (logit)");
    DeviceGuard guard(ctx->device);
    FDB_REQUIRE(guard.ok, FDB_ERR_CUDA, "cannot select device %d", ctx->device);
    fdb_kshap_plan* p = new (std::nothrow) fdb_kshap_plan();
    FDB_REQUIRE(p != nullptr, FDB_ERR_NOMEM, "out of host memory");
    p->ctx = ctx;
    p->n_bg = n_bg;
    p->nb_pad = (n_bg + 15) & ~15;
    p->n_masks = n_masks;
    p->m_pad = (n_masks + 127) & ~127;
    p->n_sets = n_sets;
    p->link = link;
    p->varying = varying;
    p->intercept = intercept;
    memcpy(p->coef, coef, sizeof(double) * FDB_F);
    const size_t SM = (size_t)n_sets * p->m_pad;

    // ---- host-side constant preparation (parameters, not rows) -----------------------------
    const double LOG2E = 1.4426950408889634074;
    std::vector<float> P((size_t)n_bg * 32, 0.f);
    double fnull = 0.0, vmax = 0.0;
    for (int b = 0; b < n_bg; ++b) {
        double z = intercept;
        for (int k = 0; k < FDB_F; ++k) {
            const double pk = coef[k] * bg[(size_t)b * FDB_F + k];
            P[(size_t)b * 32 + k] = (float)pk;
            z += pk;
            vmax = fmax(vmax, fabs(LOG2E * pk));
        }
        P[(size_t)b * 32 + 30] = (float)intercept;
        fnull += 1.0 / (1.0 + exp(-z));
    }
    vmax = fmax(vmax, fabs(LOG2E * intercept));
    p->fnull = fnull / n_bg;
    const char* force = getenv("FDB_KSHAP_FMT");
    p->fmt = (vmax < 16384.0) ? 0 : 1;                 // f16 x 2 splits unless the operand would overflow half
    if (force) p->fmt = atoi(force) ? 1 : 0;
    p->nsplit = p->fmt == 0 ? 2 : 3;
    // B operand image: [split][row-group][k-chunk][8 rows][8 elems], value -log2e * w_k * bg_bk, col 30 = -log2e * c
    std::vector<uint16_t> bimg((size_t)p->nsplit * p->nb_pad * 32, 0);
    for (int b = 0; b < n_bg; ++b)
        for (int k = 0; k < 31; ++k) {
            const double v = -LOG2E * (k < FDB_F ? coef[k] * bg[(size_t)b * FDB_F + k] : intercept);
            uint16_t parts[3];
            fdb_kshap_split(v, p->fmt, p->nsplit, parts);
            for (int sidx = 0; sidx < p->nsplit; ++sidx)
                bimg[(size_t)sidx * p->nb_pad * 32 + (size_t)(b >> 3) * 256 + (k >> 3) * 64 + (b & 7) * 8 + (k & 7)] = parts[sidx];
        }
    std::vector<uint32_t> hbits(bits ? SM : 0, 0u);
    std::vector<float> hkw(bits ? SM : 0, 0.f);
    for (int s = 0; bits && s < n_sets; ++s)
        for (int m = 0; m < n_masks; ++m) {
            // features outside `varying` follow bit 29, so that their column z_k - z_29 of the WLS is identically 0
            const uint32_t b = bits[(size_t)s * n_masks + m] & varying;
            hbits[(size_t)s * p->m_pad + m] = ((b >> 29) & 1u) ? (b | (~varying & 0x3FFFFFFFu)) : b;
            hkw[(size_t)s * p->m_pad + m] = (float)kw[(size_t)s * n_masks + m];
        }
    float wq[32] = {0};
    for (int k = 0; k < FDB_F; ++k) wq[k] = (float)coef[k];

    int rc = FDB_OK;
    cudaError_t e = cudaSuccess;
    do {
#define TRY(x) if ((e = (x)) != cudaSuccess) break
        TRY(cudaMalloc((void**)&p->d_bits, SM * 4));
        TRY(cudaMalloc((void**)&p->d_kw, SM * 4));
        TRY(cudaMalloc((void**)&p->d_P, P.size() * 4));
        TRY(cudaMalloc((void**)&p->d_wq, 32 * 4));
        TRY(cudaMalloc((void**)&p->d_bimg, bimg.size() * 2));
        if (n_sets == 1) TRY(cudaMalloc((void**)&p->d_aimg, SM * 32 * 2));
        TRY(cudaMalloc((void**)&p->d_ainv, (size_t)n_sets * kNV * 32 * 4));
        TRY(cudaMalloc((void**)&p->d_c2, (size_t)n_sets * 32 * 4));
        TRY(cudaMalloc((void**)&p->d_status, (size_t)n_sets * 4));
        p->batch_rows = FDB_KSHAP_BATCH_ROWS;          // the scratch between evaluate and solve is sized by the first calls (ensure_scratch)
        TRY(cudaMalloc((void**)&p->d_abs, 2 * 32 * 8));
        TRY(cudaMalloc((void**)&p->d_partials, (size_t)FDB_MAX_GRID * 32 * sizeof(double)));
        TRY(cudaMalloc((void**)&p->d_ticket, sizeof(unsigned int)));
        TRY(cudaMemset(p->d_ticket, 0, sizeof(unsigned int)));
        TRY(cudaMalloc((void**)&p->d_bad, sizeof(int)));
        TRY(cudaMemset(p->d_bad, 0, sizeof(int)));
        if (bits) {
            TRY(cudaMemcpy(p->d_bits, hbits.data(), SM * 4, cudaMemcpyHostToDevice));
            TRY(cudaMemcpy(p->d_kw, hkw.data(), SM * 4, cudaMemcpyHostToDevice));
        } else {
            rc = fdb_kshap_sample_on_device(p, nullptr, n_masks, seed);
            if (rc != FDB_OK) break;
            TRY(cudaDeviceSynchronize());
        }
        TRY(cudaMemcpy(p->d_P, P.data(), P.size() * 4, cudaMemcpyHostToDevice));
        TRY(cudaMemcpy(p->d_wq, wq, 32 * 4, cudaMemcpyHostToDevice));
        TRY(cudaMemcpy(p->d_bimg, bimg.data(), bimg.size() * 2, cudaMemcpyHostToDevice));
#undef TRY
    } while (0);
    if (e != cudaSuccess) {
        fdb_set_error("kernel-SHAP plan creation failed: %s", cudaGetErrorString(e));
        rc = e == cudaErrorMemoryAllocation ? FDB_ERR_NOMEM : FDB_ERR_CUDA;
    }
    if (rc == FDB_OK) {
        p->bg_host.assign(bg, bg + (size_t)n_bg * FDB_F);
        p->sampled = bits == nullptr;
        rc = refresh_set_products(p);
    }
    // A plan must have at least one fast evaluator that can run: the tensor-core one keeps the whole background
    // resident in shared memory (192 B per row in the bf16 x 3 format that large-magnitude rows need), the factored one
    // exists for single-set plans only.  Say so here, not at the first explain call.
    if (rc == FDB_OK && fdb_kshap_tc_smem_bytes(p) > 227 * 1024 && !(p->n_sets == 1 && p->d_E != nullptr)) {
        fdb_set_error("a background of %d rows in the %s operand format needs %zu bytes of shared memory per SM (limit %d): "
                      "use at most %d background rows for this model", n_bg, p->fmt == 0 ? "f16 x 2" : "bf16 x 3",
                      fdb_kshap_tc_smem_bytes(p), 227 * 1024, p->fmt == 0 ? FDB_KSHAP_MAX_BG : 800);
        rc = FDB_ERR_ARG;
    }
    if (rc != FDB_OK) { free_plan(p); return rc; }
    *out = p;
    return FDB_OK;
}

extern "C" int fdb_kshap_plan_create(fdb_ctx* ctx, const double* coef, double intercept, int n_bg, const double* bg,
                                     int n_sets, int n_masks, const uint32_t* bits, const double* kw, int link,
                                     fdb_kshap_plan** out) {
    FDB_REQUIRE(bits && kw, FDB_ERR_ARG, "bits / kw is NULL");
    return plan_create_impl(ctx, coef, intercept, n_bg, bg, n_sets, n_masks, bits, kw, link, 0ULL, 0x3FFFFFFFu, out);
}

extern "C" int fdb_kshap_plan_create_varying(fdb_ctx* ctx, const double* coef, double intercept, int n_bg, const double* bg,
                                             int n_sets, int n_masks, const uint32_t* bits, const double* kw, int link,
                                             uint32_t varying_mask, fdb_kshap_plan** out) {
    FDB_REQUIRE(bits && kw, FDB_ERR_ARG, "bits / kw is NULL");
    return plan_create_impl(ctx, coef, intercept, n_bg, bg, n_sets, n_masks, bits, kw, link, 0ULL, varying_mask, out);
}

extern "C" int fdb_kshap_plan_create_sampled(fdb_ctx* ctx, const double* coef, double intercept, int n_bg,
                                             const double* bg, int n_sets, int nsamples, uint64_t seed, int link,
                                             fdb_kshap_plan** out) {
    return plan_create_impl(ctx, coef, intercept, n_bg, bg, n_sets, fdb_kshap_resolve_nsamples(nsamples), nullptr, nullptr,
                            link, (unsigned long long)seed, 0x3FFFFFFFu, out);
}

extern "C" int fdb_kshap_plan_resample(fdb_kshap_plan* p, uint64_t seed) {
    FDB_REQUIRE(p != nullptr, FDB_ERR_ARG, "plan is NULL");
    FDB_REQUIRE(p->sampled, FDB_ERR_STATE, "only plans made by fdb_kshap_plan_create_sampled can be resampled");
    std::lock_guard<std::mutex> lock(p->ctx->host_mu);
    DeviceGuard guard(p->ctx->device);
    FDB_REQUIRE(guard.ok, FDB_ERR_CUDA, "cannot select device %d", p->ctx->device);
    FDB_CUDA(cudaDeviceSynchronize());
    int rc = fdb_kshap_sample_on_device(p, nullptr, p->n_masks, (unsigned long long)seed);
    if (rc != FDB_OK) return rc;
    FDB_CUDA(cudaDeviceSynchronize());
    return refresh_set_products(p);
}

extern "C" int fdb_kshap_plan_resample_dev(fdb_kshap_plan* p, void* stream, uint64_t seed) {
    FDB_REQUIRE(p != nullptr, FDB_ERR_ARG, "plan is NULL");
    FDB_REQUIRE(p->sampled, FDB_ERR_STATE, "only plans made by fdb_kshap_plan_create_sampled can be resampled");
    FDB_REQUIRE(p->n_sets > 1, FDB_ERR_STATE, "a single-set plan rebuilds the factored evaluator's operands on the host: use fdb_kshap_plan_resample");
    int cur = -1;
    FDB_CUDA(cudaGetDevice(&cur));
    FDB_REQUIRE(cur == p->ctx->device, FDB_ERR_STATE, "the plan lives on device %d but the current device is %d", p->ctx->device, cur);
    int rc = fdb_kshap_sample_on_device(p, (cudaStream_t)stream, p->n_masks, (unsigned long long)seed);
    if (rc != FDB_OK) return rc;
    return refresh_set_products_async(p, (cudaStream_t)stream);
}

extern "C" int fdb_kshap_plan_status(fdb_kshap_plan* p, int* n_singular) {
    FDB_REQUIRE(p != nullptr && n_singular != nullptr, FDB_ERR_ARG, "NULL argument");
    DeviceGuard guard(p->ctx->device);
    FDB_REQUIRE(guard.ok, FDB_ERR_CUDA, "cannot select device %d", p->ctx->device);
    FDB_CUDA(cudaMemcpy(n_singular, p->d_bad, sizeof(int), cudaMemcpyDeviceToHost));
    FDB_CUDA(cudaMemset(p->d_bad, 0, sizeof(int)));
    return FDB_OK;
}

extern "C" int fdb_kshap_plan_get_masks(const fdb_kshap_plan* p, int set, uint32_t* bits, float* kw) {
    FDB_REQUIRE(p && bits && kw, FDB_ERR_ARG, "NULL argument");
    FDB_REQUIRE(set >= 0 && set < p->n_sets, FDB_ERR_ARG, "set %d out of range", set);
    DeviceGuard guard(p->ctx->device);
    FDB_REQUIRE(guard.ok, FDB_ERR_CUDA, "cannot select device %d", p->ctx->device);
    FDB_CUDA(cudaMemcpy(bits, p->d_bits + (size_t)set * p->m_pad, (size_t)p->n_masks * 4, cudaMemcpyDeviceToHost));
    FDB_CUDA(cudaMemcpy(kw, p->d_kw + (size_t)set * p->m_pad, (size_t)p->n_masks * 4, cudaMemcpyDeviceToHost));
    return FDB_OK;
}

extern "C" int fdb_kshap_plan_destroy(fdb_kshap_plan* p) {
    if (p) {
        DeviceGuard guard(p->ctx->device);
        cudaDeviceSynchronize();
        free_plan(p);
    }
    return FDB_OK;
}

extern "C" double fdb_kshap_expected_value(const fdb_kshap_plan* p) {
    if (!p) return 0.0;
    return p->link == 1 ? log(p->fnull / (1.0 - p->fnull)) : p->fnull;
}

extern "C" int fdb_kshap_plan_info(const fdb_kshap_plan* p, int* info8) {
    FDB_REQUIRE(p && info8, FDB_ERR_ARG, "NULL argument");
    info8[0] = p->n_bg; info8[1] = p->nb_pad; info8[2] = p->n_masks; info8[3] = p->m_pad;
    info8[4] = p->n_sets; info8[5] = p->fmt; info8[6] = p->nsplit; info8[7] = (int)p->batch_rows;
    return FDB_OK;
}

extern "C" int fdb_kshap_plan_set_num_features(fdb_kshap_plan* p, int k) {
    FDB_REQUIRE(p != nullptr, FDB_ERR_ARG, "plan is NULL");
    FDB_REQUIRE(k >= 0 && k <= 16, FDB_ERR_ARG, "num_features must be in [0, 16] (0 = no feature selection), got %d", k);
    if (k > 0) {
        DeviceGuard guard(p->ctx->device);
        FDB_REQUIRE(guard.ok, FDB_ERR_CUDA, "cannot select device %d", p->ctx->device);
        int rc = fdb_kshap_prepare_lars(p);
        if (rc != FDB_OK) return rc;
    }
    p->num_features = k;
    p->l1_mode = k > 0 ? FDB_KSHAP_L1_NUM_FEATURES : FDB_KSHAP_L1_NONE;
    return FDB_OK;
}

extern "C" int fdb_kshap_plan_set_l1(fdb_kshap_plan* p, int mode, double param) {
    FDB_REQUIRE(p != nullptr, FDB_ERR_ARG, "plan is NULL");
    FDB_REQUIRE(mode >= FDB_KSHAP_L1_NONE && mode <= FDB_KSHAP_L1_ALPHA, FDB_ERR_ARG, "unknown l1 mode %d", mode);
    if (mode == FDB_KSHAP_L1_NONE) return fdb_kshap_plan_set_num_features(p, 0);
    if (mode == FDB_KSHAP_L1_NUM_FEATURES) {
        FDB_REQUIRE(param >= 1.0 && param <= 16.0 && param == (double)(int)param, FDB_ERR_ARG,
                    "num_features(k) needs an integer 1 <= k <= 16, got %g", param);
        return fdb_kshap_plan_set_num_features(p, (int)param);
    }
    if (mode == FDB_KSHAP_L1_ALPHA)
        FDB_REQUIRE(param > 0.0 && param < 1e300, FDB_ERR_ARG, "the Lasso penalty must be a positive finite number, got %g", param);
    DeviceGuard guard(p->ctx->device);
    FDB_REQUIRE(guard.ok, FDB_ERR_CUDA, "cannot select device %d", p->ctx->device);
    int rc = fdb_kshap_prepare_lasso(p, mode == FDB_KSHAP_L1_ALPHA ? 0 : 1);
    if (rc != FDB_OK) return rc;
    p->num_features = 0;
    p->l1_mode = mode;
    p->l1_alpha = mode == FDB_KSHAP_L1_ALPHA ? param : 0.0;
    return FDB_OK;
}

// Scratch between evaluate and solve, grown on demand and never shrunk: coalition expectations (only when they go
// through HBM: K2 / SIMT evaluators, ey_out, feature selection) and the factored evaluator's X'W y records.  Growing
// frees the old buffer, which waits for the device - it happens on the first calls only.
static int ensure_scratch(fdb_kshap_plan* p, size_t ey_floats, size_t gpart_floats) {
    if (ey_floats > p->ey_cap) {
        cudaFree(p->d_ey);
        p->d_ey = nullptr; p->ey_cap = 0;
        FDB_CUDA(cudaMalloc((void**)&p->d_ey, ey_floats * 4));
        p->ey_cap = ey_floats;
    }
    if (gpart_floats > p->gpart_cap) {
        cudaFree(p->d_gpart);
        p->d_gpart = nullptr; p->gpart_cap = 0;
        FDB_CUDA(cudaMalloc((void**)&p->d_gpart, gpart_floats * 4));
        p->gpart_cap = gpart_floats;
    }
    return FDB_OK;
}

extern "C" int fdb_kshap_explain_dev(fdb_kshap_plan* p, void* stream, int64_t n_rows, const float* x,
                                     const int32_t* set_idx, int evaluator, float* phi, float* ey_out, double* abs_sum) {
    FDB_REQUIRE(p != nullptr, FDB_ERR_ARG, "plan is NULL");
    FDB_REQUIRE(n_rows >= 0, FDB_ERR_ARG, "n_rows is negative");
    FDB_REQUIRE(n_rows == 0 || (x && phi), FDB_ERR_ARG, "x / phi is NULL");
    FDB_REQUIRE(evaluator >= FDB_KSHAP_EVAL_AUTO && evaluator <= FDB_KSHAP_EVAL_FACTORED, FDB_ERR_ARG, "unknown evaluator %d", evaluator);
    if (evaluator == FDB_KSHAP_EVAL_AUTO) evaluator = (p->n_sets == 1 && p->d_E != nullptr) ? FDB_KSHAP_EVAL_FACTORED : FDB_KSHAP_EVAL_TC;
    FDB_REQUIRE(evaluator != FDB_KSHAP_EVAL_FACTORED || p->n_sets == 1, FDB_ERR_ARG,
                "the factored evaluator needs a plan with exactly one coalition set (this plan has %d)", p->n_sets);
    cudaStream_t st = (cudaStream_t)stream;
    if (n_rows == 0) {
        if (abs_sum) FDB_CUDA(cudaMemsetAsync(abs_sum, 0, 32 * sizeof(double), st));
        return FDB_OK;
    }
    int cur = -1;
    FDB_CUDA(cudaGetDevice(&cur));
    FDB_REQUIRE(cur == p->ctx->device, FDB_ERR_STATE, "the plan lives on device %d but the current device is %d", p->ctx->device, cur);
    fdb_ctx* ctx = p->ctx;
    const bool fx = evaluator == FDB_KSHAP_EVAL_FACTORED;
    // num_features(k) selects from X'W y alone, which the factored evaluator leaves in its records; the criteria and the
    // Lasso penalty also need sum kw e^2 and the weighted mean of e, i.e. the expectations themselves
    const bool need_ey = !fx || ey_out != nullptr || p->l1_mode > FDB_KSHAP_L1_NUM_FEATURES;
    // rows per evaluate -> solve round: bounded by the ey scratch when expectations go through HBM; the factored path
    // leaves only 256 B per row behind, so it takes large rounds (every CTA gets an equal share of the units)
    int64_t batch = need_ey ? FDB_KSHAP_BATCH_ROWS : FDB_KSHAP_FX_BATCH_ROWS;
    if (fx) {                                             // equal rounds, whole row blocks
        const int64_t rounds = (n_rows + batch - 1) / batch;
        batch = ((n_rows + rounds - 1) / rounds + 127) & ~int64_t(127);
    }
    {
        const int64_t b0 = n_rows < batch ? n_rows : batch;
        size_t gfl = 0;
        if (fx) {                                         // the last round may be shorter and need more records per row
            long long units; int grid, slots;
            fdb_kshap_fx_geometry(p, (int)b0, &units, &grid, &slots);
            gfl = (size_t)b0 * slots * 32;
            const int64_t last = n_rows - ((n_rows - 1) / batch) * batch;
            fdb_kshap_fx_geometry(p, (int)last, &units, &grid, &slots);
            if ((size_t)last * slots * 32 > gfl) gfl = (size_t)last * slots * 32;
        }
        int rc = ensure_scratch(p, need_ey ? (size_t)b0 * p->m_pad : 0, gfl);
        if (rc != FDB_OK) return rc;
    }
    for (int64_t r0 = 0; r0 < n_rows; r0 += batch) {
        const int nb = (int)((n_rows - r0) < batch ? (n_rows - r0) : batch);
        const float* xb = x + r0 * FDB_F;
        const int32_t* sb = set_idx ? set_idx + r0 : nullptr;
        long long fx_units = 1; int fx_grid = 1, fx_slots = 1;
        if (evaluator == FDB_KSHAP_EVAL_TC) {
            int rc = fdb_kshap_launch_eval_tc(p, st, xb, r0, nb, sb, p->d_ey);
            if (rc != FDB_OK) return rc;
        } else if (fx) {
            fdb_kshap_fx_geometry(p, nb, &fx_units, &fx_grid, &fx_slots);
            int rc = fdb_kshap_launch_eval_fx(p, st, xb, nb, need_ey ? p->d_ey : nullptr, p->d_gpart);
            if (rc != FDB_OK) return rc;
        } else {
            dim3 grid(p->m_pad / 128, nb);
            kshap_eval_simt<<<grid, 128, 0, st>>>(xb, r0, sb, p->n_sets, p->d_bits, p->m_pad, p->d_P, p->n_bg, p->d_wq, p->d_ey, p->link);
            FDB_CUDA(cudaGetLastError());
        }
        if (ey_out) {
            const int64_t total = (int64_t)nb * p->n_masks;
            kshap_copy_ey<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(p->d_ey, p->m_pad, ey_out + r0 * p->n_masks, p->n_masks, total,
                                                                                 p->link);
            FDB_CUDA(cudaGetLastError());
        }
        if (p->l1_mode != FDB_KSHAP_L1_NONE) {       // feature selection + reduced WLS (kshap_lasso.cu)
            int rc = fdb_kshap_launch_lasso_solve(p, st, xb, r0, nb, sb, phi + r0 * FDB_F, abs_sum, r0 > 0 ? 1 : 0, n_rows,
                                                  (fx && !need_ey) ? p->d_gpart : nullptr, fx_slots, fx_units, fx_grid);
            if (rc != FDB_OK) return rc;
            continue;
        }
        int grid = (nb + 3) / 4;
        const int cap = ctx->sm_count * 8;
        if (grid > cap) grid = cap;
        if (!p->ainv_valid) {                      // sets were redrawn on a stream since the inverses were built
            int rc = ensure_inverses(p, st);
            if (rc != FDB_OK) return rc;
        }
        kshap_wls_solve<<<grid, 128, 0, st>>>(xb, r0, nb, sb, p->n_sets, p->d_bits, p->d_kw, p->m_pad, p->d_ey, p->d_ainv,
                                             p->d_c2, p->d_wq, (float)p->intercept, (float)p->fnull, p->link,
                                             phi + r0 * FDB_F, abs_sum, r0 > 0 ? 1 : 0, n_rows, p->d_partials,
                                             p->d_ticket, fx ? p->d_gpart : nullptr, fx_slots, fx_units, fx_grid, p->m_pad / 128);
        FDB_CUDA(cudaGetLastError());
    }
    return FDB_OK;
}

// Host rows in, host phi out.  Batches of FDB_KSHAP_BATCH_ROWS rows go through two staging slots and three of the
// context's streams: the copy-in of batch i + 1 and the copy-out of batch i - 1 overlap evaluate + solve of batch i
// (the plan's own scratch serialises the compute, which is the long part).
extern "C" int fdb_kshap_explain_host(fdb_kshap_plan* p, int64_t n_rows, const float* x, const int32_t* set_idx,
                                      int evaluator, float* phi, double* abs_sum) {
    FDB_REQUIRE(p != nullptr, FDB_ERR_ARG, "plan is NULL");
    FDB_REQUIRE(n_rows >= 0, FDB_ERR_ARG, "n_rows is negative");
    FDB_REQUIRE(n_rows == 0 || (x && phi), FDB_ERR_ARG, "x / phi is NULL");
    if (abs_sum) memset(abs_sum, 0, 32 * sizeof(double));
    if (n_rows == 0) return FDB_OK;
    fdb_ctx* ctx = p->ctx;
    std::lock_guard<std::mutex> lock(ctx->host_mu);
    DeviceGuard guard(ctx->device);
    FDB_REQUIRE(guard.ok, FDB_ERR_CUDA, "cannot select device %d", ctx->device);
    const int64_t batch = FDB_KSHAP_BATCH_ROWS;
    const int64_t n_batches = (n_rows + batch - 1) / batch;
    const int64_t want = n_rows < batch ? n_rows : batch;
    if (want > p->stage_rows) {
        for (int i = 0; i < 2; ++i) {
            cudaFree(p->d_x2[i]); cudaFree(p->d_phi2[i]); cudaFree(p->d_set2[i]);
            p->d_x2[i] = p->d_phi2[i] = nullptr; p->d_set2[i] = nullptr;
        }
        p->stage_rows = 0;
        const int slots = n_batches > 1 ? 2 : 1;
        const int64_t rows = n_batches > 1 ? batch : want;
        for (int i = 0; i < slots; ++i) {
            FDB_CUDA(cudaMalloc((void**)&p->d_x2[i], (size_t)rows * FDB_F * 4));
            FDB_CUDA(cudaMalloc((void**)&p->d_phi2[i], (size_t)rows * FDB_F * 4));
            FDB_CUDA(cudaMalloc((void**)&p->d_set2[i], (size_t)rows * 4));
        }
        p->stage_rows = rows;
    }
    if (n_batches > 1 && p->d_x2[1] == nullptr) {          // first multi-batch call after single-batch ones
        FDB_CUDA(cudaMalloc((void**)&p->d_x2[1], (size_t)p->stage_rows * FDB_F * 4));
        FDB_CUDA(cudaMalloc((void**)&p->d_phi2[1], (size_t)p->stage_rows * FDB_F * 4));
        FDB_CUDA(cudaMalloc((void**)&p->d_set2[1], (size_t)p->stage_rows * 4));
    }
    for (int i = 0; i < 2; ++i)
        for (int e = 0; e < 3; ++e)
            if (!p->ev[i][e]) FDB_CUDA(cudaEventCreateWithFlags(&p->ev[i][e], cudaEventDisableTiming));
    FDB_REQUIRE((size_t)n_batches * 32 * sizeof(double) <= ctx->h_stage_bytes, FDB_ERR_ARG, "too many rows for one call (%lld)", (long long)n_rows);
    double* h_abs = static_cast<double*>(ctx->h_stage);          // pinned: per-batch |phi| sums land here
    cudaStream_t s_in = ctx->slot[0].stream, s_c = ctx->slot[1].stream, s_out = ctx->slot[2].stream;
    std::vector<int32_t> tmp;
    int rc = FDB_OK;
    cudaError_t e = cudaSuccess;
    for (int64_t b = 0; b < n_batches && rc == FDB_OK && e == cudaSuccess; ++b) {
        const int sl = (int)(b & 1);
        const int64_t r0 = b * batch;
        const int64_t nb = (n_rows - r0) < batch ? (n_rows - r0) : batch;
        // copy-in: the slot's rows were last read by the compute of batch b - 2
        if (b >= 2) e = cudaStreamWaitEvent(s_in, p->ev[sl][1], 0);
        if (e == cudaSuccess) e = cudaMemcpyAsync(p->d_x2[sl], x + r0 * FDB_F, (size_t)nb * FDB_F * 4, cudaMemcpyHostToDevice, s_in);
        const int32_t* sidx = nullptr;
        if (e == cudaSuccess && set_idx) {
            e = cudaMemcpyAsync(p->d_set2[sl], set_idx + r0, (size_t)nb * 4, cudaMemcpyHostToDevice, s_in);
            sidx = p->d_set2[sl];
        } else if (e == cudaSuccess && p->n_sets > 1 && r0 > 0) {
            // the default set mapping must see global row numbers: shift through an explicit index
            tmp.resize((size_t)nb);
            for (int64_t i = 0; i < nb; ++i) tmp[(size_t)i] = (int32_t)((r0 + i) % p->n_sets);
            e = cudaMemcpyAsync(p->d_set2[sl], tmp.data(), (size_t)nb * 4, cudaMemcpyHostToDevice, s_in);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s_in);          // tmp is reused by the next batch
            sidx = p->d_set2[sl];
        }
        if (e == cudaSuccess) e = cudaEventRecord(p->ev[sl][0], s_in);
        // evaluate + solve: needs the rows, and the slot's phi buffer back from the copy-out of batch b - 2
        if (e == cudaSuccess) e = cudaStreamWaitEvent(s_c, p->ev[sl][0], 0);
        if (e == cudaSuccess && b >= 2) e = cudaStreamWaitEvent(s_c, p->ev[sl][2], 0);
        if (e != cudaSuccess) break;
        rc = fdb_kshap_explain_dev(p, s_c, nb, p->d_x2[sl], sidx, evaluator, p->d_phi2[sl], nullptr, abs_sum ? p->d_abs + 32 * sl : nullptr);
        if (rc != FDB_OK) break;
        e = cudaEventRecord(p->ev[sl][1], s_c);
        // copy-out
        if (e == cudaSuccess) e = cudaStreamWaitEvent(s_out, p->ev[sl][1], 0);
        if (e == cudaSuccess) e = cudaMemcpyAsync(phi + r0 * FDB_F, p->d_phi2[sl], (size_t)nb * FDB_F * 4, cudaMemcpyDeviceToHost, s_out);
        if (e == cudaSuccess && abs_sum) e = cudaMemcpyAsync(h_abs + b * 32, p->d_abs + 32 * sl, 32 * sizeof(double), cudaMemcpyDeviceToHost, s_out);
        if (e == cudaSuccess) e = cudaEventRecord(p->ev[sl][2], s_out);
    }
    for (auto& s : ctx->slot) {
        cudaError_t e2 = cudaStreamSynchronize(s.stream);
        if (e == cudaSuccess) e = e2;
    }
    if (rc != FDB_OK) return rc;
    FDB_CUDA(e);
    if (abs_sum) {
        for (int64_t b = 0; b < n_batches; ++b)
            for (int k = 0; k < FDB_F; ++k) abs_sum[k] += h_abs[b * 32 + k];
        abs_sum[FDB_F] = (double)n_rows;
    }
    return FDB_OK;
}
