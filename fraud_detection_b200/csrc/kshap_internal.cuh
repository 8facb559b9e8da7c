// Internal declarations of the Kernel SHAP part of libfdb200.so.
#pragma once

#include <vector>

#include <cuda_fp16.h>

#include "fdb_internal.cuh"

#define FDB_KSHAP_MAX_BG 1024          // background rows the tensor-core evaluator keeps resident in smem
#define FDB_KSHAP_BATCH_ROWS 32768     // explained rows per evaluate->solve round when coalition expectations go through HBM (bounds the ey scratch)
#define FDB_KSHAP_FX_BATCH_ROWS 262144 // ... when the factored evaluator hands K3 the finished X'W y instead (256 B of scratch per row)
#define FDB_FX_TILE_BYTES 9216         // per coalition tile of a single-set plan: x^T image (8192 B) + o_m[128] + kw_m[128]

struct fdb_kshap_plan {
    fdb_ctx* ctx = nullptr;
    int n_bg = 0, nb_pad = 0;          // background rows, padded to 16
    int n_masks = 0, m_pad = 0;        // coalitions per set, padded to 128
    int n_sets = 0;
    int link = 0;                      // 0 identity, 1 logit
    int fmt = 0, nsplit = 2;           // tcgen05 operand format: 0 = f16 x 2 splits, 1 = bf16 x 3 splits
    double coef[FDB_F] = {0};
    double intercept = 0.0;
    double fnull = 0.0;                // mean_b f(bg_b)
    int64_t batch_rows = 0;
    // device state
    uint32_t* d_bits = nullptr;        // [S][m_pad] coalition bit masks (bit k = feature k present)
    float* d_kw = nullptr;             // [S][m_pad] kernel weights (0 for padding)
    float* d_P = nullptr;              // [B][32] w_k * bg_bk, column 30 = intercept (SIMT evaluator)
    float* d_wq = nullptr;             // [32] w_k
    uint16_t* d_bimg = nullptr;        // B operand, smem image: [split][nb_pad/8][4][8][8]
    uint16_t* d_aimg = nullptr;        // single-set plans: coalition tile images [m_pad/128][16][4][8][8] for K2f (K2 expands d_bits itself)
    float* d_ainv = nullptr;           // [S][29][32] (X'WX)^-1
    float* d_c2 = nullptr;             // [S][32]     X'W z29
    int* d_status = nullptr;           // [S] 1 = singular
    float* d_ey = nullptr;             // [batch_rows][m_pad] scratch between evaluate and solve
    float* d_x2[2] = {nullptr, nullptr};     // host-API staging, two slots: copy-in, evaluate + solve and copy-out of
    float* d_phi2[2] = {nullptr, nullptr};   //   consecutive batches overlap (fdb_kshap_explain_host)
    int32_t* d_set2[2] = {nullptr, nullptr};
    cudaEvent_t ev[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};   // per slot: rows in, phi ready, phi out
    double* d_abs = nullptr;                 // [2][32]
    float* d_E = nullptr;              // factored evaluator (single set): [tile][chunk][128][cbs] 2^(e - o_m)
    float* d_eoff = nullptr;           // [m_pad] exponent centres o_m
    float fx_quad_logsum = 0.f;        // worst sum over a column quad of |log2 E| on one side of 0 (bounds the product of four denominators)
    float fx_half_spread = 0.f;        // largest half-spread (log2) of a coalition's background exponents around o_m
    unsigned char* d_fxtile = nullptr; // factored evaluator: per tile, x^T image (B operand of the fused X'W y GEMM) + o_m + kw_m
    double* d_gz = nullptr;            // [S][30][32] Z'WZ, built on demand for l1_reg = num_features(k)
    double* d_bv = nullptr;            // [S][32] B_j = sum kw s z_mj; entry 31 = rows of the augmented system (2 x coalitions with weight)
    double* d_s0 = nullptr;            // [S]
    std::vector<double> bg_host;       // background rows (host copy, for rebuilding set-dependent operands)
    bool sampled = false;              // sets were drawn on the device and may be redrawn (fdb_kshap_plan_resample)
    int num_features = 0;              // 0 = plain WLS over all features; k = LARS selection of k features first
    int l1_mode = 0;                   // FDB_KSHAP_L1_*: which feature selection runs before the WLS
    double l1_alpha = 0.0;             // FDB_KSHAP_L1_ALPHA: the Lasso penalty
    double* d_lg = nullptr;            // [S][30][32] centred (scaled) Gram of the augmented system (kshap_lasso.cu)
    double* d_lginv = nullptr;         // [S][30][32] its inverse
    double* d_lvec = nullptr;          // [S][96] column sums, 1 / sd, Q, n
    int lasso_scaled = -1;             // which variant d_lg holds: 1 unit-variance columns (criteria), 0 centred only (alpha)
    uint32_t varying = 0x3FFFFFFFu;    // features that differ from the background for the rows of this plan (bit 29 always set)
    float* d_gpart = nullptr;          // [gpart_rows][slots][32] X'W y records written by K2f, folded by K3 (grown on demand)
    size_t ey_cap = 0, gpart_cap = 0;  // floats the two scratch buffers hold
    int64_t stage_rows = 0;            // rows the host-API staging buffers hold
    void* sampler_state = nullptr;     // device-sampled plans: the uploaded sampling schedule (kshap_sampler.cu)
    int* d_bad = nullptr;              // number of coalition sets whose X'WX came out singular since the last check (resample_dev)
    bool gz_valid = true;              // d_gz / d_bv / d_s0 match the current sets (a redraw with no selection active skips them)
    bool ainv_valid = true;            // d_ainv / d_c2 match the current sets (a stream-ordered redraw leaves them stale; the
                                       //   next K3 launch rebuilds them first, the feature-selection solves never need them)
    double* d_partials = nullptr;      // [FDB_MAX_GRID][32] per-CTA |phi| column sums of K3 / K3L (the plan's own: plans on
    unsigned int* d_ticket = nullptr;  //   different streams, and K1 calls on the same context, never share an arrival ticket)
};

void fdb_kshap_split(double v, int fmt, int n, uint16_t* out);

// K2: tcgen05 evaluator (kshap_eval_tc.cu).  ey[r][m] for rows [0, n_rows) of x.
int fdb_kshap_launch_eval_tc(fdb_kshap_plan* p, cudaStream_t st, const float* x, int64_t row_offset, int n_rows,
                             const int32_t* set_idx, float* ey);

size_t fdb_kshap_tc_smem_bytes(const fdb_kshap_plan* pl);

// K2f: factored evaluator for single-set plans (kshap_eval_fx.cu).
int fdb_kshap_build_factored(fdb_kshap_plan* pl, const double* bg_host);
int fdb_kshap_launch_eval_fx(fdb_kshap_plan* pl, cudaStream_t st, const float* x, int n_rows, float* ey, float* gpart);
void fdb_kshap_fx_geometry(const fdb_kshap_plan* pl, int n_rows, long long* n_units, int* grid, int* slots);

// Coalition expectations between evaluate and solve (d_ey).  Identity link: ey itself.  Logit link: an expectation
// above 1/2 would lose its distance to 1 in float32 (1 - ey < 6e-8 already breaks logit(ey)), so the evaluators
// sum the sigmoids AND their complements and store  ey  when ey <= 1/2,  -(1 - ey)  otherwise: either way the
// small quantity keeps its relative accuracy and logit(ey) is exact to float32 in both tails.
#ifdef __CUDACC__
__device__ __forceinline__ float fdb_kshap_encode_ey(float sum_sigma, float sum_complement, float inv_bg) {
    return sum_sigma <= sum_complement ? sum_sigma * inv_bg : -(sum_complement * inv_bg);
}
__device__ __forceinline__ float fdb_kshap_decode_ey(float v) { return v < 0.f ? 1.0f + v : v; }
// link(ey) from the stored value.  A plain (unsigned) expectation that rounded to 1 - possible only from the clamped
// evaluator variants - saturates at 1 - 2^-24 instead of giving inf / NaN.
__device__ __forceinline__ float fdb_kshap_link(float v, int link) {
    if (link != 1) return v;
    if (v < 0.f) {
        const float c = fmaxf(-v, 1e-37f);            // 1 - ey
        return __logf((1.0f - c) / c);
    }
    v = fminf(fmaxf(v, 1e-37f), 0.99999994f);
    return __logf(v / (1.0f - v));
}
#endif

// Row `lane` of G = sum_m w_m v_m v_m' for one coalition set, v_m = the low NF bits of word(bits_m) (one warp).
// Also c = sum_m w_m v_m,lane flag(bits_m).  Float64 result.  Coalitions of a sampled set that were drawn once all
// carry the same weight (the smallest positive one): those are counted with bit-sliced integer arithmetic -
// T_j = ballot(bit j) over 32 coalitions, count_ij += popc(T_i & T_j) - and enter as w * count (exact in float64);
// every other weight takes the plain predicated float64 accumulation.  Any weights are handled; the split only
// decides the speed (29 DADD per coalition against ~5 integer instructions per feature per 32 coalitions).
#ifdef __CUDACC__
template <int NF, typename WordFn, typename FlagFn>
__device__ __forceinline__ void fdb_kshap_weighted_gram_row(const uint32_t* __restrict__ sb, const float* __restrict__ sk,
                                                            int m_pad, int lane, WordFn word, FlagFn flag,
                                                            double (&a)[NF], double& c) {
    const uint32_t full = 0xffffffffu;
    float wmin = 3.0e38f;
    for (int m = lane; m < m_pad; m += 32) {
        const float w = sk[m];
        if (w > 0.f) wmin = fminf(wmin, w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmin = fminf(wmin, __shfl_xor_sync(full, wmin, o));
    uint32_t cnt[NF];
#pragma unroll
    for (int j = 0; j < NF; ++j) { a[j] = 0.0; cnt[j] = 0u; }
    uint32_t cflag = 0u;
    c = 0.0;
    for (int m0 = 0; m0 < m_pad; m0 += 32) {
        const float w = sk[m0 + lane];
        const uint32_t b = sb[m0 + lane];
        const uint32_t v = word(b) & ((1u << NF) - 1u);
        const bool fl = flag(b);
        const bool fast = (w == wmin);
        const uint32_t vf = fast ? v : 0u;
        uint32_t t[NF];
        uint32_t mine = 0u;
#pragma unroll
        for (int j = 0; j < NF; ++j) {
            t[j] = __ballot_sync(full, (vf >> j) & 1u);
            if (lane == j) mine = t[j];
        }
#pragma unroll
        for (int j = 0; j < NF; ++j) cnt[j] += __popc(mine & t[j]);
        cflag += __popc(mine & __ballot_sync(full, fast && fl));
        uint32_t slow = __ballot_sync(full, !fast && w != 0.f);
        while (slow) {
            const int src = __ffs(slow) - 1;
            slow &= slow - 1u;
            const double ws = (double)__shfl_sync(full, w, src);
            const uint32_t vs = __shfl_sync(full, v, src);
            const bool fs = __shfl_sync(full, (int)fl, src) != 0;
            if ((vs >> lane) & 1u) {
#pragma unroll
                for (int j = 0; j < NF; ++j)
                    if ((vs >> j) & 1u) a[j] += ws;
                if (fs) c += ws;
            }
        }
    }
    const double wd = (double)wmin;
#pragma unroll
    for (int j = 0; j < NF; ++j) a[j] += wd * (double)cnt[j];
    c += wd * (double)cflag;
}
#endif

// The same Gram row for sets whose weights are DRAW COUNTS - every weight an integer multiple (<= 2048) of the smallest
// one, except one class of equal weights among the first 64 coalitions (the enumerated block of a sampled set): what
// the on-device sampler and shap's own sampler produce.  Then  G = w0 E + wmin R  with integer matrices
// E = sum_{enumerated} v v',  R = sum_m count_m v_m v_m',  and both are small GEMMs  V' diag(count) V  that run on the
// tensor cores (mma.sync m16n8k16, f16 operands holding 0 / 1 / count - exact - and fp32 accumulation of integers below
// 2^24 - exact).  A thread builds its operand fragments straight from four coalition words per k-step: after one shift
// the bits of its four feature rows sit in the four bytes of a word, a byte permute pairs them with the neighbouring
// coalition, one multiply turns the pair into two halves.  ~4.4 instructions per coalition for the warp against ~14 for
// the bit-sliced counting plus the predicated float64 path (which stays as the fallback for arbitrary weights: the
// function returns false and leaves a / c untouched).  The flag column rides along as feature NF.
// tile: 2 x 32 x 33 floats of shared memory owned by this warp.
#ifdef __CUDACC__
__device__ __forceinline__ void fdb_mma_m16n8k16(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                                 uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <int NF, typename WordFn, typename FlagFn>
__device__ __forceinline__ bool fdb_kshap_weighted_gram_row_mma(const uint32_t* __restrict__ sb, const float* __restrict__ sk,
                                                                int m_pad, int lane, WordFn word, FlagFn flag, float* tile,
                                                                double (&a)[NF], double& c) {
    static_assert(NF < 31, "the flag column needs a bit of its own");
    const uint32_t full = 0xffffffffu;
    float wmin = 3.0e38f;
#pragma unroll 4
    for (int m = lane; m < m_pad; m += 32) {
        const float w = __ldg(sk + m);
        if (w > 0.f) wmin = fminf(wmin, w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmin = fminf(wmin, __shfl_xor_sync(full, wmin, o));
    if (!(wmin < 3.0e38f)) return false;
    const float w0 = __ldg(sk);
    const bool has_enum = w0 > 0.f && w0 != wmin && m_pad >= 64;
    const float inv_wmin = 1.0f / wmin;
    const int g = lane >> 2, t = lane & 3;
    float acc[2][4][4];
    float worst = 0.f, largest = 0.f;       // how far a weight is from a whole multiple of wmin, and the largest multiple seen
    auto zero_acc = [&]() {
#pragma unroll
        for (int r = 0; r < 2; ++r)
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[r][q][i] = 0.f;
    };
    // one k-step: 16 coalitions (this thread's four: words bw, weights ww) starting at m0; `enumerated` selects which
    // weights count (see above)
    auto kstep = [&](int m0, bool enumerated, bool maybe_enum, const uint32_t (&bw)[4], const float (&ww)[4]) {
        uint32_t mb[4];
        float cnt[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t v = (word(bw[i]) & ((1u << NF) - 1u)) | (flag(bw[i]) ? (1u << NF) : 0u);
            mb[i] = (v >> g) & 0x01010101u;                  // features g, g + 8, g + 16, g + 24 in bytes 0..3
            const bool in_enum = maybe_enum && has_enum && m0 < 64 && ww[i] == w0;       // only the first 64 coalitions can be
            if (enumerated) {
                cnt[i] = in_enum ? 1.0f : 0.0f;
            } else {
                const float q = ww[i] * inv_wmin, rq = rintf(q);
                cnt[i] = in_enum ? 0.0f : rq;
                worst = fmaxf(worst, in_enum ? 0.f : fabsf(q - rq));       // judged once, after the loop
                largest = fmaxf(largest, cnt[i]);
            }
        }
        const __half2 hab2 = __floats2half2_rn(cnt[0], cnt[1]), hcd2 = __floats2half2_rn(cnt[2], cnt[3]);
        const uint32_t hab = *reinterpret_cast<const uint32_t*>(&hab2), hcd = *reinterpret_cast<const uint32_t*>(&hcd2);
        uint32_t aab[4], acd[4], bab4[4], bcd4[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const uint32_t sel = (uint32_t)r | ((uint32_t)r << 4) | ((uint32_t)(4 + r) << 8) | ((uint32_t)(4 + r) << 12);
            const uint32_t sab = __byte_perm(mb[0], mb[1], sel) & 0x00010001u;       // bit of coalition a | bit of coalition b << 16
            const uint32_t scd = __byte_perm(mb[2], mb[3], sel) & 0x00010001u;
            aab[r] = sab * 0x3C00u;                                                  // two halves, 0 or 1.0
            acd[r] = scd * 0x3C00u;
            bab4[r] = (sab * 0xFFFFu) & hab;                                         // two halves, 0 or count
            bcd4[r] = (scd * 0xFFFFu) & hcd;
        }
#pragma unroll
        for (int rt = 0; rt < 2; ++rt)
#pragma unroll
            for (int q = 0; q < 4; ++q)
                fdb_mma_m16n8k16(acc[rt][q], aab[2 * rt], aab[2 * rt + 1], acd[2 * rt], acd[2 * rt + 1], bab4[q], bcd4[q]);
    };
    auto load4 = [&](int m0, uint32_t (&bw)[4], float (&ww)[4]) {
        const uint2 bab = __ldg(reinterpret_cast<const uint2*>(sb + m0 + 2 * t));
        const uint2 bcd = __ldg(reinterpret_cast<const uint2*>(sb + m0 + 2 * t + 8));
        const float2 wab = __ldg(reinterpret_cast<const float2*>(sk + m0 + 2 * t));
        const float2 wcd = __ldg(reinterpret_cast<const float2*>(sk + m0 + 2 * t + 8));
        bw[0] = bab.x; bw[1] = bab.y; bw[2] = bcd.x; bw[3] = bcd.y;
        ww[0] = wab.x; ww[1] = wab.y; ww[2] = wcd.x; ww[3] = wcd.y;
    };
    // accumulators -> this warp's shared-memory tile (row = feature i, column = feature j)
    auto dump = [&](float* tl) {
#pragma unroll
        for (int rt = 0; rt < 2; ++rt)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                tl[(16 * rt + g) * 33 + 8 * q + 2 * t] = acc[rt][q][0];
                tl[(16 * rt + g) * 33 + 8 * q + 2 * t + 1] = acc[rt][q][1];
                tl[(16 * rt + g + 8) * 33 + 8 * q + 2 * t] = acc[rt][q][2];
                tl[(16 * rt + g + 8) * 33 + 8 * q + 2 * t + 1] = acc[rt][q][3];
            }
    };
    zero_acc();
    if (has_enum)
        for (int m0 = 0; m0 < 64; m0 += 16) {
            uint32_t bw[4];
            float ww[4];
            load4(m0, bw, ww);
            kstep(m0, true, true, bw, ww);
        }
    dump(tile);
    zero_acc();
    // 64 coalitions per trip (m_pad is a multiple of 128): the loads of four k-steps are issued together, ahead of the
    // fragment arithmetic, so a warp keeps 16 loads in flight instead of stalling on four per k-step
    {
        uint32_t bw[4][4];
        float ww[4][4];
#pragma unroll
        for (int q = 0; q < 4; ++q) load4(16 * q, bw[q], ww[q]);
#pragma unroll
        for (int q = 0; q < 4; ++q) kstep(16 * q, false, true, bw[q], ww[q]);
    }
    for (int m1 = 64; m1 < m_pad; m1 += 64) {                  // beyond the enumerated block: no per-coalition membership test
        uint32_t bw[4][4];
        float ww[4][4];
#pragma unroll
        for (int q = 0; q < 4; ++q) load4(m1 + 16 * q, bw[q], ww[q]);
#pragma unroll
        for (int q = 0; q < 4; ++q) kstep(m1 + 16 * q, false, false, bw[q], ww[q]);
    }
    dump(tile + 32 * 33);
    __syncwarp();
    // draw counts: every weight a whole multiple of the smallest (float32 rounding of count * scale stays below 3e-4 for
    // counts <= 2048, which the half-precision operands hold exactly); anything else takes the generic path
    if (__any_sync(full, worst > 0.01f || largest > 2048.f)) return false;
    const double dw0 = has_enum ? (double)w0 : 0.0, dwm = (double)wmin;
#pragma unroll
    for (int j = 0; j < NF; ++j) a[j] = dw0 * (double)tile[lane * 33 + j] + dwm * (double)tile[(32 + lane) * 33 + j];
    c = dw0 * (double)tile[lane * 33 + NF] + dwm * (double)tile[(32 + lane) * 33 + NF];
    return true;
}
#endif

// l1_reg = "num_features(k)": LARS selection + reduced WLS (kshap_lars.cu)
int fdb_kshap_prepare_lars(fdb_kshap_plan* p);
int fdb_kshap_refresh_lars(fdb_kshap_plan* p, cudaStream_t st);
// l1_reg = "aic" | "bic" | alpha: LARS-lasso path / coordinate descent selection + reduced WLS (kshap_lasso.cu)
int fdb_kshap_prepare_lasso(fdb_kshap_plan* p, int scale);
int fdb_kshap_refresh_lasso(fdb_kshap_plan* p, cudaStream_t st);
void fdb_kshap_free_lasso(fdb_kshap_plan* p);
int fdb_kshap_launch_lasso_solve(fdb_kshap_plan* p, cudaStream_t st, const float* x, int64_t row_offset, int n_rows,
                                 const int32_t* set_idx, float* phi, double* abs_sum, int accumulate, int64_t total_rows,
                                 const float* gpart, int gslots, long long fx_units, int fx_grid);
// on-device coalition sampler (kshap_sampler.cu)
int fdb_kshap_resolve_nsamples(int nsamples);
int fdb_kshap_sample_on_device(fdb_kshap_plan* pl, cudaStream_t stream, int nsamples, unsigned long long seed);
void fdb_kshap_sampler_free(fdb_kshap_plan* pl);
