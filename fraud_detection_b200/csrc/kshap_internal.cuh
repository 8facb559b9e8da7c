// Internal declarations of the Kernel SHAP part of libfdb200.so.
#pragma once

#include <vector>

#include "fdb_internal.cuh"

#define FDB_KSHAP_MAX_BG 1024          // background rows the tensor-core evaluator keeps resident in smem
#define FDB_KSHAP_BATCH_ROWS 32768     // explained rows per evaluate->solve round (bounds the ey / partial-sum scratch)

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
    uint16_t* d_aimg = nullptr;        // A operand tiles, smem image: [S][m_pad/128][16][4][8][8]
    float* d_ainv = nullptr;           // [S][29][32] (X'WX)^-1
    float* d_c2 = nullptr;             // [S][32]     X'W z29
    int* d_status = nullptr;           // [S] 1 = singular
    float* d_ey = nullptr;             // [batch_rows][m_pad] scratch between evaluate and solve
    float* d_x = nullptr;              // host-API staging
    float* d_phi = nullptr;
    int32_t* d_set = nullptr;
    double* d_abs = nullptr;
    float* d_E = nullptr;              // factored evaluator (single set): [tile][chunk][128][cbs] 2^(e - o_m)
    float* d_eoff = nullptr;           // [m_pad] exponent centres o_m
    float* d_xw = nullptr;             // [m_pad][32] kw_m * x_mj
    double* d_gz = nullptr;            // [S][30][32] Z'WZ, built on demand for l1_reg = num_features(k)
    double* d_bv = nullptr;            // [S][32]
    double* d_s0 = nullptr;            // [S]
    std::vector<double> bg_host;       // background rows (host copy, for rebuilding set-dependent operands)
    bool sampled = false;              // sets were drawn on the device and may be redrawn (fdb_kshap_plan_resample)
    int num_features = 0;              // 0 = plain WLS over all features; k = LARS selection of k features first
    uint32_t varying = 0x3FFFFFFFu;    // features that differ from the background for the rows of this plan (bit 29 always set)
    float* d_gpart = nullptr;          // [batch_rows][m_pad/32][32] partial X'W y written by K2f, folded by K3
};

void fdb_kshap_split(double v, int fmt, int n, uint16_t* out);

// K2: tcgen05 evaluator (kshap_eval_tc.cu).  ey[r][m] for rows [0, n_rows) of x.
int fdb_kshap_launch_eval_tc(fdb_kshap_plan* p, cudaStream_t st, const float* x, int64_t row_offset, int n_rows,
                             const int32_t* set_idx, float* ey);

// K2f: factored evaluator for single-set plans (kshap_eval_fx.cu).
int fdb_kshap_build_factored(fdb_kshap_plan* pl, const double* bg_host);
int fdb_kshap_launch_eval_fx(fdb_kshap_plan* pl, cudaStream_t st, const float* x, int n_rows, float* ey);

// Link applied to a float32 coalition expectation.  The logit keeps its argument inside [1e-30, 1 - 2^-24]: an
// expectation that rounds to 1 in float32 (|logit| beyond ~16.6) saturates there instead of giving inf / NaN.
#ifdef __CUDACC__
__device__ __forceinline__ float fdb_kshap_link(float v, int link) {
    if (link != 1) return v;
    v = fminf(fmaxf(v, 1e-30f), 0.99999994f);
    return __logf(v / (1.0f - v));
}
#endif

// l1_reg = "num_features(k)": LARS selection + reduced WLS (kshap_lars.cu)
int fdb_kshap_prepare_lars(fdb_kshap_plan* p);
int fdb_kshap_launch_lars_solve(fdb_kshap_plan* p, cudaStream_t st, const float* x, int64_t row_offset, int n_rows,
                                const int32_t* set_idx, float* phi, double* abs_sum, int accumulate, int64_t total_rows);

// on-device coalition sampler (kshap_sampler.cu)
int fdb_kshap_resolve_nsamples(int nsamples);
int fdb_kshap_sample_on_device(fdb_kshap_plan* pl, int nsamples, unsigned long long seed);
