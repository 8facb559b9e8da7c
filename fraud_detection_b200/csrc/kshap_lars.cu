// Per-set products of shap's feature-selecting regularisations (l1_reg = "num_features(k)" | "aic" | "bic" | alpha).
//
// Reference (oracle/kernel_shap.py::solve, restating shap/explainers/_kernel.py::solve):
//   w_aug   = [kw (M - s); kw s]                     s_m = number of present features of coalition m, M = number of varying features (30 unless the plan says otherwise)
//   X_aug   = sqrt(w_aug) * [Z; Z - 1],  y_aug = sqrt(w_aug) * [e; e - d]          e = link(ey) - link(fnull)
//   sel     = a selection on (X_aug, y_aug): lars_path(max_iter=k), LassoLarsIC, Lasso   (kshap_lasso.cu)
//   l       = sel[-1];  X = Z[:, sel[:-1]] - Z[:, l];  y = e - Z[:, l] d
//   w       = solve(X'WX, X'W y);  phi[sel[:-1]] = w;  phi[l] = d - sum(w);  other features 0
//
// Everything the row contributes enters through ONE 30-vector  T_j = sum_m kw_m z_mj e_m  and the scalar d:
//   X_aug' y_aug   = M T_j - d B_j - sum_i T_i + d S0           B_j = sum_m kw s z_mj,  S0 = sum_m kw s
//   X_aug' X_aug   = M Gz_ij - B_i - B_j + S0                    Gz  = Z' W Z     (per coalition set)
//   (z_i - z_l)' W (z_j - z_l) = Gz_ij - Gz_il - Gz_lj + Gz_ll ;  (z_i - z_l)' W y = T_i - T_l - d (Gz_il - Gz_ll)
// This file builds Gz, B and S0 once per coalition set (kshap_gram_prepare); the selections and the reduced solve, one
// warp per explained row in float64, are in kshap_lasso.cu.

#include <float.h>
#include <math.h>

#include "kshap_internal.cuh"

namespace {

// Per coalition set: Gz = Z'WZ (30 x 30, row stride 32), B_j = sum kw s z_j, S0 = sum kw s.  One warp per set.
__global__ void __launch_bounds__(32, 16) kshap_gram_prepare(const uint32_t* __restrict__ bits, const float* __restrict__ kw,
                                                         int m_pad, double* __restrict__ gz, double* __restrict__ bv,
                                                         double* __restrict__ s0, uint32_t varying) {
    const int s = blockIdx.x, lane = threadIdx.x;
    const uint32_t* sb = bits + (size_t)s * m_pad;
    const float* sk = kw + (size_t)s * m_pad;
    // B_i = sum_m kw s_m z_mi = sum_j Gz_ij and S0 = sum_m kw s_m = trace(Gz), because s_m = sum_j z_mj
    double a[FDB_F];
    double unused;
    __shared__ float s_tile[2 * 32 * 33];
    auto word = [varying](uint32_t z) { return z & varying; };
    auto flag = [](uint32_t) { return false; };
    // draw-count weights (sampled sets, shap's own sets) take the tensor-core Gram, anything else the exact generic path
    if (!fdb_kshap_weighted_gram_row_mma<FDB_F>(sb, sk, m_pad, lane, word, flag, s_tile, a, unused))
        fdb_kshap_weighted_gram_row<FDB_F>(sb, sk, m_pad, lane, word, flag, a, unused);
    double b = 0.0;
#pragma unroll
    for (int j = 0; j < FDB_F; ++j) b += a[j];
    double diag = 0.0;
#pragma unroll
    for (int j = 0; j < FDB_F; ++j)
        if (j == lane) diag = a[j];
    if (lane >= FDB_F) diag = 0.0;
    double tot = diag;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    if (lane < FDB_F) {
#pragma unroll
        for (int j = 0; j < FDB_F; ++j) gz[((size_t)s * FDB_F + lane) * 32 + j] = a[j];
        gz[((size_t)s * FDB_F + lane) * 32 + 30] = 0.0;
        gz[((size_t)s * FDB_F + lane) * 32 + 31] = 0.0;
    }
    // rows of the augmented system: two per coalition that carries weight (padding and merged duplicates carry none)
    int cnt = 0;
    for (int m = lane; m < m_pad; m += 32) cnt += sk[m] > 0.f ? 1 : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    bv[(size_t)s * 32 + lane] = lane < FDB_F ? b : (lane == 31 ? 2.0 * (double)cnt : 0.0);
    if (lane == 0) s0[s] = tot;
}


}  // namespace

int fdb_kshap_prepare_lars(fdb_kshap_plan* p) {
    if (p->d_gz != nullptr && p->gz_valid) return FDB_OK;
    if (p->d_gz == nullptr) {
        FDB_CUDA(cudaMalloc((void**)&p->d_gz, (size_t)p->n_sets * FDB_F * 32 * 8));
        FDB_CUDA(cudaMalloc((void**)&p->d_bv, (size_t)p->n_sets * 32 * 8));
        FDB_CUDA(cudaMalloc((void**)&p->d_s0, (size_t)p->n_sets * 8));
    }
    p->gz_valid = true;
    kshap_gram_prepare<<<p->n_sets, 32>>>(p->d_bits, p->d_kw, p->m_pad, p->d_gz, p->d_bv, p->d_s0, p->varying);
    FDB_CUDA(cudaGetLastError());
    FDB_CUDA(cudaDeviceSynchronize());
    return FDB_OK;
}

// after a resample: the same products for the new sets, stream-ordered, buffers kept
int fdb_kshap_refresh_lars(fdb_kshap_plan* p, cudaStream_t st) {
    if (p->d_gz == nullptr) return FDB_OK;
    kshap_gram_prepare<<<p->n_sets, 32, 0, st>>>(p->d_bits, p->d_kw, p->m_pad, p->d_gz, p->d_bv, p->d_s0, p->varying);
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}
