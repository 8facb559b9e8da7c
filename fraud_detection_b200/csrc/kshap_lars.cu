// Kernel SHAP solve with shap's feature-selecting regularisation  l1_reg = "num_features(k)"
// (what an unmodified KernelExplainer.shap_values(X) call runs [recalled default]).
//
// Reference (oracle/kernel_shap.py::solve, restating shap/explainers/_kernel.py::solve):
//   w_aug   = [kw (M - s); kw s]                     s_m = number of present features of coalition m, M = number of varying features (30 unless the plan says otherwise)
//   X_aug   = sqrt(w_aug) * [Z; Z - 1],  y_aug = sqrt(w_aug) * [e; e - d]          e = link(ey) - link(fnull)
//   sel     = sklearn.linear_model.lars_path(X_aug, y_aug, max_iter=k)[1]          (method "lar", Gram form)
//   l       = sel[-1];  X = Z[:, sel[:-1]] - Z[:, l];  y = e - Z[:, l] d
//   w       = solve(X'WX, X'W y);  phi[sel[:-1]] = w;  phi[l] = d - sum(w);  other features 0
//
// Everything the row contributes enters through ONE 30-vector  T_j = sum_m kw_m z_mj e_m  and the scalar d:
//   X_aug' y_aug   = 30 T_j - d B_j - sum_i T_i + d S0          B_j = sum_m kw s z_mj,  S0 = sum_m kw s
//   X_aug' X_aug   = 30 Gz_ij - B_i - B_j + S0                   Gz  = Z' W Z     (per coalition set)
//   (z_i - z_l)' W (z_j - z_l) = Gz_ij - Gz_il - Gz_lj + Gz_ll ;  (z_i - z_l)' W y = T_i - T_l - d (Gz_il - Gz_ll)
// so the per-set matrices are built once (kshap_gram_prepare) and each explained row is one warp:
// T from ey (lanes own coalitions, butterfly transpose-reduce), then the LARS iteration and the reduced
// solve in float64 with lane j = feature j.  The selection is discrete, hence float64 and sklearn's exact
// step rules (argmax |Cov|, Cholesky update with the 1e-7 degeneracy test, equiangular direction, the two
// min_pos step candidates, rounding of corr_eq_dir to 15 decimals, the "drop" flag that skips the next add).

#include <float.h>
#include <math.h>

#include "kshap_internal.cuh"

namespace {

constexpr int kMaxSel = 16;   // num_features(k) supported up to k = 16 (shap's recalled default is 10)

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
__device__ __forceinline__ double shfl_xor_d(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }

__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, shfl_xor_d(v, o));
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += shfl_xor_d(v, o);
    return v;
}

struct WarpScratch {
    double L[kMaxSel][kMaxSel + 1];   // Cholesky factor of the active Gram block (row-major, lower)
    double ls[kMaxSel];               // equiangular direction
    double sgn[kMaxSel];
    double mat[kMaxSel][kMaxSel + 1]; // reduced WLS system [A | rhs]
    int act[kMaxSel];
};

// Per coalition set: Gz = Z'WZ (30 x 30, row stride 32), B_j = sum kw s z_j, S0 = sum kw s.  One warp per set.
__global__ void __launch_bounds__(32) kshap_gram_prepare(const uint32_t* __restrict__ bits, const float* __restrict__ kw,
                                                         int m_pad, double* __restrict__ gz, double* __restrict__ bv,
                                                         double* __restrict__ s0, uint32_t varying) {
    const int s = blockIdx.x, lane = threadIdx.x;
    const uint32_t* sb = bits + (size_t)s * m_pad;
    const float* sk = kw + (size_t)s * m_pad;
    // B_i = sum_m kw s_m z_mi = sum_j Gz_ij and S0 = sum_m kw s_m = trace(Gz), because s_m = sum_j z_mj
    double a[FDB_F];
    double unused;
    fdb_kshap_weighted_gram_row<FDB_F>(
        sb, sk, m_pad, lane, [varying](uint32_t z) { return z & varying; }, [](uint32_t) { return false; }, a, unused);
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
    bv[(size_t)s * 32 + lane] = lane < FDB_F ? b : 0.0;
    if (lane == 0) s0[s] = tot;
}


__device__ __forceinline__ double min_pos(double v) { return v > 0.0 ? v : DBL_MAX; }

__global__ void __launch_bounds__(128) kshap_lars_solve(
    const float* __restrict__ x, int64_t row_offset, int n_rows, const int32_t* __restrict__ set_idx, int n_sets,
    const uint32_t* __restrict__ bits, const float* __restrict__ kw, int m_pad, const float* __restrict__ ey,
    const double* __restrict__ gz, const double* __restrict__ bv, const double* __restrict__ s0v,
    const float* __restrict__ wcoef, double intercept, double fnull, int link, int kfeat, double n_samples_aug,
    float* __restrict__ phi, double* __restrict__ abs_sum, int accumulate, int64_t total_rows,
    double* __restrict__ partials, unsigned int* __restrict__ ticket, uint32_t varying) {
    __shared__ WarpScratch scr[4];
    const bool is_var = (varying >> (threadIdx.x & 31)) & 1u;   // lane j = feature j takes part in the regression
    const int n_var = __popc(varying);
    const double Mf = (double)n_var;                            // the reference's M (number of varying features)
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    WarpScratch& S = scr[warp];
    const double lnull = link == 1 ? log(fnull / (1.0 - fnull)) : fnull;
    const float lnull_f = (float)lnull;
    const double wk = lane < FDB_F ? (double)wcoef[lane] : 0.0;
    const double eps = DBL_EPSILON, tiny32 = (double)FLT_MIN, eq_tol = (double)FLT_EPSILON;
    double abs_acc = 0.0;

    for (int r = blockIdx.x * 4 + warp; r < n_rows; r += gridDim.x * 4) {
        const int s = set_idx ? set_idx[r] : (n_sets == 1 ? 0 : (int)((row_offset + r) % n_sets));
        const double zl = warp_sum(lane < FDB_F ? wk * (double)x[(size_t)r * FDB_F + lane] : 0.0) + intercept;
        const double d = (link == 1 ? zl : 1.0 / (1.0 + exp(-zl))) - lnull;
        // ---- T_j = sum_m kw_m z_mj e_m : lanes own coalitions, then transpose-reduce so lane j holds T_j
        const uint32_t* sb = bits + (size_t)s * m_pad;
        const float* sk = kw + (size_t)s * m_pad;
        const float* er = ey + (size_t)r * m_pad;
        float gj[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) gj[j] = 0.f;
        for (int m0 = 0; m0 < m_pad; m0 += 32) {
            const uint32_t b = sb[m0 + lane];
            const float kwv = sk[m0 + lane];
            const float ev = er[m0 + lane];
            const float v = kwv != 0.f ? kwv * (fdb_kshap_link(ev, link) - lnull_f) : 0.f;
#pragma unroll
            for (int j = 0; j < FDB_F; ++j) gj[j] += ((b >> j) & 1u) ? v : 0.f;
        }
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
        const double T = is_var ? (double)gj[0] : 0.0;
        const double* G = gz + (size_t)s * FDB_F * 32;
        const double* Bs = bv + (size_t)s * 32;
        const double Bj = Bs[lane];
        const double S0 = s0v[s];
        const double Tsum = warp_sum(T);
        auto gram = [&](int i, int j) { return Mf * G[i * 32 + j] - Bs[i] - Bs[j] + S0; };

        // ---- LARS (method "lar", Gram form) over the augmented system; lane j = feature j
        double Cov = is_var ? Mf * T - d * Bj - Tsum + d * S0 : 0.0;
        bool inactive = is_var;
        double coef = 0.0;
        int n_active = 0, n_iter = 0;
        bool drop = false;
        while (true) {
            double best = inactive ? fabs(Cov) : -1.0;
            int bidx = lane;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double ob = shfl_xor_d(best, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
                if (ob > best || (ob == best && oi < bidx)) { best = ob; bidx = oi; }
            }
            const double C = best >= 0.0 ? best : 0.0;
            if (C / n_samples_aug <= eq_tol) break;                      // alpha <= alpha_min + tol
            if (n_iter >= kfeat || n_active >= n_var || n_active >= kMaxSel) break;
            if (!drop) {
                const double Csigned = shfl_d(Cov, bidx);
                bool degenerate = false;
                if (lane == 0) {
                    const int n = n_active;
                    const double c = gram(bidx, bidx);
                    double v = 0.0;
                    for (int a = 0; a < n; ++a) {
                        double t = gram(bidx, S.act[a]);
                        for (int b = 0; b < a; ++b) t -= S.L[n][b] * S.L[a][b];
                        t /= S.L[a][a];
                        S.L[n][a] = t;
                        v += t * t;
                    }
                    const double diag = fmax(sqrt(fabs(c - v)), eps);
                    S.L[n][n] = diag;
                    degenerate = diag < 1e-7;
                    if (!degenerate) {
                        S.act[n] = bidx;
                        S.sgn[n] = Csigned > 0.0 ? 1.0 : (Csigned < 0.0 ? -1.0 : 0.0);
                    }
                }
                degenerate = __shfl_sync(0xffffffffu, (int)degenerate, 0) != 0;
                __syncwarp();
                if (degenerate) {                                        // sklearn: zero its Cov and try the next one
                    if (lane == bidx) Cov = 0.0;
                    continue;
                }
                if (lane == bidx) { inactive = false; coef = 0.0; }
                ++n_active;
            }
            // equiangular direction: solve (L L') ls = sign, AA = 1 / sqrt(ls . sign)
            double AA = 0.0;
            if (lane == 0) {
                const int n = n_active;
                double doff = 0.0;
                for (int attempt = 0;; ++attempt) {
                    for (int a = 0; a < n; ++a) {
                        double t = S.sgn[a];
                        for (int b = 0; b < a; ++b) t -= S.L[a][b] * S.ls[b];
                        S.ls[a] = t / (S.L[a][a] + doff);
                    }
                    for (int a = n - 1; a >= 0; --a) {
                        double t = S.ls[a];
                        for (int b = a + 1; b < n; ++b) t -= S.L[b][a] * S.ls[b];
                        S.ls[a] = t / (S.L[a][a] + doff);
                    }
                    if (attempt == 0 && n == 1 && S.ls[0] == 0.0) { S.ls[0] = 1.0; AA = 1.0; break; }
                    double tmp = 0.0;
                    for (int a = 0; a < n; ++a) tmp += S.ls[a] * S.sgn[a];
                    if (attempt > 0) tmp = fmax(tmp, eps);
                    AA = 1.0 / sqrt(tmp);
                    if (isfinite(AA) || attempt > 60) {
                        for (int a = 0; a < n; ++a) S.ls[a] *= AA;
                        break;
                    }
                    doff += ldexp(eps, attempt);                          // L_.diag += 2^i eps, cumulative
                }
            }
            AA = shfl_d(AA, 0);
            __syncwarp();
            double corr = 0.0;
            if (inactive) {
                for (int a = 0; a < n_active; ++a) corr += gram(S.act[a], lane) * S.ls[a];
                corr = rint(corr * 1e15) / 1e15;                          // np.around(..., decimals=15)
            }
            const double g1 = warp_min(inactive ? min_pos((C - Cov) / (AA - corr + tiny32)) : DBL_MAX);
            const double g2 = warp_min(inactive ? min_pos((C + Cov) / (AA + corr + tiny32)) : DBL_MAX);
            const double gamma = fmin(fmin(g1, g2), C / AA);
            // sign-change test on the active coefficients (sets the flag that skips the next add)
            double zz = DBL_MAX;
            int my_a = -1;
            for (int a = 0; a < n_active; ++a)
                if (S.act[a] == lane) my_a = a;
            if (my_a >= 0) zz = min_pos(-coef / (S.ls[my_a] + tiny32));
            const double z_pos = warp_min(zz);
            drop = false;
            if (z_pos < gamma) {
                if (my_a >= 0 && zz == z_pos) S.sgn[my_a] = -S.sgn[my_a];
                drop = true;
            }
            ++n_iter;
            if (my_a >= 0) coef += gamma * S.ls[my_a];
            if (inactive) Cov -= gamma * corr;
            __syncwarp();
        }

        // ---- reduced, additivity-constrained WLS over the selected features (order of entry; last one eliminated)
        double p = 0.0;
        if (n_active == 1) {
            if (lane == S.act[0]) p = d;
        } else if (n_active > 1) {
            const int kk = n_active - 1;
            const int l = S.act[kk];
            const double Tl = shfl_d(T, l);
            double Ti = 0.0;
            for (int a = 0; a < kk; ++a) {
                const double t = shfl_d(T, S.act[a]);
                if (lane == a) Ti = t;
            }
            if (lane < kk) {
                const int ia = S.act[lane];
                for (int b = 0; b < kk; ++b) {
                    const int ib = S.act[b];
                    S.mat[lane][b] = G[ia * 32 + ib] - G[ia * 32 + l] - G[l * 32 + ib] + G[l * 32 + l];
                }
                S.mat[lane][kk] = Ti - Tl - d * (G[ia * 32 + l] - G[l * 32 + l]);
            }
            __syncwarp();
            if (lane == 0) {                                             // Gaussian elimination, partial pivoting
                for (int c = 0; c < kk; ++c) {
                    int piv = c;
                    for (int rr = c + 1; rr < kk; ++rr)
                        if (fabs(S.mat[rr][c]) > fabs(S.mat[piv][c])) piv = rr;
                    if (piv != c)
                        for (int q = c; q <= kk; ++q) { const double t = S.mat[c][q]; S.mat[c][q] = S.mat[piv][q]; S.mat[piv][q] = t; }
                    const double inv = 1.0 / S.mat[c][c];
                    for (int rr = c + 1; rr < kk; ++rr) {
                        const double f = S.mat[rr][c] * inv;
                        for (int q = c; q <= kk; ++q) S.mat[rr][q] -= f * S.mat[c][q];
                    }
                }
                for (int c = kk - 1; c >= 0; --c) {
                    double t = S.mat[c][kk];
                    for (int q = c + 1; q < kk; ++q) t -= S.mat[c][q] * S.ls[q];
                    S.ls[c] = t / S.mat[c][c];                            // reuse ls[] for the solution
                }
            }
            __syncwarp();
            double sumw = 0.0;
            for (int a = 0; a < kk; ++a) {
                sumw += S.ls[a];
                if (lane == S.act[a]) p = S.ls[a];
            }
            if (lane == l) p = d - sumw;
            __syncwarp();
        }
        if (fabs(p) < 1e-10) p = 0.0;
        if (lane < FDB_F) phi[(size_t)r * FDB_F + lane] = (float)p;
        abs_acc += fabs(p);
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

}  // namespace

int fdb_kshap_prepare_lars(fdb_kshap_plan* p) {
    if (p->d_gz != nullptr) return FDB_OK;
    FDB_CUDA(cudaMalloc((void**)&p->d_gz, (size_t)p->n_sets * FDB_F * 32 * 8));
    FDB_CUDA(cudaMalloc((void**)&p->d_bv, (size_t)p->n_sets * 32 * 8));
    FDB_CUDA(cudaMalloc((void**)&p->d_s0, (size_t)p->n_sets * 8));
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

int fdb_kshap_launch_lars_solve(fdb_kshap_plan* p, cudaStream_t st, const float* x, int64_t row_offset, int n_rows,
                                const int32_t* set_idx, float* phi, double* abs_sum, int accumulate, int64_t total_rows) {
    int grid = (n_rows + 3) / 4;
    const int cap = p->ctx->sm_count * 8;
    if (grid > cap) grid = cap;
    kshap_lars_solve<<<grid, 128, 0, st>>>(x, row_offset, n_rows, set_idx, p->n_sets, p->d_bits, p->d_kw, p->m_pad, p->d_ey,
                                          p->d_gz, p->d_bv, p->d_s0, p->d_wq, p->intercept, p->fnull, p->link,
                                          p->num_features, 2.0 * (double)p->n_masks, phi, abs_sum, accumulate, total_rows,
                                          p->d_partials, p->d_ticket, p->varying);
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}
