// Kernel SHAP solve with shap's feature selections:  l1_reg = "num_features(k)" | "aic" | "bic" | "auto" | <alpha>.
//
// Reference (oracle/kernel_shap.py::solve, restating shap/explainers/_kernel.py::solve [recalled]):
//   X_aug = sqrt(w_aug) * [Z; Z - 1],  y_aug = sqrt(w_aug) * [e; e - d],  w_aug = [kw (M - s); kw s]      (2 n_masks rows)
//   "num_features(k)": sel = lars_path(X_aug, y_aug, max_iter=k)[1]      (method "lar"; shap's default, k = 10 [recalled])
//   "aic" / "bic":  sel = nonzero(make_pipeline(StandardScaler(with_mean=False), LassoLarsIC(criterion)).fit(X_aug, y_aug).coef_)
//   alpha        :  sel = nonzero(Lasso(alpha).fit(X_aug, y_aug).coef_)
//   "auto"       :  "aic" when nsamples / (2^M - 2) < 0.2, no selection otherwise (decided by the caller)
//   then the additivity-constrained WLS over `sel` only, every other feature gets phi = 0.
// sklearn's part (1.7.2 pinned by the reference, 1.9.0 in this image; sklearn/linear_model/_least_angle.py):
//   LassoLarsIC.fit centres X and y, walks lars_path(method="lasso") to alpha = 0, and keeps the path point that minimises
//   n log(2 pi s2) + RSS / s2 + factor * df   (factor 2 | log n,  df = #|coef| > eps,  s2 = RSS_ols / (n - M - 1));
//   Lasso.fit centres X and y and runs cyclic coordinate descent on (1 / 2n) |y - Xw|^2 + alpha |w|_1 to a duality gap.
//
// The kernel never forms the 2 n_masks x M matrix.  Per coalition set (kshap_lasso_prepare, on top of the Z'WZ products of
// kshap_lars.cu): column sums and variances of X_aug, the centred (and for the criteria unit-variance) Gram
//   G_ij = (A_ij - n mu_i mu_j) / (sd_i sd_j),   A = X_aug'X_aug = M Gz_ij - B_i - B_j + S0,   mu_j = (R_j - Q) / n,
//   R_j = sum_m (sqrt(kw (M - s)) + sqrt(kw s)) z_mj,   Q = sum_m sqrt(kw s),   sd_j^2 = A_jj / n - mu_j^2
// and its inverse.  Per explained row ONE pass over the coalition expectations gives T_j = sum kw z_mj e_m,
// E2 = sum kw e^2 and P1 = sum (sqrt(kw (M - s)) + sqrt(kw s)) e, from which
//   X'y_j = (M T_j - d B_j - sum T + d S0 - (R_j - Q) (P1 - d Q) / n) / sd_j,   y'y = M E2 - 2 d sum T + d^2 S0 - (P1 - d Q)^2 / n.
// num_features(k) needs T and d only (raw Gram, nothing centred); after the factored evaluator T comes from the records of
// its fused X'W y GEMM and no pass over the expectations is made at all.
// The LARS walk (adds, sign-change drops / flips, RSS and support recorded at every path point of the lasso variant), the
// criterion, the coordinate descent and the reduced WLS all run in float64, one warp per row, lane j = feature j
// (tests/lasso_gram_model.py is the same algorithm in numpy, checked against sklearn on the CPU).

#include <float.h>
#include <math.h>

#include "kshap_internal.cuh"

namespace {

constexpr int kMaxSel = 16;           // num_features(k): k <= 16
constexpr int kMaxPath = 96;          // path points kept per row (sklearn: max_iter = 500; 30 features need ~31 + drops)
constexpr int kVecStride = 96;        // per set: [0..31] column sums c_j, [32..63] 1 / sd_j, [64] Q, [65] n

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
__device__ __forceinline__ double shfl_xor_d(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, shfl_xor_d(v, o));
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += shfl_xor_d(v, o);
    return v;
}
__device__ __forceinline__ double min_pos(double v) { return v > 0.0 ? v : DBL_MAX; }
// Non-negative doubles order like their bit patterns: the minimum / maximum over the warp takes two integer REDUX
// instructions (high words, then low words among the lanes that hold the extreme high word) instead of a ten-shuffle
// butterfly, and it returns one lane's value bit for bit (the callers compare for equality afterwards).
__device__ __forceinline__ double warp_min_nonneg(double v) {
    const unsigned hi = (unsigned)__double2hiint(v), lo = (unsigned)__double2loint(v);
    const unsigned mh = __reduce_min_sync(0xffffffffu, hi);
    const unsigned ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
    return __hiloint2double((int)mh, (int)ml);
}
// maximum of non-negative doubles and the lowest lane that holds it
__device__ __forceinline__ double warp_argmax_nonneg(double v, int& who) {
    const unsigned hi = (unsigned)__double2hiint(v), lo = (unsigned)__double2loint(v);
    const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
    const unsigned ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
    who = __ffs(__ballot_sync(0xffffffffu, hi == mh && lo == ml)) - 1;
    return __hiloint2double((int)mh, (int)ml);
}

// lanes own coalitions and hold 32 per-feature partial sums each; afterwards lane j holds the total of feature j in v[0]
__device__ __forceinline__ void transpose_reduce(double (&v)[32], int lane) {
#pragma unroll
    for (int half = 16; half >= 1; half >>= 1) {
        const bool up = (lane & half) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            const double send = up ? v[i] : v[i + half];
            const double keep = up ? v[i + half] : v[i];
            v[i] = keep + shfl_xor_d(send, half);
        }
    }
}

// NMAX: most features ever active (30; num_features(k) stops at kMaxSel = 16); NPATH: path points recorded (criteria only)
template <int NMAX, int NPATH>
struct LassoScratchT {
    double L[NMAX][NMAX + 1];         // Cholesky factor of the active Gram block; afterwards the reduced WLS system [A | rhs]
    double ls[NMAX];                  // equiangular direction, by active position; afterwards the WLS solution
    double sgn[NMAX];
    double dinv[NMAX];                // 1 / L[a][a]
    int act[NMAX];
    double rss[NPATH + 1];
    uint32_t nz[NPATH + 1];
};

// Per coalition set: Gram of the augmented matrix - scale 1: centred, unit-variance columns (criteria); 0: centred only
// (alpha) - its inverse, column sums, 1 / sd.  (num_features(k) walks the raw Gram, read straight from the Z'WZ products.)
__global__ void __launch_bounds__(32) kshap_lasso_prepare(const uint32_t* __restrict__ bits, const float* __restrict__ kw,
                                                          int m_pad, const double* __restrict__ gz, const double* __restrict__ bv,
                                                          const double* __restrict__ s0, double* __restrict__ lg,
                                                          double* __restrict__ lginv, double* __restrict__ lvec,
                                                          uint32_t varying, int scale) {
    const int s = blockIdx.x, lane = threadIdx.x;
    const uint32_t* sb = bits + (size_t)s * m_pad;
    const float* sk = kw + (size_t)s * m_pad;
    const double Mf = (double)__popc(varying);
    const bool is_var = (varying >> lane) & 1u;
    double g[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) g[j] = 0.0;
    double qacc = 0.0, cnt = 0.0;
    for (int m = lane; m < m_pad; m += 32) {
        const double w = (double)sk[m];
        if (!(w > 0.0)) continue;
        cnt += 1.0;
        const uint32_t b = sb[m] & varying;
        const double sc = (double)__popc(b);
        const double q = sqrt(w * sc);
        const double r = sqrt(w * (Mf - sc)) + q;
        qacc += q;
#pragma unroll
        for (int j = 0; j < FDB_F; ++j) g[j] += ((b >> j) & 1u) ? r : 0.0;
    }
    transpose_reduce(g, lane);
    const double Q = warp_sum(qacc);
    const double n = 2.0 * warp_sum(cnt);
    const double* G0 = gz + (size_t)s * FDB_F * 32;
    const double Bj = bv[(size_t)s * 32 + lane];
    const double S0 = s0[s];
    const double c = is_var ? g[0] - Q : 0.0;
    const double mu = c / n;
    const double ajj = is_var ? Mf * G0[lane * 32 + lane] - 2.0 * Bj + S0 : 0.0;
    const double var = ajj / n - mu * mu;
    double isd = 1.0;
    if (scale == 1 && is_var && var > 10.0 * DBL_EPSILON) isd = 1.0 / sqrt(var);       // StandardScaler: near-constant columns keep scale 1
    // a[i] = G[i][lane]
    double a[FDB_F], inv[FDB_F];
#pragma unroll
    for (int i = 0; i < FDB_F; ++i) {
        const double mu_i = shfl_d(mu, i), isd_i = shfl_d(isd, i), B_i = shfl_d(Bj, i);
        const bool var_i = (varying >> i) & 1u;
        double v = 0.0;
        if (lane < FDB_F) {
            if (var_i && is_var) v = (Mf * G0[i * 32 + lane] - B_i - Bj + S0 - n * mu_i * mu) * isd_i * isd;
            else v = (i == lane) ? 1.0 : 0.0;                                      // features outside `varying` never enter
        }
        a[i] = v;
        inv[i] = (i == lane) ? 1.0 : 0.0;
        if (lane < FDB_F) lg[((size_t)s * FDB_F + i) * 32 + lane] = v;
    }
    // Gauss-Jordan without pivoting (G is symmetric positive definite for any set the WLS itself can use)
    bool singular = false;
#pragma unroll
    for (int k = 0; k < FDB_F; ++k) {
        const double pk = shfl_d(a[k], k);
        if (!(fabs(pk) > 1e-300)) { singular = true; break; }
        const double r = 1.0 / pk;
        const double ak = a[k] * r, ik = inv[k] * r;
#pragma unroll
        for (int i = 0; i < FDB_F; ++i) {
            const double f = shfl_d(a[i], k);
            if (i != k) {
                a[i] = fma(-f, ak, a[i]);
                inv[i] = fma(-f, ik, inv[i]);
            }
        }
        a[k] = ak;
        inv[k] = ik;
    }
    if (lane < FDB_F) {
#pragma unroll
        for (int i = 0; i < FDB_F; ++i) lginv[((size_t)s * FDB_F + i) * 32 + lane] = singular ? 0.0 : inv[i];
    }
    double* V = lvec + (size_t)s * kVecStride;
    V[lane] = c;
    V[32 + lane] = isd;
    V[64 + lane] = lane == 0 ? Q : (lane == 1 ? n : 0.0);
}

// The Gram matrix the walk reads: stored per set (criteria, alpha: centred / scaled, kshap_lasso_prepare) or, for
// num_features(k), the raw X_aug'X_aug = M Gz_ij - B_i - B_j + S0 put together on the fly from the Z'WZ products.
struct GramStored {
    const double* g;
    __device__ __forceinline__ double operator()(int i, int j) const { return g[i * 32 + j]; }
};
struct GramRaw {
    const double* gz;
    const double* bv;
    double Mf, S0;
    __device__ __forceinline__ double operator()(int i, int j) const { return Mf * gz[i * 32 + j] - bv[i] - bv[j] + S0; }
};

struct RowProducts {
    double xy;      // lane j: centred (scaled) X'y of feature j, 0 outside `varying`
    double yy;      // centred y'y
    double T;       // lane j: sum_m kw z_mj e_m
    double d;       // link(f(x)) - link(fnull)
};

// LARS walk on (G, X'y).  LAR = false (method "lasso", to alpha = 0): a coefficient that reaches zero leaves the active set;
// RSS and support of every path point are recorded in S.rss / S.nz; returns the point count - 1.  LAR = true (method
// "lar", lars_path(max_iter)): a sign change only flips the stored sign and skips the next add; stops after max_iter
// steps; returns the number of active features (S.act).
//
// sklearn keeps the Cholesky factor of the active Gram block and solves two triangular systems per path point - 2n
// dependent steps for one warp.  Here the INVERSE of the active block lives in S.L (lane i = row of active position i):
//   add feature b:  g = G[act, b],  u = Ainv g,  s = G_bb - g.u  (sqrt|s| is sklearn's new Cholesky pivot, same 1e-7
//                   degeneracy test),  Ainv <- [[Ainv + u u'/s, -u/s], [-u'/s, 1/s]]
//                   and the unnormalised direction Ainv.sign follows without a solve:  ls0 <- [ls0 + u t, -t],
//                   t = (u.sign - sign_b) / s
//   drop position k: Ainv <- Ainv - c c'/Ainv_kk without row / column k (c = column k); the last position moves into k
// - every step is a lane-parallel sweep over one row, nothing is sequential across lanes.  tests/lasso_gram_model.py
// holds both forms (sklearn's factor, this inverse); tests/test_lasso_gram_cpu.py checks both against lars_path.
template <bool LAR, class SC, class GM>
__device__ int lasso_path(SC& S, const GM& G, const RowProducts& R, double n_aug, bool is_var,
                          int n_var, int lane, int max_iter) {
    const double eps = DBL_EPSILON, tiny32 = (double)FLT_MIN, eq_tol = (double)FLT_EPSILON;
    double Cov = is_var ? R.xy : 0.0;
    bool inactive = is_var;
    double coef = 0.0;
    int my_a = -1;                           // position of this lane's feature in the active list
    int n_active = 0, n_iter = 0;
    bool drop = false;
    double prev_C = DBL_MAX;
    double ls0 = 0.0, sgv = 0.0;             // lane a: (Ainv sign)_a and sign_a of active position a
    if (!LAR && lane == 0) { S.rss[0] = R.yy; S.nz[0] = 0u; }
    // ls0 = Ainv sign from scratch (after a drop or a sign flip)
    auto direction_from_scratch = [&](int n) {
        double v = 0.0;
        if (lane < n)
            for (int a = 0; a < n; ++a) v += S.L[lane][a] * S.sgn[a];
        ls0 = v;
    };
    while (true) {
        int bidx;
        const double C = warp_argmax_nonneg(inactive ? fabs(Cov) : 0.0, bidx);      // ties: the lowest feature index
        if (C <= eq_tol * n_aug) break;                                  // alpha = C / n <= alpha_min + tolerance
        if (n_iter >= max_iter || n_active >= n_var) break;
        if (!drop) {
            const int n = n_active;
            const double g = lane < n ? G(bidx, S.act[lane]) : 0.0;
            if (lane < n) S.dinv[lane] = g;                              // dinv[] / ls[] double as broadcast buffers here
            __syncwarp();
            double u = 0.0;
            if (lane < n)
                for (int a = 0; a < n; ++a) u += S.L[lane][a] * S.dinv[a];
            const double sres = G(bidx, bidx) - warp_sum(g * u);
            const double diag = fmax(sqrt(fabs(sres)), eps);
            if (diag < 1e-7) {                                           // sklearn: degenerate regressor, zero its Cov and go on
                if (lane == bidx) Cov = 0.0;
                __syncwarp();
                continue;
            }
            const double sinv = 1.0 / (diag * diag);
            const double csig = shfl_d(Cov, bidx);
            const double sg_n = csig > 0.0 ? 1.0 : (csig < 0.0 ? -1.0 : 0.0);
            const double t = (warp_sum(u * sgv) - sg_n) * sinv;
            if (lane < n) S.ls[lane] = u;
            __syncwarp();
            if (lane < n) {
                const double us = u * sinv;
                for (int a = 0; a < n; ++a) S.L[lane][a] += us * S.ls[a];
                S.L[lane][n] = -us;
                S.L[n][lane] = -us;
                ls0 += u * t;
            } else if (lane == n) {
                S.L[n][n] = sinv;
                S.act[n] = bidx;
                S.sgn[n] = sg_n;
                ls0 = -t;
                sgv = sg_n;
            }
            if (lane == bidx) { inactive = false; coef = 0.0; my_a = n; }
            ++n_active;
            __syncwarp();
        }
        if (!LAR && n_iter > 0 && prev_C < C) break;                     // alpha (= C / n) no longer decreasing: numerical end of the path
        // equiangular direction ls = AA Ainv sign, AA = 1 / sqrt(sign . Ainv sign)
        const int n = n_active;
        double lsv, AA;
        if (n == 1 && shfl_d(ls0, 0) == 0.0) {
            lsv = lane == 0 ? 1.0 : 0.0;
            AA = 1.0;
        } else {
            double tmp = warp_sum(ls0 * sgv);
            if (!(tmp > 0.0)) tmp = eps;                                 // sklearn retries on a perturbed factor; same floor
            AA = 1.0 / sqrt(tmp);
            lsv = ls0 * AA;
        }
        if (lane < n) S.ls[lane] = lsv;
        __syncwarp();
        double corr = 0.0;
        if (inactive) {
            for (int a = 0; a < n; ++a) corr += G(S.act[a], lane) * S.ls[a];
            corr = rint(corr * 1e15) / 1e15;                             // np.around(..., decimals=15)
        }
        const double g12 = inactive ? fmin(min_pos((C - Cov) / (AA - corr + tiny32)), min_pos((C + Cov) / (AA + corr + tiny32))) : DBL_MAX;
        double gamma = fmin(warp_min_nonneg(g12), C / AA);
        const double my_ls = my_a >= 0 ? S.ls[my_a] : 0.0;
        const double zz = my_a >= 0 ? min_pos(-coef / (my_ls + tiny32)) : DBL_MAX;
        const double z_pos = warp_min_nonneg(zz);
        drop = false;
        bool leaving = false;
        if (z_pos < gamma) {
            leaving = my_a >= 0 && zz == z_pos;
            if (!LAR) gamma = z_pos;
            drop = true;
        }
        ++n_iter;
        if (my_a >= 0) coef += gamma * my_ls;
        if (inactive) Cov -= gamma * corr;
        prev_C = C;
        if (LAR) {
            if (drop) {                                                  // sign flip only; the direction has to follow it
                if (leaving) S.sgn[my_a] = -S.sgn[my_a];
                __syncwarp();
                if (lane < n) sgv = S.sgn[lane];
                direction_from_scratch(n);
            }
            __syncwarp();
            continue;
        }
        // path point n_iter: q = G coef, RSS = y'y - 2 coef.X'y + coef.q, support = |coef| > eps
        double q = 0.0;
        for (int a = 0; a < n; ++a) {
            const int f = S.act[a];
            q += G(f, lane) * shfl_d(coef, f);
        }
        const double rss = R.yy - warp_sum(coef * (2.0 * R.xy - q));
        const uint32_t nz = __ballot_sync(0xffffffffu, fabs(coef) > eps);
        if (lane == 0) { S.rss[n_iter] = rss; S.nz[n_iter] = nz; }
        if (drop) {
            // the coefficients that reached zero leave, one position at a time
            if (leaving) { Cov = R.xy - q; coef = 0.0; inactive = true; }
            uint32_t out = __ballot_sync(0xffffffffu, leaving);          // by FEATURE lane
            __syncwarp();
            while (out) {
                const int f = __ffs(out) - 1;
                out &= out - 1;
                const int k = __shfl_sync(0xffffffffu, my_a, f);         // its active position
                const int nn = n_active, last = nn - 1;
                const double pinv = 1.0 / S.L[k][k];
                if (lane < nn && lane != k) {
                    const double ck = S.L[lane][k] * pinv;
                    for (int a = 0; a < nn; ++a) S.L[lane][a] -= ck * S.L[k][a];
                }
                __syncwarp();
                if (k != last) {                                         // the last position takes the freed slot
                    if (lane == k)
                        for (int a = 0; a < nn; ++a) S.L[k][a] = S.L[last][a];
                    __syncwarp();
                    if (lane < nn) S.L[lane][k] = S.L[lane][last];
                    __syncwarp();
                    if (lane == k) S.L[k][k] = S.L[last][last];
                    const int fl = S.act[last];
                    __syncwarp();
                    if (lane == 0) { S.act[k] = fl; S.sgn[k] = S.sgn[last]; }
                    if (lane == fl) my_a = k;
                }
                if (lane == f) my_a = -1;
                n_active = last;
                __syncwarp();
            }
            sgv = lane < n_active ? S.sgn[lane] : 0.0;
            direction_from_scratch(n_active);
            if (n_active == 0) drop = false;
        }
        __syncwarp();
    }
    return LAR ? n_active : n_iter;
}

// Lasso(alpha) support: cyclic coordinate descent on the Gram form from w = 0 with sklearn's stopping rule
// (largest update / largest coefficient < tol, then duality gap < tol * y'y; tol = 1e-4, max_iter = 1000).
__device__ uint32_t lasso_cd_support(const double* __restrict__ G, const RowProducts& R, double lam, bool is_var, int lane) {
    const double tol = 1e-4;
    const double gjj = lane < FDB_F ? G[lane * 32 + lane] : 1.0;
    double w = 0.0, q = 0.0;                 // lane j: w_j and (G w)_j
    const bool ok = is_var && gjj != 0.0;
    for (int it = 0; it < 1000; ++it) {
        // One cyclic sweep.  A coordinate whose update returns the value it already has changes nothing (sklearn skips
        // the residual update too), so the sweep jumps from one CHANGING coordinate to the next: every lane evaluates
        // its own update from its current (G w)_j, the lowest changing lane at or after the cursor is applied, repeat.
        // The iterates are those of the coordinate-by-coordinate loop; a sweep costs (changes + 1) steps instead of 30.
        double d_w_max = 0.0;
        int pos = 0;
        while (true) {
            const double t = R.xy - q + gjj * w;
            const double mag = fabs(t) - lam;
            const double nw = (ok && mag > 0.0) ? copysign(mag, t) / gjj : 0.0;
            const unsigned m = __ballot_sync(0xffffffffu, ok && lane >= pos && nw != w);
            if (m == 0u) break;
            const int j = __ffs(m) - 1;
            const double dw = shfl_d(nw - w, j);
            if (lane < FDB_F) q += G[j * 32 + lane] * dw;
            if (lane == j) w = nw;
            d_w_max = fmax(d_w_max, fabs(dw));
            pos = j + 1;
        }
        const double w_max = warp_max(fabs(w));
        if (w_max == 0.0 || d_w_max / w_max < tol || it == 999) {
            const double dual = warp_max(is_var ? fabs(R.xy - q) : 0.0);
            const double wxy = warp_sum(w * R.xy), wq = warp_sum(w * q), l1 = warp_sum(fabs(w));
            const double r2 = R.yy - 2.0 * wxy + wq;
            double cst = 1.0, gap = r2;
            if (dual > lam) { cst = lam / dual; gap = 0.5 * (r2 + r2 * cst * cst); }
            gap += lam * l1 - cst * (R.yy - wxy);
            if (gap < tol * R.yy) break;
        }
    }
    return __ballot_sync(0xffffffffu, w != 0.0) & 0x3FFFFFFFu;
}

struct Coef30 { double w[FDB_F]; };

// MODE 0: information criterion (factor = 2 for AIC, log n for BIC, chosen by `bic`); MODE 1: Lasso(alpha);
// MODE 2: num_features(kfeat), LAR walk on the raw Gram - it needs T only, which the factored evaluator's fused X'W y GEMM
// has already left in `gpart` (column j < 29: T_j - T_29, column 29: T_29) when that evaluator ran: no pass over ey then.
template <int MODE>
__global__ void __launch_bounds__(128) kshap_lasso_solve(
    const float* __restrict__ x, int64_t row_offset, int n_rows, const int32_t* __restrict__ set_idx, int n_sets,
    const uint32_t* __restrict__ bits, const float* __restrict__ kw, int m_pad, const float* __restrict__ ey,
    const double* __restrict__ gz, const double* __restrict__ bv, const double* __restrict__ s0v,
    const double* __restrict__ lg, const double* __restrict__ lginv, const double* __restrict__ lvec,
    const Coef30 wc, double intercept, double fnull, int link, int bic, double alpha_l1,
    float* __restrict__ phi, double* __restrict__ abs_sum, int accumulate, int64_t total_rows,
    double* __restrict__ partials, unsigned int* __restrict__ ticket, uint32_t varying, int kfeat,
    const float* __restrict__ gpart, int gslots, long long fx_units, int fx_grid, int fx_tiles) {
    using Scratch = LassoScratchT<MODE == 2 ? kMaxSel : FDB_F, MODE == 0 ? kMaxPath : 0>;
    __shared__ Scratch scr[4];
    __shared__ float s_tab[32];                                   // sqrt(M - s) + sqrt(s) by coalition size
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const bool is_var = (varying >> lane) & 1u;
    const int n_var = __popc(varying);
    const double Mf = (double)n_var;
    if (threadIdx.x < 32) s_tab[lane] = lane <= n_var ? (float)(sqrt(Mf - (double)lane) + sqrt((double)lane)) : 0.f;
    __syncthreads();
    Scratch& S = scr[warp];
    const double lnull = link == 1 ? log(fnull / (1.0 - fnull)) : fnull;
    double wk = 0.0;
#pragma unroll
    for (int j = 0; j < FDB_F; ++j)
        if (j == lane) wk = wc.w[j];
    double abs_acc = 0.0;

    for (int r = blockIdx.x * 4 + warp; r < n_rows; r += gridDim.x * 4) {
        const int s = set_idx ? set_idx[r] : (n_sets == 1 ? 0 : (int)((row_offset + r) % n_sets));
        const double zl = warp_sum(lane < FDB_F ? wk * (double)x[(size_t)r * FDB_F + lane] : 0.0) + intercept;
        RowProducts R;
        R.d = (link == 1 ? zl : 1.0 / (1.0 + exp(-zl))) - lnull;
        // ---- one pass over the row's coalition expectations: T_j, E2 = sum kw e^2, P1 = sum r_m e  (float64 sums)
        const uint32_t* sb = bits + (size_t)s * m_pad;
        const float* sk = kw + (size_t)s * m_pad;
        const float* er = ey + (size_t)r * m_pad;
        double e2 = 0.0, p1 = 0.0;
        if (MODE == 2 && gpart != nullptr) {
            // records of the CTAs that worked on this row's block of 128 rows, folded in CTA order (as K3 does)
            const long long ub = (long long)(r >> 7) * fx_tiles;
            const int nseg = (int)((((ub + fx_tiles) * fx_grid - 1) / fx_units) - (((ub + 1) * fx_grid - 1) / fx_units)) + 1;
            const float* gp = gpart + (size_t)r * gslots * 32 + lane;
            float g = 0.f;
            for (int q = 0; q < nseg; ++q) g += gp[(size_t)q * 32];
            const double t29 = shfl_d((double)g, FDB_F - 1);
            R.T = !is_var ? 0.0 : (lane < FDB_F - 1 ? (double)g + t29 : t29);
        } else if (MODE == 2) {
            // num_features(k) needs T only; float32 sums (as K3's own X'W y pass), a third of the float64 pass's time
            const float lnull_f = (float)lnull;
            float gj[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) gj[j] = 0.f;
            for (int m0 = 0; m0 < m_pad; m0 += 32) {
                const uint32_t b = sb[m0 + lane] & varying;
                const float kwv = sk[m0 + lane];
                const float v = kwv != 0.f ? kwv * (fdb_kshap_link(er[m0 + lane], link) - lnull_f) : 0.f;
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
            R.T = is_var ? (double)gj[0] : 0.0;
        } else {
            double gj[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) gj[j] = 0.0;
            for (int m0 = 0; m0 < m_pad; m0 += 32) {
                const uint32_t b = sb[m0 + lane] & varying;
                const float kwv = sk[m0 + lane];
                if (kwv != 0.f) {                                   // lanes diverge only on the padding of the last words
                    const double e = (double)fdb_kshap_link(er[m0 + lane], link) - lnull;
                    const double v = (double)kwv * e;
                    e2 = fma(v, e, e2);
                    p1 = fma((double)(sqrtf(kwv) * s_tab[__popc(b)]), e, p1);
#pragma unroll
                    for (int j = 0; j < FDB_F; ++j) gj[j] += ((b >> j) & 1u) ? v : 0.0;
                }
            }
            transpose_reduce(gj, lane);
            R.T = is_var ? gj[0] : 0.0;
        }
        const double E2 = warp_sum(e2), P1 = warp_sum(p1), Tsum = warp_sum(R.T);
        const double* Bs = bv + (size_t)s * 32;
        const double Bj = lane < FDB_F ? Bs[lane] : 0.0, S0 = s0v[s];
        const double* G = nullptr;
        double n_aug;
        if (MODE == 2) {                                            // raw system: nothing centred, nothing scaled
            n_aug = Bs[31];
            R.xy = is_var ? Mf * R.T - R.d * Bj - Tsum + R.d * S0 : 0.0;
            R.yy = 0.0;
        } else {
            G = lg + (size_t)s * FDB_F * 32;
            const double* V = lvec + (size_t)s * kVecStride;
            const double cj = V[lane], isd = V[32 + lane], Q = V[64];
            n_aug = V[65];
            const double ysum = P1 - R.d * Q;
            R.xy = is_var ? (Mf * R.T - R.d * Bj - Tsum + R.d * S0 - cj * ysum / n_aug) * isd : 0.0;
            R.yy = Mf * E2 - 2.0 * R.d * Tsum + R.d * R.d * S0 - ysum * ysum / n_aug;
        }

        uint32_t sel;
        if (MODE == 0) {
            // noise variance of LassoLarsIC: OLS residual / (n - M - 1), RSS_ols = y'y - X'y . G^-1 X'y
            const double* Gi = lginv + (size_t)s * FDB_F * 32;
            double u = 0.0;
            for (int i = 0; i < FDB_F; ++i) u += (lane < FDB_F ? Gi[i * 32 + lane] : 0.0) * shfl_d(R.xy, i);
            const double sigma2 = (R.yy - warp_sum(R.xy * u)) / (n_aug - Mf - 1.0);
            const int n_pts = lasso_path<false>(S, GramStored{G}, R, n_aug, is_var, n_var, lane, kMaxPath);
            __syncwarp();
            const double factor = bic ? log(n_aug) : 2.0;
            double best = DBL_MAX;
            int bk = n_pts;                                         // no usable noise estimate: keep the end of the path
            if (sigma2 > 0.0) {
                for (int k = lane; k <= n_pts; k += 32) {
                    const double c = S.rss[k] / sigma2 + factor * (double)__popc(S.nz[k]);
                    if (c < best) { best = c; bk = k; }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const double ob = shfl_xor_d(best, o);
                    const int ok = __shfl_xor_sync(0xffffffffu, bk, o);
                    if (ob < best || (ob == best && ok < bk)) { best = ob; bk = ok; }
                }
            }
            sel = S.nz[bk] & varying;
        } else if (MODE == 1) {
            sel = lasso_cd_support(G, R, alpha_l1 * n_aug, is_var, lane) & varying;
        } else {
            const int n_act = lasso_path<true>(S, GramRaw{gz + (size_t)s * FDB_F * 32, Bs, Mf, S0}, R, n_aug, is_var, n_var, lane, kfeat);
            __syncwarp();
            sel = 0u;
            for (int a = 0; a < n_act; ++a) sel |= 1u << S.act[a];
            sel &= varying;
        }
        __syncwarp();

        // ---- additivity-constrained WLS over the selected features (ascending; the last one is eliminated)
        const int n_sel = __popc(sel);
        double p = 0.0;
        if (n_sel == 1) {
            if ((sel >> lane) & 1u) p = R.d;
        } else if (n_sel > 1) {
            const int kk = n_sel - 1;
            const int l = 31 - __clz(sel);
            const double* Gz = gz + (size_t)s * FDB_F * 32;
            const int pos = __popc(sel & ((1u << lane) - 1u));       // position of feature `lane` in the selection
            if (((sel >> lane) & 1u) && lane != l) S.act[pos] = lane;
            __syncwarp();
            const double Tl = shfl_d(R.T, l);
            double rhs = 0.0;
            if (lane < kk) {
                const int ia = S.act[lane];
                for (int b = 0; b < kk; ++b) {
                    const int ib = S.act[b];
                    S.L[lane][b] = Gz[ia * 32 + ib] - Gz[ia * 32 + l] - Gz[l * 32 + ib] + Gz[l * 32 + l];
                }
            }
            for (int a = 0; a < kk; ++a) {
                const double t = shfl_d(R.T, S.act[a]);
                if (lane == a) rhs = t - Tl - R.d * (Gz[S.act[a] * 32 + l] - Gz[l * 32 + l]);
            }
            __syncwarp();
            // elimination without pivoting (X'WX is symmetric positive definite), lane = row
            for (int c = 0; c < kk; ++c) {
                const double piv = S.L[c][c];
                const double rc = shfl_d(rhs, c);
                if (lane > c && lane < kk) {
                    const double f = S.L[lane][c] / piv;
                    for (int q = c + 1; q < kk; ++q) S.L[lane][q] -= f * S.L[c][q];
                    rhs -= f * rc;
                }
                __syncwarp();
            }
            double sol = 0.0;
            for (int c = kk - 1; c >= 0; --c) {
                const double xc = shfl_d(rhs, c) / S.L[c][c];
                if (lane == c) sol = xc;
                if (lane < c) rhs -= S.L[lane][c] * xc;
            }
            if (lane < kk) S.ls[lane] = sol;
            const double sumw = warp_sum(lane < kk ? sol : 0.0);
            __syncwarp();
            if ((sel >> lane) & 1u) p = lane == l ? R.d - sumw : S.ls[pos];
            __syncwarp();
        }
        if (fabs(p) < 1e-10) p = 0.0;
        if (lane < FDB_F) phi[(size_t)r * FDB_F + lane] = (float)p;
        abs_acc += fabs((double)(float)p);          // of the value as stored, so that a recount over phi reproduces the sums
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

static int launch_prepare(fdb_kshap_plan* p, cudaStream_t st) {
    kshap_lasso_prepare<<<p->n_sets, 32, 0, st>>>(p->d_bits, p->d_kw, p->m_pad, p->d_gz, p->d_bv, p->d_s0, p->d_lg, p->d_lginv,
                                                   p->d_lvec, p->varying, p->lasso_scaled);
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}

// scale = 1: unit-variance columns (the StandardScaler step in front of LassoLarsIC); 0: plain centring (Lasso(alpha))
int fdb_kshap_prepare_lasso(fdb_kshap_plan* p, int scale) {
    int rc = fdb_kshap_prepare_lars(p);                       // Z'WZ, B, S0
    if (rc != FDB_OK) return rc;
    if (p->d_lg != nullptr && p->lasso_scaled == scale) return FDB_OK;
    if (p->d_lg == nullptr) {
        FDB_CUDA(cudaMalloc((void**)&p->d_lg, (size_t)p->n_sets * FDB_F * 32 * 8));
        FDB_CUDA(cudaMalloc((void**)&p->d_lginv, (size_t)p->n_sets * FDB_F * 32 * 8));
        FDB_CUDA(cudaMalloc((void**)&p->d_lvec, (size_t)p->n_sets * kVecStride * 8));
    }
    p->lasso_scaled = scale;
    rc = launch_prepare(p, 0);
    if (rc != FDB_OK) return rc;
    FDB_CUDA(cudaDeviceSynchronize());
    return FDB_OK;
}

// after a resample (stream-ordered, after fdb_kshap_refresh_lars on the same stream)
int fdb_kshap_refresh_lasso(fdb_kshap_plan* p, cudaStream_t st) {
    if (p->d_lg == nullptr) return FDB_OK;
    return launch_prepare(p, st);
}

void fdb_kshap_free_lasso(fdb_kshap_plan* p) {
    cudaFree(p->d_lg); cudaFree(p->d_lginv); cudaFree(p->d_lvec);
    p->d_lg = nullptr; p->d_lginv = nullptr; p->d_lvec = nullptr;
    p->lasso_scaled = -1;
}

int fdb_kshap_launch_lasso_solve(fdb_kshap_plan* p, cudaStream_t st, const float* x, int64_t row_offset, int n_rows,
                                 const int32_t* set_idx, float* phi, double* abs_sum, int accumulate, int64_t total_rows,
                                 const float* gpart, int gslots, long long fx_units, int fx_grid) {
    int grid = (n_rows + 3) / 4;
    const int cap = p->ctx->sm_count * 8;
    if (grid > cap) grid = cap;
    Coef30 wc;
    for (int j = 0; j < FDB_F; ++j) wc.w[j] = p->coef[j];
    const int tiles = p->m_pad / 128;
#define FDB_LASSO_ARGS(bic, alpha, kfeat, gp)                                                                                   \
    x, row_offset, n_rows, set_idx, p->n_sets, p->d_bits, p->d_kw, p->m_pad, p->d_ey, p->d_gz, p->d_bv, p->d_s0, p->d_lg,       \
        p->d_lginv, p->d_lvec, wc, p->intercept, p->fnull, p->link, bic, alpha, phi, abs_sum, accumulate, total_rows,           \
        p->d_partials, p->d_ticket, p->varying, kfeat, gp, gslots, fx_units, fx_grid, tiles
    if (p->l1_mode == FDB_KSHAP_L1_ALPHA) {
        FDB_REQUIRE(p->lasso_scaled == 0, FDB_ERR_STATE, "selection products were prepared for another l1 mode");
        kshap_lasso_solve<1><<<grid, 128, 0, st>>>(FDB_LASSO_ARGS(0, p->l1_alpha, 0, nullptr));
    } else if (p->l1_mode == FDB_KSHAP_L1_NUM_FEATURES) {
        kshap_lasso_solve<2><<<grid, 128, 0, st>>>(FDB_LASSO_ARGS(0, 0.0, p->num_features, gpart));
    } else {
        FDB_REQUIRE(p->lasso_scaled == 1, FDB_ERR_STATE, "selection products were prepared for another l1 mode");
        kshap_lasso_solve<0><<<grid, 128, 0, st>>>(FDB_LASSO_ARGS(p->l1_mode == FDB_KSHAP_L1_BIC ? 1 : 0, 0.0, 0, nullptr));
    }
#undef FDB_LASSO_ARGS
    FDB_CUDA(cudaGetLastError());
    return FDB_OK;
}
