"""Gram-form model of the device's `l1_reg = "aic" | "bic" | alpha` feature selection (csrc/kshap_lasso.cu), in numpy.

Test infrastructure: the kernel never sees the 2n x M augmented matrix shap hands to sklearn; it works from per-set
products (Z'WZ and friends) and ONE pass over a row's coalition expectations.  This file states those identities and the
LARS-lasso walk in that form, so `tests/test_lasso_gram_cpu.py` can check them against sklearn on the CPU (the oracle
calls the real `LassoLarsIC` / `Lasso`), and a GPU failure can be told apart from a wrong identity.

Reference: shap `_kernel.py::solve` [recalled] -> `make_pipeline(StandardScaler(with_mean=False), LassoLarsIC(criterion=c))`
on `mask_aug`, `eyAdj_aug`; sklearn `_least_angle.py::_lars_path_solver(method="lasso")`, `LassoLarsIC.fit`.
"""

from __future__ import annotations

import numpy as np

EQ_TOL = float(np.finfo(np.float32).eps)
TINY32 = float(np.finfo(np.float32).tiny)
EPS = float(np.finfo(np.float64).eps)


def set_products(Z: np.ndarray, kw: np.ndarray, scale: bool = True) -> dict:
    """Everything that depends on the coalition set only (kshap_lasso_prepare): the centred (and, when `scale`, unit
    variance) Gram of the augmented matrix, its inverse, and the vectors that turn a row's sums into X'y."""
    Z = np.asarray(Z, dtype=np.float64)
    kw = np.asarray(kw, dtype=np.float64)
    n_masks, M = Z.shape
    n = 2.0 * n_masks
    s = Z.sum(1)
    Gz = Z.T @ (kw[:, None] * Z)
    B = Gz.sum(1)                      # sum_m kw s z_mj
    S0 = float(np.trace(Gz))           # sum_m kw s
    A = M * Gz - B[:, None] - B[None, :] + S0          # X_aug' X_aug
    r = np.sqrt(kw * (M - s)) + np.sqrt(kw * s)
    q = np.sqrt(kw * s)
    R = Z.T @ r
    Q = float(q.sum())
    c = R - Q                           # column sums of X_aug
    mu = c / n
    var = np.diag(A) / n - mu * mu
    sd = np.sqrt(var) if scale else np.ones(M)
    G = (A - n * np.outer(mu, mu)) / np.outer(sd, sd)
    return dict(M=M, n=n, Gz=Gz, B=B, S0=S0, c=c, sd=sd, Q=Q, r=r, G=G, Ginv=np.linalg.inv(G))


def row_products(P: dict, Z: np.ndarray, kw: np.ndarray, e: np.ndarray, d: float) -> tuple[np.ndarray, float]:
    """One pass over a row's e_m = link(ey_m) - link(fnull): centred X'y (length M) and centred y'y."""
    M, n = P["M"], P["n"]
    T = Z.T @ (kw * e)
    Tsum = float(T.sum())
    E2 = float(np.sum(kw * e * e))
    P1 = float(np.sum(P["r"] * e))
    ysum = P1 - d * P["Q"]
    xy_raw = M * T - d * P["B"] - Tsum + d * P["S0"]
    xy = (xy_raw - P["c"] * ysum / n) / P["sd"]
    yy = M * E2 - 2.0 * d * Tsum + d * d * P["S0"] - ysum * ysum / n
    return xy, yy


def lasso_path_gram(G: np.ndarray, xy: np.ndarray, yy: float, n: float, max_iter: int = 96):
    """LARS-lasso on (G, X'y): returns [(rss, nonzero mask)] per path point, as the kernel records them."""
    M = len(xy)
    cov = xy.copy()
    inactive = np.ones(M, bool)
    coef = np.zeros(M)
    active: list[int] = []
    sign: list[float] = []
    L = np.zeros((M, M))
    points = [(yy, np.zeros(M, bool))]
    n_iter, drop, prev_alpha = 0, False, np.inf
    while True:
        if inactive.any():
            cand = np.where(inactive, np.abs(cov), -1.0)
            j = int(np.argmax(cand))
            C = float(cand[j])
        else:
            j, C = -1, 0.0
        alpha = C / n
        if alpha <= EQ_TOL:
            break
        if n_iter >= max_iter or len(active) >= M:
            break
        if not drop:
            k = len(active)
            row = np.array([G[j, a] for a in active])
            w = np.linalg.solve(L[:k, :k], row) if k else np.zeros(0)      # lower triangular by construction
            diag = max(np.sqrt(abs(G[j, j] - w @ w)), EPS)
            if diag < 1e-7:
                cov[j] = 0.0
                continue
            L[k, :k] = w
            L[k, k] = diag
            active.append(j)
            sign.append(float(np.sign(cov[j])))
            inactive[j] = False
        if n_iter > 0 and prev_alpha < alpha:
            break
        k = len(active)
        sg = np.array(sign)
        ls = np.linalg.solve(L[:k, :k].T, np.linalg.solve(L[:k, :k], sg))
        if k == 1 and ls[0] == 0:
            ls[0], AA = 1.0, 1.0
        else:
            AA = 1.0 / np.sqrt(np.sum(ls * sg))
            ls = ls * AA
        corr = np.around(G[active].T @ ls, 15)
        ia = np.nonzero(inactive)[0]

        def min_pos(v):
            v = v[v > 0]
            return float(v.min()) if len(v) else np.finfo(np.float64).max

        g1 = min_pos((C - cov[ia]) / (AA - corr[ia] + TINY32))
        g2 = min_pos((C + cov[ia]) / (AA + corr[ia] + TINY32))
        gamma = min(g1, g2, C / AA)
        z = -coef[active] / (ls + TINY32)
        z_pos = min_pos(z)
        drop = False
        out = []
        if z_pos < gamma:
            out = [int(i) for i in np.nonzero(z == z_pos)[0][::-1]]
            gamma = z_pos
            drop = True
        n_iter += 1
        coef[active] += gamma * ls
        cov[ia] -= gamma * corr[ia]
        prev_alpha = alpha
        qv = G @ coef
        points.append((float(yy - 2.0 * coef @ xy + coef @ qv), np.abs(coef) > EPS))
        if drop:
            for i in out:
                f = active.pop(i)
                sign.pop(i)
                cov[f] = xy[f] - qv[f]
                coef[f] = 0.0
                inactive[f] = True
            k = len(active)
            L[:] = 0.0
            if k:
                L[:k, :k] = np.linalg.cholesky(G[np.ix_(active, active)])
    return points


def lasso_path_inv(G: np.ndarray, xy: np.ndarray, yy: float, n: float, max_iter: int = 96, lar: bool = False):
    """The device's own form of the walk (csrc/kshap_lasso.cu::lasso_path): the INVERSE of the active Gram block instead
    of its Cholesky factor.  Appending a feature is a rank-one border update, the unnormalised direction Ainv.sign
    follows without a solve, a drop is a rank-one downdate with the last position moving into the freed slot.
    `lar=False`: method "lasso", returns [(rss, nonzero mask)] like `lasso_path_gram`; `lar=True`: method "lar" stopped
    after `max_iter` steps (num_features(k)), returns the active features in order of entry."""
    M = len(xy)
    cov = xy.copy()
    inactive = np.ones(M, bool)
    coef = np.zeros(M)
    active: list[int] = []
    sign: list[float] = []
    Ainv = np.zeros((M, M))
    ls0 = np.zeros(M)
    points = [(yy, np.zeros(M, bool))]
    n_iter, drop, prev_C = 0, False, np.inf

    def min_pos(v):
        v = v[v > 0]
        return float(v.min()) if len(v) else np.finfo(np.float64).max

    while True:
        if inactive.any():
            cand = np.where(inactive, np.abs(cov), 0.0)
            j = int(np.argmax(cand))
            C = float(cand[j])
        else:
            j, C = -1, 0.0
        if C <= EQ_TOL * n:
            break
        if n_iter >= max_iter or len(active) >= M:
            break
        if not drop:
            k = len(active)
            g = np.array([G[j, a] for a in active])
            u = Ainv[:k, :k] @ g
            diag = max(np.sqrt(abs(G[j, j] - g @ u)), EPS)            # sklearn's new Cholesky pivot
            if diag < 1e-7:
                cov[j] = 0.0
                continue
            s_inv = 1.0 / (diag * diag)
            sg_new = float(np.sign(cov[j]))
            t = ((u @ np.array(sign) if k else 0.0) - sg_new) * s_inv
            Ainv[:k, :k] += np.outer(u, u) * s_inv
            Ainv[:k, k] = Ainv[k, :k] = -u * s_inv
            Ainv[k, k] = s_inv
            ls0[:k] += u * t
            ls0[k] = -t
            active.append(j)
            sign.append(sg_new)
            inactive[j] = False
        if not lar and n_iter > 0 and prev_C < C:
            break
        k = len(active)
        sg = np.array(sign)
        if k == 1 and ls0[0] == 0:
            ls, AA = np.array([1.0]), 1.0
        else:
            tmp = float(ls0[:k] @ sg)
            AA = 1.0 / np.sqrt(tmp if tmp > 0 else EPS)
            ls = ls0[:k] * AA
        corr = np.around(G[active].T @ ls, 15)
        ia = np.nonzero(inactive)[0]
        g1 = min_pos((C - cov[ia]) / (AA - corr[ia] + TINY32))
        g2 = min_pos((C + cov[ia]) / (AA + corr[ia] + TINY32))
        gamma = min(g1, g2, C / AA)
        z = -coef[active] / (ls + TINY32)
        z_pos = min_pos(z)
        drop, out = False, []
        if z_pos < gamma:
            out = [int(i) for i in np.nonzero(z == z_pos)[0][::-1]]
            if not lar:
                gamma = z_pos
            drop = True
        n_iter += 1
        coef[active] += gamma * ls
        cov[ia] -= gamma * corr[ia]
        prev_C = C
        if lar:
            if drop:                                                    # sign flip only; the direction follows it
                for i in out:
                    sign[i] = -sign[i]
                ls0[:k] = Ainv[:k, :k] @ np.array(sign)
            continue
        qv = G @ coef
        points.append((float(yy - 2.0 * coef @ xy + coef @ qv), np.abs(coef) > EPS))
        if drop:
            for i in out:
                k = len(active)
                f = active[i]
                cov[f] = xy[f] - qv[f]
                coef[f] = 0.0
                inactive[f] = True
                c = Ainv[:k, i].copy()
                Ainv[:k, :k] -= np.outer(c, c) / Ainv[i, i]
                last = k - 1
                if i != last:                                           # the last position takes the freed slot
                    Ainv[i, :k] = Ainv[last, :k]
                    Ainv[:k, i] = Ainv[:k, last]
                    Ainv[i, i] = Ainv[last, last]
                    active[i], sign[i] = active[last], sign[last]
                active.pop()
                sign.pop()
                Ainv[last, :] = 0.0
                Ainv[:, last] = 0.0
            k = len(active)
            ls0[:] = 0.0
            if k:
                ls0[:k] = Ainv[:k, :k] @ np.array(sign)
    return active if lar else points


def select_ic(P: dict, xy: np.ndarray, yy: float, criterion: str) -> np.ndarray:
    """Indices LassoLarsIC keeps: argmin over the path of n log(2 pi s2) + rss / s2 + factor * df."""
    n, M = P["n"], P["M"]
    s2 = (yy - xy @ (P["Ginv"] @ xy)) / (n - M - 1.0)
    factor = 2.0 if criterion == "aic" else np.log(n)
    pts = lasso_path_gram(P["G"], xy, yy, n)
    crit = [n * np.log(2.0 * np.pi * s2) + rss / s2 + factor * mask.sum() for rss, mask in pts]
    return np.nonzero(pts[int(np.argmin(crit))][1])[0]


def select_alpha(P: dict, xy: np.ndarray, yy: float, alpha: float, max_iter: int = 1000, tol: float = 1e-4) -> np.ndarray:
    """Indices sklearn's Lasso(alpha) keeps: cyclic coordinate descent on the Gram form from w = 0, stopping on the
    duality gap like `enet_coordinate_descent` (objective (1 / 2n) |y - Xw|^2 + alpha |w|_1, X and y centred)."""
    G, n, M = P["G"], P["n"], P["M"]
    lam = alpha * n
    w = np.zeros(M)
    q = np.zeros(M)                     # G w
    tol_abs = tol * yy
    diag = np.diag(G)
    for _ in range(max_iter):
        w_max = d_w_max = 0.0
        for j in range(M):
            if diag[j] == 0.0:
                continue
            wj = w[j]
            t = xy[j] - q[j] + diag[j] * wj
            new = np.sign(t) * max(abs(t) - lam, 0.0) / diag[j]
            if new != wj:
                q += G[:, j] * (new - wj)
                w[j] = new
            d_w_max = max(d_w_max, abs(new - wj))
            w_max = max(w_max, abs(new))
        if w_max == 0.0 or d_w_max / w_max < tol:
            xta = xy - q
            dual = float(np.max(np.abs(xta)))
            r2 = yy - 2.0 * w @ xy + w @ q
            if dual > lam:
                const = lam / dual
                gap = 0.5 * (r2 + r2 * const * const)
            else:
                const = 1.0
                gap = r2
            gap += lam * np.abs(w).sum() - const * (yy - w @ xy)
            if gap < tol_abs:
                break
    return np.nonzero(w)[0]
