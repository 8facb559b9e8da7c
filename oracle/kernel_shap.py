"""Kernel SHAP oracle (test infrastructure).  PARITY UNPINNED - see oracle/__init__.py.

CPU restatement of ``shap.KernelExplainer`` (``shap==0.49.1``, pinned at the reference's
``requirements.txt:21``; the reference itself holds no KernelExplainer call site).  The
package source is not available in this image, so every step below is [recalled] from
upstream ``shap/explainers/_kernel.py`` (functions ``explain``, ``addsample``, ``run``,
``solve``) and follows SURVEY.md section 8(a) row a9, steps 1-8.  It deliberately keeps
the reference's shape: float64, tiled synthetic data, one model call per explained row,
Python loops - it is the checker and the timed CPU baseline, not a fast path.

The pieces are exposed separately so the GPU path can be fed the *identical*
coalition masks and kernel weights:

    sample_coalitions(M, nsamples)            -> (Z, kw)          steps 2-5
    evaluate(f, x, background, Z)             -> ey               step 6
    solve(Z, kw, ey, fx, fnull, link, l1_reg) -> phi              steps 7-8
    explain_row / shap_values                 -> the composition

Known-answer anchors (tests/test_oracle_kernel_shap.py): identity-link Kernel SHAP
of a linear function equals (x - mean(bg)) * coef for any full-rank mask set; full
enumeration on 2-4 feature toys equals brute-force Shapley values; additivity;
the schedule constants of SURVEY.md 8(a) step 4 (60 enumerated masks of weight
4.352066e-3 and weight_left 0.738876 at M=30, nsamples 2048/4096).
"""

from __future__ import annotations

import copy
import itertools
import math
import warnings

import numpy as np
from scipy.special import binom


# ----------------------------------------------------------------------------- links
def _identity(x):
    return x


def _logit(x):
    return np.log(x / (1.0 - x))


LINKS = {"identity": _identity, "logit": _logit}


# ------------------------------------------------------------------- steps 2-5: masks
def size_weights(M: int) -> tuple[np.ndarray, int, int]:
    """Step 3: Shapley-kernel weight of each coalition size 1..ceil((M-1)/2)."""
    num_subset_sizes = int(np.ceil((M - 1) / 2.0))
    num_paired_subset_sizes = int(np.floor((M - 1) / 2.0))
    weight_vector = np.array([(M - 1.0) / (i * (M - i)) for i in range(1, num_subset_sizes + 1)])
    weight_vector[:num_paired_subset_sizes] *= 2
    weight_vector /= np.sum(weight_vector)
    return weight_vector, num_subset_sizes, num_paired_subset_sizes


def resolve_nsamples(M: int, nsamples="auto") -> tuple[int, int]:
    """Step 2: 'auto' -> 2M + 2048, capped at 2^M - 2 when M <= 30."""
    if nsamples == "auto":
        nsamples = 2 * M + 2 ** 11
    max_samples = 2 ** 30
    if M <= 30:
        max_samples = 2 ** M - 2
        if nsamples > max_samples:
            nsamples = max_samples
    return int(nsamples), int(max_samples)


def sample_coalitions(M: int, nsamples="auto", random=None) -> tuple[np.ndarray, np.ndarray, dict]:
    """Steps 2-5.  Returns (maskMatrix (n, M) float64 of 0/1, kernelWeights (n,), info).

    ``random`` is the module/object supplying ``choice`` and ``permutation``; the
    reference draws from the GLOBAL ``np.random`` state (so successive rows get
    different masks), which is the default here too.
    """
    random = np.random if random is None else random
    nsamples, max_samples = resolve_nsamples(M, nsamples)
    masks: list[np.ndarray] = []
    weights: list[float] = []

    def addsample(m, w):
        masks.append(m.copy())
        weights.append(float(w))

    weight_vector, num_subset_sizes, num_paired_subset_sizes = size_weights(M)

    # step 4: enumerate whole coalition sizes while the budget allows
    num_full_subsets = 0
    num_samples_left = nsamples
    group_inds = np.arange(M, dtype="int64")
    mask = np.zeros(M)
    remaining_weight_vector = copy.copy(weight_vector)
    for subset_size in range(1, num_subset_sizes + 1):
        nsubsets = binom(M, subset_size)
        if subset_size <= num_paired_subset_sizes:
            nsubsets *= 2
        if num_samples_left * remaining_weight_vector[subset_size - 1] / nsubsets >= 1.0 - 1e-8:
            num_full_subsets += 1
            num_samples_left -= nsubsets
            if remaining_weight_vector[subset_size - 1] < 1.0:
                remaining_weight_vector /= (1 - remaining_weight_vector[subset_size - 1])
            w = weight_vector[subset_size - 1] / binom(M, subset_size)
            if subset_size <= num_paired_subset_sizes:
                w /= 2.0
            for inds in itertools.combinations(group_inds, subset_size):
                mask[:] = 0.0
                mask[np.array(inds, dtype="int64")] = 1.0
                addsample(mask, w)
                if subset_size <= num_paired_subset_sizes:
                    mask[:] = np.abs(mask - 1)
                    addsample(mask, w)
        else:
            break

    # step 5: random paired draws over the sizes that were not enumerated
    nfixed_samples = len(masks)
    samples_left = nsamples - len(masks)
    weight_left = 0.0
    if num_full_subsets != num_subset_sizes:
        remaining_weight_vector = copy.copy(weight_vector)
        remaining_weight_vector[:num_paired_subset_sizes] /= 2
        remaining_weight_vector = remaining_weight_vector[num_full_subsets:]
        remaining_weight_vector /= np.sum(remaining_weight_vector)
        ind_set = random.choice(len(remaining_weight_vector), 4 * samples_left, p=remaining_weight_vector)
        ind_set_pos = 0
        used_masks: dict[tuple, int] = {}
        while samples_left > 0 and ind_set_pos < len(ind_set):
            mask.fill(0.0)
            ind = ind_set[ind_set_pos]
            ind_set_pos += 1
            subset_size = ind + num_full_subsets + 1
            mask[random.permutation(M)[:subset_size]] = 1.0
            mask_tuple = tuple(mask)
            new_sample = False
            if mask_tuple not in used_masks:
                new_sample = True
                used_masks[mask_tuple] = len(masks)
                samples_left -= 1
                addsample(mask, 1.0)
            else:
                weights[used_masks[mask_tuple]] += 1.0
            if samples_left > 0 and subset_size <= num_paired_subset_sizes:
                mask[:] = np.abs(mask - 1)
                if new_sample:
                    samples_left -= 1
                    addsample(mask, 1.0)
                else:
                    weights[used_masks[mask_tuple] + 1] += 1.0
        weight_left = float(np.sum(weight_vector[num_full_subsets:]))
        kw = np.array(weights)
        kw[nfixed_samples:] *= weight_left / kw[nfixed_samples:].sum()
    else:
        kw = np.array(weights)

    Z = np.array(masks).reshape(len(masks), M)
    info = {"nsamples": nsamples, "max_samples": max_samples, "num_full_subsets": num_full_subsets,
            "nfixed_samples": nfixed_samples, "weight_left": weight_left,
            "weight_vector": weight_vector}
    return Z, kw, info


# ------------------------------------------------------------------ step 6: evaluate
def evaluate(f, x: np.ndarray, background: np.ndarray, Z: np.ndarray,
             varying: np.ndarray | None = None) -> np.ndarray:
    """ey[m] = mean_b f(x*z_m + b*(1-z_m)) by tiling the background (addsample + run)."""
    background = np.asarray(background, dtype=np.float64)
    x = np.asarray(x, dtype=np.float64).reshape(-1)
    N = background.shape[0]
    n = Z.shape[0]
    varying = np.arange(background.shape[1]) if varying is None else np.asarray(varying)
    synth = np.tile(background, (n, 1))
    for m in range(n):
        groups = varying[Z[m] == 1.0]
        synth[m * N:(m + 1) * N, groups] = x[groups]
    y = np.asarray(f(synth), dtype=np.float64).reshape(n, N)
    bg_w = np.ones(N) / N
    ey = np.zeros(n)
    for j in range(N):
        ey += y[:, j] * bg_w[j]
    return ey


# ------------------------------------------------------------------ steps 7-8: solve
def solve(Z: np.ndarray, kw: np.ndarray, ey: np.ndarray, fx: float, fnull: float,
          link: str = "identity", l1_reg=False, fraction_evaluated: float = 0.0,
          ic_standardize: bool = True) -> np.ndarray:
    """Constrained weighted least squares of ``KernelExplainer.solve``.

    ``l1_reg`` [recalled]: ``"num_features(k)"`` -> ``lars_path(max_iter=k)``; ``"aic"`` / ``"bic"`` ->
    ``make_pipeline(StandardScaler(with_mean=False), LassoLarsIC(criterion=c))`` (the scaler step replaced the
    ``normalize=False`` argument sklearn 1.2 deprecated; ``ic_standardize=False`` restates the older plain
    ``LassoLarsIC``); ``"auto"`` -> ``"aic"`` when ``fraction_evaluated`` (= nsamples / (2^M - 2)) is below 0.2, no
    selection otherwise; a float -> ``Lasso(alpha=l1_reg)``; ``False`` / 0 -> no selection.
    """
    lf = LINKS[link]
    M = Z.shape[1]
    eyAdj = lf(ey) - lf(fnull)
    d = lf(fx) - lf(fnull)
    s = np.sum(Z, 1)
    nonzero_inds = np.arange(M)
    if (l1_reg not in ["auto", False, 0]) or (fraction_evaluated < 0.2 and l1_reg == "auto"):
        w_aug = np.hstack((kw * (M - s), kw * s))
        w_sqrt_aug = np.sqrt(w_aug)
        eyAdj_aug = np.hstack((eyAdj, eyAdj - d))
        eyAdj_aug *= w_sqrt_aug
        mask_aug = np.transpose(w_sqrt_aug * np.transpose(np.vstack((Z, Z - 1))))
        if isinstance(l1_reg, str) and l1_reg.startswith("num_features("):
            from sklearn.linear_model import lars_path
            r = int(l1_reg[len("num_features("):-1])
            nonzero_inds = lars_path(mask_aug, eyAdj_aug, max_iter=r)[1]
        elif l1_reg in ("auto", "bic", "aic"):
            from sklearn.linear_model import LassoLarsIC
            from sklearn.pipeline import make_pipeline
            from sklearn.preprocessing import StandardScaler
            c = "aic" if l1_reg == "auto" else l1_reg
            if ic_standardize:
                model = make_pipeline(StandardScaler(with_mean=False), LassoLarsIC(criterion=c)).fit(mask_aug, eyAdj_aug)[1]
            else:
                model = LassoLarsIC(criterion=c).fit(mask_aug, eyAdj_aug)
            nonzero_inds = np.nonzero(model.coef_)[0]
        else:
            from sklearn.linear_model import Lasso
            nonzero_inds = np.nonzero(Lasso(alpha=l1_reg).fit(mask_aug, eyAdj_aug).coef_)[0]
    if len(nonzero_inds) == 0:
        return np.zeros(M)

    nonzero_inds = np.asarray(nonzero_inds)
    eyAdj2 = eyAdj - Z[:, nonzero_inds[-1]] * d
    etmp = np.transpose(np.transpose(Z[:, nonzero_inds[:-1]]) - Z[:, nonzero_inds[-1]])
    y = np.asarray(eyAdj2)
    X = etmp
    WX = kw[:, None] * X
    try:
        w = np.linalg.solve(X.T @ WX, WX.T @ y)
    except np.linalg.LinAlgError:
        warnings.warn("Linear regression equation is singular, a least squares solutions is used instead.")
        sqrt_W = np.sqrt(kw)
        w = np.linalg.lstsq(sqrt_W[:, None] * X, sqrt_W * y, rcond=None)[0]
    phi = np.zeros(M)
    phi[nonzero_inds[:-1]] = w
    phi[nonzero_inds[-1]] = d - sum(w)
    phi[np.abs(phi) < 1e-10] = 0
    return phi


# ------------------------------------------------------------------ composition
def varying_features(x: np.ndarray, background: np.ndarray) -> np.ndarray:
    """Step 1: features where x differs from at least one background row."""
    x = np.asarray(x, dtype=np.float64).reshape(-1)
    mismatch = ~np.isclose(background, x[None, :], equal_nan=True)
    return np.nonzero(mismatch.sum(0) > 0)[0]


def explain_row(f, x, background, nsamples="auto", l1_reg=False, link="identity",
                masks: tuple[np.ndarray, np.ndarray] | None = None, random=None,
                return_parts: bool = False):
    """One ``KernelExplainer.explain`` call.  ``masks=(Z, kw)`` bypasses the sampler."""
    background = np.asarray(background, dtype=np.float64)
    x = np.asarray(x, dtype=np.float64).reshape(-1)
    P = background.shape[1]
    lf = LINKS[link]
    fnull = float(np.mean(np.asarray(f(background), dtype=np.float64)))
    fx = float(np.asarray(f(x[None, :]), dtype=np.float64).reshape(-1)[0])
    varying = varying_features(x, background)
    M = len(varying)
    phi = np.zeros(P)
    parts = {"fx": fx, "fnull": fnull, "varying": varying}
    if M == 0:
        pass
    elif M == 1:
        phi[varying[0]] = lf(fx) - lf(fnull)
    else:
        if masks is None:
            Z, kw, info = sample_coalitions(M, nsamples, random=random)
            parts["info"] = info
            n_used, max_samples = resolve_nsamples(M, nsamples)
        else:
            Z, kw = masks
            n_used, max_samples = len(Z), resolve_nsamples(M, len(Z))[1]
        # explain(): self.solve(self.nsamples / self.max_samples, d), max_samples = 2^M - 2 for M <= 30 else 2^30 [recalled]
        fraction = n_used / float(max_samples)
        ey = evaluate(f, x, background, Z, varying)
        phi[varying] = solve(Z, kw, ey, fx, fnull, link=link, l1_reg=l1_reg, fraction_evaluated=fraction)
        parts.update(Z=Z, kw=kw, ey=ey)
    return (phi, parts) if return_parts else phi


def shap_values(f, X, background, nsamples="auto", l1_reg=False, link="identity",
                masks=None, random=None) -> np.ndarray:
    """``KernelExplainer.shap_values(X)`` for a 2-D X: one explain() per row.

    ``masks`` is None (draw per row from the global state, like the reference), one
    ``(Z, kw)`` pair shared by all rows, or a list with one pair per row.
    """
    X = np.asarray(X.values if hasattr(X, "values") else X, dtype=np.float64)
    if X.ndim == 1:
        X = X[None, :]
    out = np.zeros((X.shape[0], X.shape[1]))
    for i in range(X.shape[0]):
        mk = masks[i] if isinstance(masks, list) else masks
        out[i] = explain_row(f, X[i], background, nsamples, l1_reg, link, mk, random)
    return out


def expected_value(f, background, link="identity") -> float:
    return float(LINKS[link](np.mean(np.asarray(f(np.asarray(background, dtype=np.float64))))))


# ------------------------------------------------------------------ brute force KAT
def exact_shapley(f, x, background, players=None) -> np.ndarray:
    """Exact interventional Shapley values by enumerating all coalitions of ``players`` (default: every feature;
    pass ``varying_features(x, background)`` to leave out the null players, whose value is 0)."""
    x = np.asarray(x, dtype=np.float64).reshape(-1)
    background = np.asarray(background, dtype=np.float64)
    players = list(range(x.shape[0])) if players is None else [int(j) for j in players]
    M = len(players)
    cache: dict = {}

    def v(S):
        key = frozenset(S)
        if key not in cache:
            synth = background.copy()
            idx = list(S)
            synth[:, idx] = x[idx]
            cache[key] = float(np.mean(f(synth)))
        return cache[key]

    phi = np.zeros(x.shape[0])
    for i in players:
        rest = [j for j in players if j != i]
        for k in range(M):
            for S in itertools.combinations(rest, k):
                wgt = math.factorial(k) * math.factorial(M - k - 1) / math.factorial(M)
                phi[i] += wgt * (v(S + (i,)) - v(S))
    return phi
