"""Coalition sampler of Kernel SHAP (host logic of the product path).

Reproduces the sampling schedule of ``shap.KernelExplainer.explain`` [recalled from upstream
shap/explainers/_kernel.py; shap==0.49.1 pinned at the reference's requirements.txt:21]:
Shapley-kernel weights per coalition size, complete enumeration of the smallest/largest
sizes while the sample budget allows, then random paired draws (a coalition and its
complement) with duplicate merging, and renormalisation of the random part to the weight
left over.  It consumes the random stream exactly like the reference (one ``choice`` call,
then one ``permutation`` per draw, from the GLOBAL ``np.random`` state unless a
``RandomState`` is given), so the same seed gives the same masks.

Coalitions are returned as 30-bit integers (bit k set = feature k takes the explained
row's value) - the form ``fdb_kshap_plan_create`` uploads - plus float64 kernel weights.
"""

from __future__ import annotations

import itertools
from math import comb

import numpy as np


def resolve_nsamples(M: int, nsamples="auto") -> int:
    if nsamples == "auto":
        nsamples = 2 * M + 2 ** 11
    nsamples = int(nsamples)
    if M <= 30:
        nsamples = min(nsamples, 2 ** M - 2)
    return nsamples


def size_weights(M: int):
    n_sizes = int(np.ceil((M - 1) / 2.0))
    n_paired = int(np.floor((M - 1) / 2.0))
    w = np.array([(M - 1.0) / (i * (M - i)) for i in range(1, n_sizes + 1)])
    w[:n_paired] *= 2
    w /= np.sum(w)
    return w, n_sizes, n_paired


def sample_coalitions(M: int, nsamples="auto", random=None):
    """-> (bits uint32 (n,), weights float64 (n,)) over M varying features (bit j = j-th varying feature)."""
    random = np.random if random is None else random
    nsamples = resolve_nsamples(M, nsamples)
    full = (1 << M) - 1
    wv, n_sizes, n_paired = size_weights(M)
    bits: list[int] = []
    weights: list[float] = []

    # complete enumeration of sizes 1, 2, ... (with complements) while the budget allows
    n_full = 0
    left = nsamples
    remaining = wv.copy()
    for size in range(1, n_sizes + 1):
        nsub = comb(M, size) * (2 if size <= n_paired else 1)
        if left * remaining[size - 1] / nsub >= 1.0 - 1e-8:
            n_full += 1
            left -= nsub
            if remaining[size - 1] < 1.0:
                remaining /= (1 - remaining[size - 1])
            w = wv[size - 1] / comb(M, size)
            if size <= n_paired:
                w /= 2.0
            for inds in itertools.combinations(range(M), size):
                b = 0
                for i in inds:
                    b |= 1 << i
                bits.append(b)
                weights.append(w)
                if size <= n_paired:
                    bits.append(full ^ b)
                    weights.append(w)
        else:
            break

    n_fixed = len(bits)
    left = nsamples - n_fixed
    if n_full != n_sizes:
        rem = wv.copy()
        rem[:n_paired] /= 2
        rem = rem[n_full:]
        rem /= np.sum(rem)
        draws = random.choice(len(rem), 4 * left, p=rem)
        seen: dict[int, int] = {}
        pos = 0
        while left > 0 and pos < len(draws):
            size = int(draws[pos]) + n_full + 1
            pos += 1
            b = 0
            for i in random.permutation(M)[:size]:
                b |= 1 << int(i)
            new = b not in seen
            if new:
                seen[b] = len(bits)
                left -= 1
                bits.append(b)
                weights.append(1.0)
            else:
                weights[seen[b]] += 1.0
            if left > 0 and size <= n_paired:
                if new:
                    left -= 1
                    bits.append(full ^ b)
                    weights.append(1.0)
                else:
                    weights[seen[b] + 1] += 1.0
        kw = np.array(weights, dtype=np.float64)
        weight_left = float(np.sum(wv[n_full:]))
        kw[n_fixed:] *= weight_left / kw[n_fixed:].sum()
    else:
        kw = np.array(weights, dtype=np.float64)
    return np.array(bits, dtype=np.uint32), kw


def masks_to_bits(Z) -> np.ndarray:
    """(n, 30) 0/1 mask matrix (the reference's ``maskMatrix``) -> (n,) uint32 bit masks."""
    Z = np.asarray(Z)
    if Z.ndim != 2 or Z.shape[1] > 30:
        raise ValueError("mask matrix must be (n, <=30)")
    pw = (1 << np.arange(Z.shape[1], dtype=np.uint64))
    return ((Z != 0).astype(np.uint64) * pw).sum(1).astype(np.uint32)


def bits_to_masks(bits, M: int = 30) -> np.ndarray:
    bits = np.asarray(bits, dtype=np.uint32)
    return ((bits[:, None] >> np.arange(M, dtype=np.uint32)) & 1).astype(np.float64)
