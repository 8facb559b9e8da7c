"""Credit-card-shaped synthetic rows (oracle side; test infrastructure).

Restates ``scripts/generate_synthetic_data.py:6-27`` of the reference: global
``np.random.seed(42)``; ``V1..V28 = randn(n, 28)``; ``Time = sort(U(0, 172800))``;
``Amount = exp(N(3, 1))``; column order ``Time, V1..V28, Amount``.  The draw order
is kept so the same seed yields the same float64 matrix as the reference
function (checked in ``tools/make_golden.py`` by importing the reference here).
"""

from __future__ import annotations

import numpy as np

FEATURE_NAMES = ["Time"] + [f"V{i}" for i in range(1, 29)] + ["Amount"]
N_FEATURES = 30


def reference_rows(n_samples: int, seed: int = 42) -> np.ndarray:
    """(n_samples, 30) float64, identical to the reference generator for seed 42."""
    state = np.random.get_state()
    try:
        np.random.seed(seed)
        v = np.random.randn(n_samples, N_FEATURES - 2)
        time = np.sort(np.random.uniform(0, 172800, n_samples))
        amount = np.exp(np.random.normal(3, 1, n_samples))
    finally:
        np.random.set_state(state)
    return np.column_stack([time, v, amount])


def rows(n_samples: int, seed: int) -> np.ndarray:
    """Same three distributions from a private Generator (Time left unsorted)."""
    rng = np.random.default_rng(seed)
    out = np.empty((n_samples, N_FEATURES), dtype=np.float64)
    out[:, 1:29] = rng.standard_normal((n_samples, 28))
    out[:, 0] = rng.uniform(0.0, 172800.0, n_samples)
    out[:, 29] = np.exp(rng.normal(3.0, 1.0, n_samples))
    return out
