"""Linear SHAP oracle (test infrastructure).  PARITY UNPINNED - see oracle/__init__.py.

Restates shap.LinearExplainer with ``feature_perturbation="interventional"`` as the
reference uses it:

* ``explain_model.py:24``  ``LinearExplainer(model, X_test_scaled, "interventional")`` -
  the data is wrapped in ``maskers.Independent(data, max_samples=100)``; more than 100
  rows are subsampled with ``sklearn.utils.resample(data, n_samples=100, random_state=0,
  replace=False)``; ``mean = background.mean(0)``; ``coef = model.coef_[0]``;
  ``expected_value = coef . mean + intercept``  [recalled from shap/explainers/_linear.py,
  shap/maskers/_tabular.py, shap/utils/_legacy.py::sample].
* ``explain_model.py:27`` / ``api/worker.py:75``  ``shap_values(X) = (X - mean) * coef``
  (DataFrame -> ``.values`` first); explains the LOGIT (quirk Q5 in SURVEY.md).
* ``explain_model.py:39-40``  ``mean_abs = abs(phi).mean(0)``; ``top3 = argsort()[-3:][::-1]``.
* ``xai_tasks.py:106-107``  ``coef_.ravel() * x`` (the same with a zero background).
"""

from __future__ import annotations

import numpy as np

MAX_BACKGROUND = 100


def background_subsample(data: np.ndarray, max_samples: int = MAX_BACKGROUND) -> np.ndarray:
    data = np.asarray(data.values if hasattr(data, "values") else data)
    if data.shape[0] <= max_samples:
        return data
    from sklearn.utils import resample
    return resample(data, n_samples=max_samples, random_state=0, replace=False)


def explainer_state(coef: np.ndarray, intercept: float, data) -> tuple[np.ndarray, float]:
    """(mean, expected_value) as LinearExplainer.__init__ computes them."""
    bg = np.asarray(background_subsample(data), dtype=np.float64)
    mean = np.array(np.mean(bg, 0)).flatten()
    return mean, float(np.dot(coef, mean) + intercept)


def shap_values(coef: np.ndarray, mean: np.ndarray, X) -> np.ndarray:
    X = np.asarray(X.values if hasattr(X, "values") else X, dtype=np.float64)
    return np.array(X - mean) * coef


def coef_times_x(coef: np.ndarray, x: np.ndarray) -> np.ndarray:
    """xai_tasks.py:106-107."""
    return np.asarray(coef).ravel() * np.asarray(x, dtype=np.float64)


def global_summary(phi: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """explain_model.py:39-40 -> (mean |phi| per feature, top-3 indices, largest first)."""
    mean_abs = np.abs(phi).mean(axis=0)
    return mean_abs, mean_abs.argsort()[-3:][::-1]
