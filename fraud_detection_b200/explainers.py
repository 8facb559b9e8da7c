"""Drop-in SHAP explainers for the reference's explanation path, executed on the B200.

    EXPL = shap.LinearExplainer(MODEL, X_sample)                         api/worker.py:52-53
    explainer = shap.LinearExplainer(model, X_s, feature_perturbation="interventional")
    shap_values = explainer.shap_values(X_s)                             explain_model.py:24-27
    np.abs(shap_values).mean(axis=0); argsort()[-3:][::-1]               explain_model.py:39-40

``B200LinearExplainer`` keeps that construction and call surface; ``shap_values`` returns a
host ``(N, 30)`` array computed by the fused linear-pass kernel.  ``shap`` itself is not
installed in this image, so behaviour is mirrored from the published algorithm (see
oracle/linear_shap.py for the citations); parity against the shap package is unpinned.
"""

from __future__ import annotations


import numpy as np

from . import _lib
from ._lib import N_FEATURES

MAX_BACKGROUND = 100   # shap.maskers.Independent(max_samples=100)


def _parse_model(model):
    """shap's LinearExplainer._parse_model for sklearn-style linear models / (coef, intercept)."""
    if isinstance(model, tuple) and len(model) == 2:
        coef, intercept = model
        return np.asarray(coef, dtype=np.float64).reshape(-1), float(np.asarray(intercept).reshape(-1)[0])
    if hasattr(model, "coef_") and hasattr(model, "intercept_"):
        coef = np.asarray(model.coef_, dtype=np.float64)
        if coef.ndim == 2:
            if coef.shape[0] != 1:
                raise ValueError("multi-output linear models are not supported on the B200 path")
            coef = coef[0]
        return coef.copy(), float(np.asarray(model.intercept_).reshape(-1)[0])
    raise ValueError("An unknown model type was passed: " + str(type(model)))


def _background(data, max_samples=MAX_BACKGROUND) -> np.ndarray:
    """maskers.Independent: keep <= max_samples rows (resample without replacement, seed 0)."""
    data = data.values if hasattr(data, "values") else data
    data = np.asarray(data, dtype=np.float64)
    if data.ndim != 2:
        raise ValueError("background data must be 2-D")
    if data.shape[0] > max_samples:
        from sklearn.utils import resample
        data = resample(data, n_samples=max_samples, random_state=0, replace=False)
    return data


def _plain_rows(X) -> np.ndarray:
    """shap accepts DataFrame / Series / ndarray and does no sklearn-style validation."""
    if hasattr(X, "values"):
        X = X.values
    A = np.asarray(X)
    if A.dtype not in (np.float32, np.float64):
        A = A.astype(np.float64)
    single = A.ndim == 1
    if single:
        A = A.reshape(1, -1)
    if A.ndim != 2 or A.shape[1] != N_FEATURES:
        raise ValueError(f"X must have {N_FEATURES} features per row, got shape {A.shape}")
    return np.ascontiguousarray(A), single


class B200LinearExplainer:
    """``shap.LinearExplainer`` with ``feature_perturbation="interventional"``.

    phi = (X - mean(background)) * coef explains the LOGIT (``expected_value`` =
    coef . mean + intercept), exactly as the reference uses it.
    """

    def __init__(self, model, data, feature_perturbation="interventional", device=None, **kwargs):
        if feature_perturbation not in (None, "interventional"):
            raise NotImplementedError("only feature_perturbation='interventional' is on the reference path")
        self.feature_perturbation = "interventional"
        self.coef, self.intercept = _parse_model(model)
        if self.coef.shape[0] != N_FEATURES:
            raise ValueError(f"the B200 path is built for {N_FEATURES} features, got {self.coef.shape[0]}")
        if isinstance(data, tuple) and len(data) == 2:          # (mean, cov) form
            self.mean = np.asarray(data[0], dtype=np.float64).reshape(-1)
        else:
            self.mean = np.array(np.mean(_background(data), 0)).flatten()
        self.expected_value = float(np.dot(self.coef, self.mean) + self.intercept)
        self.M = len(self.mean)
        self.device = device
        self._mean_c = _lib.dvec(self.mean)
        self._coef_c = _lib.dvec(self.coef)

    def shap_values(self, X):
        A, single = _plain_rows(X)
        phi = np.empty_like(A)
        _lib.check(_lib.lib.fdb_linear_pass_host(
            _lib.get_ctx(self.device), _lib.dtype_code(A.dtype), A.shape[0], _lib.np_ptr(A), self._mean_c,
            self._coef_c, self.expected_value, None, 1, None, _lib.np_ptr(phi), None, None))
        phi = phi.astype(np.float64, copy=False)
        return phi[0] if single else phi

    def explain_with_summary(self, X):
        """(phi, mean|phi| per feature, top-3 indices) in one pass - explain_model.py:27,39-40."""
        A, _ = _plain_rows(X)
        phi = np.empty_like(A)
        abs_sum = np.zeros(32, dtype=np.float64)
        _lib.check(_lib.lib.fdb_linear_pass_host(
            _lib.get_ctx(self.device), _lib.dtype_code(A.dtype), A.shape[0], _lib.np_ptr(A), self._mean_c,
            self._coef_c, self.expected_value, None, 1, None, _lib.np_ptr(phi), None,
            _lib.np_ptr(abs_sum)))
        mean_abs = abs_sum[:N_FEATURES] / max(A.shape[0], 1)
        return phi.astype(np.float64, copy=False), mean_abs, mean_abs.argsort()[-3:][::-1]
