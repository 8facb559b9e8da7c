"""Drop-in estimators for the reference's scoring path, executed on the B200.

The reference keeps sklearn objects in module globals and calls three methods on them
(SURVEY.md 8(b)):

    SCALER.transform(X)                       api/app.py:194, explain_model.py:19, predict_single.py:25
    MODEL.predict(X) / MODEL.predict_proba(X) api/app.py:209,223; xai_tasks.py:97; evaluate_model.py:26-27
    model.coef_, SCALER.mean_                 xai_tasks.py:106, api/app.py:185

``wrap(joblib.load(path))`` returns an object with the same attributes, return types
(host numpy: ``transform`` keeps the float dtype, ``predict`` int64 ``(N,)``,
``predict_proba`` float64 ``(N, 2)``) and error behaviour (``ValueError`` on a wrong
feature count / NaN / infinity / 1-D input, sklearn's feature-name warnings), but every
row is computed by the hand-written CUDA kernel behind ``libfdb200.so``.  float64 input is
computed in float64 on the device (parity ~1e-15); float32 input in float32 (parity
~1e-6 against sklearn's float64).  No CPU fallback exists.
"""

from __future__ import annotations

import warnings

import numpy as np

from . import _lib
from ._lib import N_FEATURES

# host-side finiteness scan only below this many elements; larger batches rely on the
# |phi| column sums produced by the kernel itself (a NaN/inf input poisons its column sum)
_HOST_SCAN_ELEMS = 1 << 20


def _feature_names_of(X):
    """String column labels of a DataFrame as an object array, else None (sklearn ignores non-string labels:
    api/app.py:184 passes ints).  ``inferred_type`` is cached by pandas and ``to_numpy`` avoids the 50 us that
    ``list(columns)`` costs on an Arrow-backed string Index - this runs on every serving request."""
    cols = getattr(X, "columns", None)
    if cols is None:
        return None
    kind = getattr(cols, "inferred_type", None)
    if kind is None:                                   # not a pandas Index: the generic route
        names = list(cols)
        if len(names) and all(isinstance(c, str) for c in names):
            return np.asarray(names, dtype=object)
        return None
    if kind != "string" or len(cols) == 0:
        return None
    return cols.to_numpy(dtype=object)


def _names_match_fast(X, est) -> bool:
    """True when X carries exactly the fitted feature names (one C-level Index comparison, the serving hot path)."""
    cols = getattr(X, "columns", None)
    fitted = getattr(est, "feature_names_in_", None)
    if cols is None or fitted is None or not hasattr(cols, "equals"):
        return False
    fi = getattr(est, "_fitted_index", None)
    if fi is None:
        import pandas as pd
        fi = est._fitted_index = pd.Index(fitted)
    return bool(cols.equals(fi))


def _rows(X, est, *, allow_nan: bool = False):
    """sklearn-style validation -> C-contiguous float32/float64 (N, 30) host array."""
    name = type(est).sk_name
    if not _names_match_fast(X, est):
        fitted = getattr(est, "feature_names_in_", None)
        cols = getattr(X, "columns", None)
        if fitted is None and getattr(cols, "inferred_type", None) is not None:
            given = True if (cols.inferred_type == "string" and len(cols)) else None      # only "has names?" matters here
        else:
            given = _feature_names_of(X)
        if fitted is not None and given is None:
            warnings.warn(f"X does not have valid feature names, but {name} was fitted with feature names", UserWarning)
        elif fitted is None and given is not None:
            warnings.warn(f"X has feature names, but {name} was fitted without feature names", UserWarning)
        elif fitted is not None and given is not None and (len(fitted) != len(given) or np.any(fitted != given)):
            raise ValueError("The feature names should match those that were passed during fit.")
    if hasattr(X, "to_numpy"):
        X = X.to_numpy()
    A = np.asarray(X)
    if A.dtype not in (np.float32, np.float64):
        try:
            A = A.astype(np.float64)
        except (TypeError, ValueError) as exc:
            raise ValueError(f"could not convert input to float: {exc}") from None
    if A.ndim == 1:
        raise ValueError(
            f"Expected 2D array, got 1D array instead:\narray={A}.\nReshape your data either using "
            "array.reshape(-1, 1) if your data has a single feature or array.reshape(1, -1) if it "
            "contains a single sample.")
    if A.ndim != 2:
        raise ValueError(f"Found array with dim {A.ndim}. {name} expected <= 2.")
    if A.shape[1] != est.n_features_in_:
        raise ValueError(f"X has {A.shape[1]} features, but {name} is expecting {est.n_features_in_} features as input.")
    if A.size <= _HOST_SCAN_ELEMS:
        _raise_if_not_finite(A, allow_nan)
    return np.ascontiguousarray(A)


def _raise_if_not_finite(A: np.ndarray, allow_nan: bool) -> None:
    if A.size == 0 or np.isfinite(A.sum()):
        return
    if not allow_nan and np.isnan(A).any():
        raise ValueError("Input X contains NaN.")
    if np.isinf(A).any():
        raise ValueError(f"Input X contains infinity or a value too large for dtype('{A.dtype}').")


def _check_abs_sum(abs_sum: np.ndarray, A: np.ndarray) -> None:
    """Large batches: the kernel's |phi| column sums flag NaN / infinity in the input."""
    if A.size > _HOST_SCAN_ELEMS and not np.all(np.isfinite(abs_sum[:N_FEATURES])):
        _raise_if_not_finite(A, allow_nan=False)
        raise ValueError("Input X contains infinity or a value too large for dtype('%s')." % A.dtype)


class B200StandardScaler:
    """``sklearn.preprocessing.StandardScaler`` (fitted) with ``transform`` on the GPU.

    Mirrors SK:preprocessing/_data.py:1092-1135: ``(X - mean_) / scale_`` on a copy, same
    float dtype out as in, NaN passes through, infinity raises.
    """

    sk_name = "StandardScaler"

    def __init__(self, mean, scale, var=None, n_samples_seen=None, feature_names_in=None, device=None):
        self.mean_ = np.asarray(mean, dtype=np.float64).copy()
        self.scale_ = np.asarray(scale, dtype=np.float64).copy()
        self.var_ = None if var is None else np.asarray(var, dtype=np.float64).copy()
        self.n_features_in_ = int(self.mean_.shape[0])
        if self.n_features_in_ != N_FEATURES:
            raise ValueError(f"the B200 path is built for {N_FEATURES} features, got {self.n_features_in_}")
        self.n_samples_seen_ = n_samples_seen
        if feature_names_in is not None:
            self.feature_names_in_ = np.asarray(feature_names_in, dtype=object)
        self.with_mean = True
        self.with_std = True
        self.copy = True
        self.device = device
        self._mean_c = _lib.dvec(self.mean_)
        self._scale_c = _lib.dvec(self.scale_)

    @classmethod
    def wrap(cls, sk, device=None):
        if getattr(sk, "with_mean", True) is not True or getattr(sk, "with_std", True) is not True:
            raise ValueError("only with_mean=True, with_std=True scalers are supported")
        return cls(sk.mean_, sk.scale_, getattr(sk, "var_", None), getattr(sk, "n_samples_seen_", None),
                   getattr(sk, "feature_names_in_", None), device=device)

    def get_feature_names_out(self, input_features=None):
        names = getattr(self, "feature_names_in_", None)
        return names if names is not None else np.asarray([f"x{i}" for i in range(self.n_features_in_)], dtype=object)

    def transform(self, X, copy=None):
        A = _rows(X, self, allow_nan=True)
        out = np.empty_like(A)
        _lib.check(_lib.lib.fdb_transform_host(_lib.get_ctx(self.device), _lib.dtype_code(A.dtype), A.shape[0],
                                               _lib.np_ptr(A), self._mean_c, self._scale_c, _lib.np_ptr(out)))
        return out


class B200LogisticRegression:
    """Binary ``sklearn.linear_model.LogisticRegression`` (fitted) scored on the GPU.

    ``decision_function`` / ``predict`` / ``predict_proba`` follow
    SK:linear_model/_base.py:366-440 and _logistic.py:1595-1625: ``z = X @ coef_.T +
    intercept_``, ``classes_[z > 0]``, ``[1 - expit(z), expit(z)]``.  ``score_rows`` is an
    addition: both outputs from ONE pass (the reference calls predict and predict_proba back
    to back, api/app.py:209-223, computing z twice).
    """

    sk_name = "LogisticRegression"

    def __init__(self, coef, intercept, classes=(0, 1), n_iter=None, feature_names_in=None, device=None):
        self.coef_ = np.asarray(coef, dtype=np.float64).reshape(1, -1).copy()
        self.intercept_ = np.asarray(intercept, dtype=np.float64).reshape(1).copy()
        self.classes_ = np.asarray(classes).copy()
        if self.classes_.shape[0] != 2:
            raise ValueError("only binary LogisticRegression is supported")
        self.n_features_in_ = int(self.coef_.shape[1])
        if self.n_features_in_ != N_FEATURES:
            raise ValueError(f"the B200 path is built for {N_FEATURES} features, got {self.n_features_in_}")
        if n_iter is not None:
            self.n_iter_ = np.asarray(n_iter)
        if feature_names_in is not None:
            self.feature_names_in_ = np.asarray(feature_names_in, dtype=object)
        self.device = device
        self._coef_c = _lib.dvec(self.coef_[0])

    @classmethod
    def wrap(cls, sk, device=None):
        if np.asarray(sk.coef_).shape[0] != 1:
            raise ValueError("only binary LogisticRegression (coef_.shape[0] == 1) is supported")
        return cls(sk.coef_, sk.intercept_, sk.classes_, getattr(sk, "n_iter_", None),
                   getattr(sk, "feature_names_in_", None), device=device)

    # -- one kernel launch; which outputs are materialised is chosen per method ------------
    #: Request-size batches (up to this many rows) always get proba, logit and class from ONE launch, and the result
    #: is kept for the next call on byte-identical rows: the reference calls ``predict`` and ``predict_proba`` back
    #: to back on the same request (api/app.py:209,223), which sklearn answers with two full passes.
    pair_cache_rows = 512

    def _pass(self, X, *, proba=False, logit=False, cls=False):
        A = _rows(X, self)
        n = A.shape[0]
        dt = A.dtype
        small = 0 < n <= self.pair_cache_rows
        if small:
            # same request as the previous call?  cheap tests first, then one pass over the rows (no copy, no hashing)
            key = (dt.str, n, id(self._coef_c), float(self.intercept_[0]))
            last = getattr(self, "_last_pass", None)
            if last is not None and last[0] == key and np.array_equal(last[1], A):
                _, _, p, z, c = last
                return (p.copy() if proba else None), (z.copy() if logit else None), (c.copy() if cls else None)
            proba_k = logit_k = cls_k = True
        else:
            proba_k, logit_k, cls_k = proba, logit, cls
        out_p = np.empty((n, 2), dtype=dt) if proba_k else None
        out_z = np.empty((n,), dtype=dt) if logit_k else None
        out_c = np.empty((n,), dtype=np.uint8) if cls_k else None
        abs_sum = np.zeros(32, dtype=np.float64)
        _lib.check(_lib.lib.fdb_linear_pass_host(
            _lib.get_ctx(self.device), _lib.dtype_code(dt), n, _lib.np_ptr(A), None, self._coef_c,
            float(self.intercept_[0]), _lib.np_ptr(out_p), 2, _lib.np_ptr(out_z), None, _lib.np_ptr(out_c),
            _lib.np_ptr(abs_sum)))
        _check_abs_sum(abs_sum, A)
        if small:
            self._last_pass = (key, A.copy(), out_p, out_z, out_c)      # one reference assignment: safe against concurrent readers
            return (out_p.copy() if proba else None), (out_z.copy() if logit else None), (out_c.copy() if cls else None)
        return out_p, out_z, out_c

    def decision_function(self, X):
        return self._pass(X, logit=True)[1].astype(np.float64, copy=False)

    def predict(self, X):
        return self.classes_.take(self._pass(X, cls=True)[2].astype(np.intp), axis=0)

    def predict_proba(self, X):
        return self._pass(X, proba=True)[0].astype(np.float64, copy=False)

    def predict_log_proba(self, X):
        return np.log(self.predict_proba(X))

    def score_rows(self, X):
        """(predict(X), predict_proba(X)) from a single pass over X."""
        p, _, c = self._pass(X, proba=True, cls=True)
        return self.classes_.take(c.astype(np.intp), axis=0), p.astype(np.float64, copy=False)


def wrap(obj, device=None):
    """Replace a fitted sklearn StandardScaler / LogisticRegression by its B200 twin.

    One line changes at each reference site, e.g. api/app.py:44-47::

        MODEL  = wrap(joblib.load('models/logistic_model.joblib'))
        SCALER = wrap(joblib.load('models/scaler.joblib'))
    """
    kind = type(obj).__name__
    if kind == "StandardScaler":
        return B200StandardScaler.wrap(obj, device=device)
    if kind == "LogisticRegression":
        return B200LogisticRegression.wrap(obj, device=device)
    if isinstance(obj, (B200StandardScaler, B200LogisticRegression)):
        return obj
    raise TypeError(f"cannot wrap {kind}: expected a fitted StandardScaler or LogisticRegression")


def from_joblib(model_path: str | None = None, scaler_path: str | None = None, device=None):
    """Load the reference artefacts (same env vars / default paths as api/utils.py:10-12 and
    api/app.py:44-47) and wrap them: returns ``(model, scaler)``."""
    import os

    import joblib
    model_path = model_path or os.getenv("MODEL_PATH", "./models/logistic_model.joblib")
    scaler_path = scaler_path or os.getenv("SCALER_PATH", "./models/scaler.joblib")
    if not os.path.exists(model_path):
        raise FileNotFoundError(f"Model not found at {model_path}")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        model = joblib.load(model_path)
        scaler = joblib.load(scaler_path) if os.path.exists(scaler_path) else None
    return wrap(model, device=device), (wrap(scaler, device=device) if scaler is not None else None)
