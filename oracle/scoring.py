"""Scoring oracle (test infrastructure): StandardScaler + LogisticRegression in float64.

Two forms are offered and cross-checked in ``tests/test_oracle_scoring.py``:

* a plain numpy restatement (``transform``, ``decision_function``, ``predict_proba``,
  ``predict``) of ``SK:preprocessing/_data.py:1092-1135`` (``X -= mean_; X /= scale_`` on a
  copy), ``SK:linear_model/_base.py:366-396`` (``X @ coef_.T + intercept_``),
  ``SK:linear_model/_base.py:429-440`` / ``_logistic.py:1595-1625`` (``expit`` then
  ``stack([1-p, p])``) and ``_base.py:398-423`` (``classes_[(z > 0)]``);
* ``live_estimators`` - real sklearn objects: the reference's own pickles when
  ``oracle/_ref`` holds them (see ``build_ref.py``), else estimators rebuilt from the
  committed parameters.  These execute sklearn's own code and are the timed CPU
  baseline for the scoring legs.

PINNED by ``tests/golden/scoring_golden.npz`` (made by ``tools/make_golden.py`` from the
reference's ``models/*.joblib`` in this container).
"""

from __future__ import annotations

import json
import os
import warnings
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_GOLDEN = os.path.join(os.path.dirname(_HERE), "tests", "golden")
REF_DIR = os.path.join(_HERE, "_ref")


@dataclass
class Params:
    coef: np.ndarray        # (30,)  LogisticRegression.coef_[0]
    intercept: float        # LogisticRegression.intercept_[0]
    mean: np.ndarray        # (30,)  StandardScaler.mean_
    scale: np.ndarray       # (30,)  StandardScaler.scale_
    var: np.ndarray         # (30,)  StandardScaler.var_
    feature_names: list

    @property
    def n_features(self) -> int:
        return int(self.coef.shape[0])


def load_params(path: str | None = None) -> Params:
    """Parameters exported bit-exactly (float.hex) from the reference's joblib artefacts."""
    path = path or os.path.join(_GOLDEN, "model_params.json")
    with open(path, "r", encoding="utf-8") as fh:
        raw = json.load(fh)

    def arr(key):
        return np.array([float.fromhex(h) for h in raw[key]], dtype=np.float64)

    return Params(coef=arr("coef_hex"), intercept=float.fromhex(raw["intercept_hex"]),
                  mean=arr("mean_hex"), scale=arr("scale_hex"), var=arr("var_hex"),
                  feature_names=list(raw["feature_names"]))


# ---------------------------------------------------------------- numpy restatement
def _as_2d(X) -> np.ndarray:
    if hasattr(X, "values"):
        X = X.values
    X = np.asarray(X)
    if X.ndim != 2:
        raise ValueError(f"Expected 2D array, got {X.ndim}D array instead")
    return X


def transform(p: Params, X) -> np.ndarray:
    X = np.array(_as_2d(X), dtype=np.float64, copy=True)
    X -= p.mean
    X /= p.scale
    return X


def decision_function(p: Params, Xs) -> np.ndarray:
    Xs = np.asarray(_as_2d(Xs), dtype=np.float64)
    return Xs @ p.coef + p.intercept


def _expit(z: np.ndarray) -> np.ndarray:
    out = np.empty_like(z)
    pos = z >= 0
    out[pos] = 1.0 / (1.0 + np.exp(-z[pos]))
    ez = np.exp(z[~pos])
    out[~pos] = ez / (1.0 + ez)
    return out


def predict_proba(p: Params, Xs) -> np.ndarray:
    prob = _expit(decision_function(p, Xs))
    return np.stack([1.0 - prob, prob], axis=1)


def predict(p: Params, Xs) -> np.ndarray:
    return (decision_function(p, Xs) > 0).astype(np.int64)


# ---------------------------------------------------------------- live sklearn
def live_estimators(p: Params | None = None):
    """(scaler, model, origin).  origin is 'joblib' (reference pickles from oracle/_ref)
    or 'rebuilt' (sklearn estimators with the exported parameters assigned)."""
    from sklearn.linear_model import LogisticRegression
    from sklearn.preprocessing import StandardScaler

    sp = os.path.join(REF_DIR, "scaler.joblib")
    mp = os.path.join(REF_DIR, "logistic_model.joblib")
    if os.path.exists(sp) and os.path.exists(mp):
        import joblib
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            return joblib.load(sp), joblib.load(mp), "joblib"
    p = p or load_params()
    scaler = StandardScaler()
    scaler.mean_ = p.mean.copy()
    scaler.scale_ = p.scale.copy()
    scaler.var_ = p.var.copy()
    scaler.n_features_in_ = p.n_features
    scaler.n_samples_seen_ = 284807
    model = LogisticRegression(C=1.0, max_iter=1000, random_state=42)
    model.coef_ = p.coef.reshape(1, -1).copy()
    model.intercept_ = np.array([p.intercept])
    model.classes_ = np.array([0, 1])
    model.n_features_in_ = p.n_features
    model.n_iter_ = np.array([39], dtype=np.int32)
    return scaler, model, "rebuilt"
