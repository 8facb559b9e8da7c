"""CPU-only checks of the C-ABI boundary: the library loads, exports every symbol that
include/fdb200.h declares, reports errors through return codes (never aborts), and the
product path fails loudly without a GPU (no CPU fallback)."""

import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    with open(os.path.join(ROOT, "include", "fdb200.h"), "r", encoding="utf-8") as fh:
        text = re.sub(r"/\*.*?\*/", "", fh.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(fdb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported():
    from fraud_detection_b200 import _lib
    names = _declared_symbols()
    assert len(names) >= 10
    for n in names:
        assert hasattr(_lib.lib, n), f"{n} declared in include/fdb200.h but not exported"
    assert _lib.lib.fdb_version() >= 100


def test_every_declared_symbol_is_bound_with_a_signature():
    from fraud_detection_b200 import _lib
    for n in _declared_symbols():
        assert getattr(_lib.lib, n).argtypes is not None, f"{n} has no ctypes signature"


def test_bad_arguments_return_codes_not_aborts():
    from fraud_detection_b200 import _lib
    out = ctypes.c_void_p()
    rc = _lib.lib.fdb_ctx_create(0, None)
    assert rc == _lib.ERR_ARG and b"NULL" in _lib.lib.fdb_last_error()
    rc = _lib.lib.fdb_linear_pass_host(None, 0, 1, None, None, None, 0.0, None, 1, None, None, None, None)
    assert rc == _lib.ERR_ARG
    with pytest.raises(ValueError):
        _lib.check(rc)
    assert _lib.lib.fdb_ctx_destroy(None) == 0
    del out


def test_no_cpu_fallback_without_gpu(params):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from fraud_detection_b200 import B200LogisticRegression, FdbError
    model = B200LogisticRegression(params.coef, params.intercept)
    with pytest.raises(FdbError, match="no CPU fallback"):
        model.predict_proba(np.zeros((2, 30)))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "fraud_detection_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dirpath, f), "r", encoding="utf-8") as fh:
                    src = fh.read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports oracle"


def test_host_validation_matches_sklearn_messages(params):
    """Argument checking happens before any device call, so it is testable without a GPU."""
    import warnings
    import pandas as pd
    from fraud_detection_b200 import B200LogisticRegression, B200StandardScaler
    from fraud_detection_b200.estimators import _rows
    from oracle import scoring
    scaler_sk, model_sk, _ = scoring.live_estimators(params)
    scaler, model = B200StandardScaler.wrap(scaler_sk), B200LogisticRegression.wrap(model_sk)
    assert np.array_equal(scaler.mean_, scaler_sk.mean_) and np.array_equal(model.coef_, model_sk.coef_)
    assert model.classes_.tolist() == [0, 1] and model.n_features_in_ == 30
    for bad in (np.zeros(30), np.zeros((2, 29)), np.full((1, 30), np.nan), np.full((1, 30), np.inf)):
        with pytest.raises(ValueError) as ours:
            _rows(bad, model)
        with pytest.raises(ValueError) as theirs:
            model_sk.predict_proba(bad)
        assert str(ours.value).split("\n")[0][:40] == str(theirs.value).split("\n")[0][:40]
    # StandardScaler lets NaN through but not infinity
    _rows(np.full((1, 30), np.nan), scaler, allow_nan=True)
    with pytest.raises(ValueError, match="infinity"):
        _rows(np.full((1, 30), np.inf), scaler, allow_nan=True)
    # feature-name warnings: int-labelled DataFrame (api/app.py:184) and named DataFrame on the model
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        _rows(pd.DataFrame([[0.1] * 30]), scaler, allow_nan=True)
        _rows(pd.DataFrame([[0.1] * 30], columns=params.feature_names), model)
    msgs = [str(x.message) for x in w]
    assert any("does not have valid feature names, but StandardScaler" in m for m in msgs)
    assert any("has feature names, but LogisticRegression was fitted without" in m for m in msgs)
    with pytest.raises(ValueError, match="feature names should match"):
        _rows(pd.DataFrame([[0.1] * 30], columns=[f"f{i}" for i in range(30)]), scaler, allow_nan=True)
