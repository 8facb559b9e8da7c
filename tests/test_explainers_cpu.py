"""Host logic of the drop-in explainers that needs no GPU: identifying a logistic-linear model behind an opaque
callable (what a shap user hands to KernelExplainer), and the background-subsample rule of the linear explainer."""

import numpy as np
import pytest

from fraud_detection_b200.explainers import SUBSAMPLE_RULES, _background, _identify_logistic, _linear_model_of
from oracle import linear_shap, scoring, synth


def test_callable_is_identified_as_the_logistic_model(params):
    Xs = scoring.transform(params, synth.rows(300, seed=3))
    f1 = lambda A: scoring.predict_proba(params, A)[:, 1]  # noqa: E731
    f2 = lambda A: scoring.predict_proba(params, A)        # noqa: E731  (both columns, like the bound predict_proba)
    for f, n_out in ((f1, 1), (f2, 2)):
        coef, intercept, n = _linear_model_of(f, Xs[:100])
        assert n == n_out
        np.testing.assert_allclose(coef, params.coef, rtol=1e-9, atol=1e-12)
        assert abs(intercept - params.intercept) < 1e-9
    # a one-row, all-zero background (api/worker.py:52) still identifies the model
    coef, intercept, _ = _identify_logistic(f1, np.zeros((1, 30)))
    np.testing.assert_allclose(coef, params.coef, rtol=1e-9, atol=1e-12)


def test_estimators_tuples_and_bound_methods_are_taken_as_they_are(params):
    scaler, model, _ = scoring.live_estimators(params)
    for m, n_out in ((model, 1), ((params.coef, params.intercept), 1), (model.predict_proba, 2)):
        coef, intercept, n = _linear_model_of(m, None)
        assert n == n_out and np.array_equal(coef, params.coef) and intercept == params.intercept


def test_non_logistic_callables_are_refused(params):
    Xs = scoring.transform(params, synth.rows(100, seed=4))
    with pytest.raises(ValueError, match="not sigmoid"):
        _identify_logistic(lambda A: 1.0 / (1.0 + np.exp(-(A @ params.coef + 0.3 * A[:, 1] ** 2))), Xs)
    with pytest.raises(ValueError, match="binary"):
        _identify_logistic(lambda A: np.full((A.shape[0], 3), 1.0 / 3.0), Xs)


def test_background_subsample_rules(params):
    data = scoring.transform(params, synth.rows(1000, seed=5))
    assert _background(data[:100]) is not None and _background(data[:100]).shape == (100, 30)      # small backgrounds are kept whole
    default = _background(data)
    np.testing.assert_array_equal(default, linear_shap.background_subsample(data))                 # the oracle's (recalled) rule
    assert len(np.unique(default, axis=0)) == 100
    with_repl = _background(data, subsample="resample_with_replacement")
    assert with_repl.shape == (100, 30) and not np.array_equal(with_repl, default)
    np.testing.assert_array_equal(_background(data, subsample="first"), data[:100])
    assert _background(data, subsample="none").shape == (1000, 30)
    np.testing.assert_array_equal(_background(data, subsample=lambda d, k: d[-k:]), data[-100:])
    assert _background(data, max_samples=10).shape == (10, 30)
    with pytest.raises(ValueError):
        _background(data, subsample="bogus")
    assert "resample_without_replacement" in SUBSAMPLE_RULES
