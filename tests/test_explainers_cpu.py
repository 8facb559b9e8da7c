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


def test_l1_reg_values_are_parsed_like_shap_takes_them():
    """Every l1_reg value KernelExplainer.shap_values accepts maps to a device mode; "auto" is decided per row from
    M = number of varying features: AIC when the coalitions cover under 20 % of the 2^M - 2 possible ones."""
    from fraud_detection_b200.explainers import L1_AIC, L1_ALPHA, L1_BIC, L1_NONE, L1_NUM_FEATURES, _L1Spec
    assert _L1Spec(False).resolve(30, "auto") == (L1_NONE, 0.0) == _L1Spec(0).resolve(30, 2048)
    assert _L1Spec("num_features(10)").resolve(30, "auto") == (L1_NUM_FEATURES, 10.0)
    assert _L1Spec("aic").resolve(5, "auto") == (L1_AIC, 0.0) and _L1Spec("bic").resolve(30, "auto") == (L1_BIC, 0.0)
    assert _L1Spec(0.01).resolve(30, "auto") == (L1_ALPHA, 0.01)
    auto = _L1Spec("auto")
    assert auto.resolve(30, "auto") == (L1_AIC, 0.0)            # 2108 of 2^30 - 2
    assert auto.resolve(12, 1024) == (L1_NONE, 0.0)             # 1024 of 4094 = 0.25
    assert auto.resolve(13, 1024) == (L1_AIC, 0.0)              # 1024 of 8190 = 0.125
    assert auto.resolve(6, "auto") == (L1_NONE, 0.0)            # capped at the 62 possible coalitions: fraction 1
    for bad in ("ridge", "num_features(0)", "num_features(17)", -1.0, float("nan")):
        with pytest.raises(ValueError):
            _L1Spec(bad)


def test_oracle_auto_rule_follows_the_evaluated_fraction(params):
    """The oracle's own "auto": same rule, through explain_row (no GPU, real sklearn)."""
    import warnings
    from oracle import kernel_shap as ks
    Xs = scoring.transform(params, synth.rows(60, seed=5))
    bg, x = Xs[:20], Xs[40]
    f = lambda A: scoring.predict_proba(params, A)[:, 1]  # noqa: E731
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        auto = ks.explain_row(f, x, bg, nsamples=300, l1_reg="auto", random=np.random.RandomState(1))
        aic = ks.explain_row(f, x, bg, nsamples=300, l1_reg="aic", random=np.random.RandomState(1))
        none = ks.explain_row(f, x, bg, nsamples=300, l1_reg=False, random=np.random.RandomState(1))
    np.testing.assert_array_equal(auto, aic)                     # 300 of 2^30 - 2 coalitions
    assert np.count_nonzero(none) == 30 and np.count_nonzero(aic) < 30
    # few varying features: everything is enumerated, "auto" selects nothing
    x6 = bg[0].copy()
    x6[:6] = x[:6]
    bg1 = bg[:1]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        a = ks.explain_row(f, x6, bg1, nsamples=300, l1_reg="auto", random=np.random.RandomState(2))
        n = ks.explain_row(f, x6, bg1, nsamples=300, l1_reg=False, random=np.random.RandomState(2))
    np.testing.assert_array_equal(a, n)
