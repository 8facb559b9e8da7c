"""Pin the scoring oracle to the reference's own artefacts (CPU only).

Golden vectors come from the reference's models/*.joblib run through sklearn in the
build container (tools/make_golden.py).  Both oracle forms - the numpy restatement and
the live sklearn estimators - must reproduce them.
"""

import numpy as np

from oracle import linear_shap, scoring, synth


def test_params_roundtrip(params):
    assert params.coef.shape == (30,) and params.mean.shape == (30,) and params.scale.shape == (30,)
    assert params.feature_names == synth.FEATURE_NAMES
    assert abs(params.intercept - (-4.23085)) < 1e-5
    np.testing.assert_allclose(params.scale, np.sqrt(params.var), rtol=1e-15)


def test_generator_matches_reference_rows(golden):
    assert np.array_equal(synth.reference_rows(512), golden["X"])


def test_numpy_restatement_vs_golden(params, golden):
    X = golden["X"]
    Xs = scoring.transform(params, X)
    assert np.array_equal(Xs, golden["X_scaled"])           # same two in-place passes -> bit exact
    z = scoring.decision_function(params, Xs)
    np.testing.assert_allclose(z, golden["decision"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(scoring.predict_proba(params, Xs), golden["proba"], rtol=0, atol=1e-15)
    assert np.array_equal(scoring.predict(params, Xs), golden["pred"])


def test_live_estimators_vs_golden(params, golden):
    scaler, model, origin = scoring.live_estimators(params)
    assert origin in ("joblib", "rebuilt")
    Xs = scaler.transform(golden["X"])
    assert np.array_equal(Xs, golden["X_scaled"])
    np.testing.assert_allclose(model.decision_function(Xs), golden["decision"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(model.predict_proba(Xs), golden["proba"], rtol=0, atol=1e-15)
    assert np.array_equal(model.predict(Xs), golden["pred"])


def test_rebuilt_estimators_equal_restatement(params, golden, monkeypatch):
    monkeypatch.setattr(scoring, "REF_DIR", "/nonexistent")
    scaler, model, origin = scoring.live_estimators(params)
    assert origin == "rebuilt"
    Xs = scaler.transform(golden["X"])
    assert np.array_equal(Xs, golden["X_scaled"])
    np.testing.assert_allclose(model.predict_proba(Xs), golden["proba"], rtol=0, atol=1e-15)


def test_full_config1_checksums(params, golden):
    """Config 1 at full size (284,807 x 30): float64 checksums of the reference outputs."""
    X = synth.reference_rows(int(golden["full_n"]))
    Xs = scoring.transform(params, X)
    np.testing.assert_allclose(Xs.sum(0), golden["full_scaled_colsum"], rtol=1e-12)
    z = scoring.decision_function(params, Xs)
    np.testing.assert_allclose(z[:8], golden["full_decision_head"], atol=1e-12)
    np.testing.assert_allclose(z[-8:], golden["full_decision_tail"], atol=1e-12)
    np.testing.assert_allclose(z.sum(), golden["full_decision_sum"], rtol=1e-12)
    np.testing.assert_allclose(scoring.predict_proba(params, Xs)[:, 1].sum(), golden["full_proba1_sum"], rtol=1e-12)
    assert int(scoring.predict(params, Xs).sum()) == int(golden["full_pred_sum"])


def test_edge_inputs(params):
    assert scoring.transform(params, np.zeros((0, 30))).shape == (0, 30)
    assert scoring.predict_proba(params, np.zeros((0, 30))).shape == (0, 2)
    one = scoring.predict_proba(params, np.zeros((1, 30)))
    np.testing.assert_allclose(one[0, 1], 1.0 / (1.0 + np.exp(-params.intercept)), rtol=1e-15)
    big = scoring.predict_proba(params, np.full((1, 30), 1e6))
    assert np.all(np.isfinite(big))


def test_linear_shap_additivity_on_logit(params, golden):
    """Quirk Q5: LinearExplainer explains the logit; sum(phi) + E = decision_function(x)."""
    Xs = golden["X_scaled"]
    mean, ev = linear_shap.explainer_state(params.coef, params.intercept, Xs)
    assert mean.shape == (30,)
    phi = linear_shap.shap_values(params.coef, mean, Xs)
    np.testing.assert_allclose(phi.sum(1) + ev, golden["decision"], atol=1e-10)
    # api/worker.py:52-53: a single all-zero background row -> phi = coef * x (xai_tasks.py:106-107)
    mean0, ev0 = linear_shap.explainer_state(params.coef, params.intercept, np.zeros((1, 30)))
    assert np.all(mean0 == 0) and ev0 == params.intercept
    np.testing.assert_array_equal(linear_shap.shap_values(params.coef, mean0, Xs[:4]),
                                  np.stack([linear_shap.coef_times_x(params.coef, r) for r in Xs[:4]]))
    mean_abs, top3 = linear_shap.global_summary(phi)
    assert mean_abs.shape == (30,) and len(top3) == 3 and mean_abs[top3[0]] == mean_abs.max()


def test_background_subsample_rule(golden):
    Xs = golden["X_scaled"]
    assert linear_shap.background_subsample(Xs[:100]) is not None
    assert linear_shap.background_subsample(Xs[:100]).shape == (100, 30)
    sub = linear_shap.background_subsample(Xs)
    assert sub.shape == (100, 30)
    # without replacement, deterministic (random_state=0)
    assert len({tuple(r) for r in sub}) == 100
    assert np.array_equal(sub, linear_shap.background_subsample(Xs))
