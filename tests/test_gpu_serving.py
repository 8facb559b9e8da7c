"""GPU: the micro-batcher and the XAI-task helper return what the one-row reference calls return."""

import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import linear_shap, scoring, synth  # noqa: E402


def test_microbatcher_equals_one_row_calls(params):
    import warnings
    from fraud_detection_b200 import B200LinearExplainer, wrap
    from fraud_detection_b200.serving import MicroBatcher
    warnings.filterwarnings("ignore")
    scaler_sk, model_sk, _ = scoring.live_estimators(params)
    scaler, model = wrap(scaler_sk), wrap(model_sk)
    expl = B200LinearExplainer(model, np.zeros((1, 30)))
    X = synth.rows(512, seed=4)
    Xs = scaler_sk.transform(X)
    want_p, want_c = model_sk.predict_proba(Xs), model_sk.predict(Xs)
    want_phi = linear_shap.shap_values(model_sk.coef_[0], np.zeros(30), Xs)
    res = [None] * len(X)
    with MicroBatcher(scaler, model, explainer=expl, max_batch=256, max_wait_us=5000) as mb:
        def client(lo, hi):
            for i in range(lo, hi):
                res[i] = mb.submit(X[i]).result(30)
        ts = [threading.Thread(target=client, args=(k * 64, (k + 1) * 64)) for k in range(8)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        assert mb.batches < len(X)
    for i, r in enumerate(res):
        assert r.prediction == want_c[i]
        np.testing.assert_allclose(r.proba, want_p[i], atol=1e-12)
        np.testing.assert_allclose(r.shap_values, want_phi[i], atol=1e-10)


def test_xai_task_helper_keeps_the_reference_quirks(params):
    from fraud_detection_b200 import wrap
    from fraud_detection_b200.serving import attributions_for_task
    scaler_sk, model_sk, _ = scoring.live_estimators(params)
    scaler, model = wrap(scaler_sk), wrap(model_sk)
    names = params.feature_names
    row = synth.rows(1, seed=5)[0]
    # keys the task actually receives from api/app.py:243 -> every lookup misses -> all-zero row (quirk Q2)
    p0, attr0 = attributions_for_task(model, names, {f"feature_{i}": float(v) for i, v in enumerate(row)})
    assert abs(p0 - 1.0 / (1.0 + np.exp(-params.intercept))) < 1e-12 and all(v == 0.0 for v in attr0.values())
    # matching keys: score of the UNSCALED row and coef * x, as xai_tasks.py:97,106-107 compute them
    data = dict(zip(names, row.tolist()))
    p1, attr1 = attributions_for_task(model, names, data)
    assert abs(p1 - model_sk.predict_proba(row[None])[0, 1]) < 1e-12
    np.testing.assert_allclose([attr1[n] for n in names], linear_shap.coef_times_x(model_sk.coef_, row), rtol=1e-12)
    # opt-in fix: apply the scaler first
    p2, attr2 = attributions_for_task(model, names, data, scaler=scaler, apply_scaler=True)
    xs = scaler_sk.transform(row[None])
    assert abs(p2 - model_sk.predict_proba(xs)[0, 1]) < 1e-12
    np.testing.assert_allclose([attr2[n] for n in names], model_sk.coef_[0] * xs[0], rtol=1e-10, atol=1e-12)


def test_score_file_end_to_end(params, tmp_path):
    """CSV -> pinned rows -> fused raw-row pass, against pandas + sklearn + numpy linear SHAP."""
    import pandas as pd
    from fraud_detection_b200 import io as fio
    from fraud_detection_b200.device import FraudScorer
    X = synth.reference_rows(50000)
    path = tmp_path / "creditcard.csv"
    pd.DataFrame(np.column_stack([X, np.zeros(len(X))]), columns=synth.FEATURE_NAMES + ["Class"]).to_csv(path, index=False)
    scorer = FraudScorer(params.coef, params.intercept, params.mean, params.scale)
    out = fio.score_file(str(path), scorer, dtype=np.float64)
    Xs = scoring.transform(params, X)                      # the CSV round-trips exactly (correctly rounded parser)
    np.testing.assert_allclose(out["proba"], scoring.predict_proba(params, Xs)[:, 1], atol=1e-12)
    phi_ref = linear_shap.shap_values(params.coef, np.zeros(30), Xs)
    np.testing.assert_allclose(out["phi"], phi_ref, atol=1e-9)
    ref_abs, ref_top3 = linear_shap.global_summary(phi_ref)
    np.testing.assert_allclose(out["mean_abs_shap"], ref_abs, rtol=1e-10)
    assert list(out["top3"]) == list(ref_top3) and out["rows"] == 50000


def test_paired_predict_calls_share_one_pass(params):
    """api/app.py:209,223 calls predict and predict_proba on the same request: request-size batches keep the fused
    result for byte-identical rows; returned arrays are the caller's own and different rows never hit the cache."""
    import warnings
    from fraud_detection_b200 import wrap
    warnings.filterwarnings("ignore")
    _, model_sk, _ = scoring.live_estimators(params)
    model = wrap(model_sk)
    rs = np.random.RandomState(5)
    for n in (1, 7, 64, 65, 300, 512, 513):
        Xs = rs.randn(n, 30) * 2.0
        c1 = model.predict(Xs)
        p1 = model.predict_proba(Xs)
        z1 = model.decision_function(Xs)
        np.testing.assert_array_equal(c1, model_sk.predict(Xs))
        np.testing.assert_allclose(p1, model_sk.predict_proba(Xs), atol=1e-12)
        np.testing.assert_allclose(z1, model_sk.decision_function(Xs), atol=1e-12)
        p1[:] = 7.0                                   # the caller owns what it got
        np.testing.assert_allclose(model.predict_proba(Xs), model_sk.predict_proba(Xs), atol=1e-12)
        Xs2 = Xs.copy()
        Xs2[0, 3] += 1.0                              # different bytes -> a fresh pass
        np.testing.assert_allclose(model.predict_proba(Xs2), model_sk.predict_proba(Xs2), atol=1e-12)
        X32 = Xs.astype(np.float32)                   # same values, other dtype -> its own pass and dtype rules
        np.testing.assert_allclose(model.predict_proba(X32), model_sk.predict_proba(X32), atol=1e-5)


def test_pipe_probes_report_plausible_rates():
    """fdb_probe_pipes (the measured denominators of the Kernel SHAP rooflines): on a B200 the SFU does 16 results per
    clock per SM, the FP32 pipe 128, and K2f's quad loop cannot beat the issue bound of 3.5 slots per sigmoid."""
    from fraud_detection_b200 import _lib
    p = _lib.probe_pipes(0)
    sms = p["sm_count"]
    assert sms >= 100
    per_clk = lambda rate, ghz: rate / (sms * ghz * 1e9)  # noqa: E731
    assert 10.0 < per_clk(p["ex2_per_s"], 1.965) < 17.0 and 10.0 < per_clk(p["rcp_per_s"], 1.965) < 17.0
    assert 90.0 < per_clk(p["ffma_per_s"], 1.965) < 130.0
    bound = p["ffma_per_s"] / 3.5
    assert 0.3 * bound < p["k2f_quad_16warps_sigmoids_per_s"] <= p["k2f_quad_lds_sigmoids_per_s"] * 1.05 < 1.1 * bound
