"""GPU parity: K1 (fused scale + dot + sigmoid + linear SHAP) against the oracle.

Every case goes through the C ABI of libfdb200.so - either the host-buffer entry points
(via the drop-in classes) or the device-pointer entry points (via fraud_detection_b200.device).
Tolerances follow BASELINE.json's north_star: probabilities 1e-5 absolute, linear SHAP 1e-5
(float32 outputs cannot hold 1e-5 absolute once |phi| >= 128, so the float32 bar is
max(1e-5, 2 ulp_f32(|phi|)) as SURVEY.md section 7 derives; float64 runs are checked at 1e-12).
"""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import linear_shap, scoring, synth  # noqa: E402


@pytest.fixture(scope="module")
def wrapped(params):
    from fraud_detection_b200 import wrap
    scaler_sk, model_sk, _ = scoring.live_estimators(params)
    return wrap(scaler_sk), wrap(model_sk), scaler_sk, model_sk


def _phi_tol(ref):
    return np.maximum(1e-5, 2 * np.spacing(np.abs(ref).astype(np.float32)).astype(np.float64))


# --------------------------------------------------------------------------- golden vectors
def test_golden_vectors_float64(wrapped, golden):
    scaler, model, _, _ = wrapped
    Xs = scaler.transform(golden["X"])
    assert Xs.dtype == np.float64
    np.testing.assert_allclose(Xs, golden["X_scaled"], rtol=0, atol=1e-12)
    proba = model.predict_proba(golden["X_scaled"])
    assert proba.shape == (512, 2) and proba.dtype == np.float64
    np.testing.assert_allclose(proba, golden["proba"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(model.decision_function(golden["X_scaled"]), golden["decision"], rtol=0, atol=1e-10)
    pred = model.predict(golden["X_scaled"])
    assert pred.dtype == np.int64 and np.array_equal(pred, golden["pred"])


def test_golden_vectors_float32(wrapped, golden):
    scaler, model, _, _ = wrapped
    X32 = golden["X"].astype(np.float32)
    Xs = scaler.transform(X32)
    assert Xs.dtype == np.float32
    np.testing.assert_allclose(Xs, scoring.transform(scoring.load_params(), X32.astype(np.float64)), rtol=2e-6, atol=1e-6)
    Xs32 = golden["X_scaled"].astype(np.float32)
    proba = model.predict_proba(Xs32)
    assert proba.dtype == np.float64
    want = scoring.predict_proba(scoring.load_params(), Xs32.astype(np.float64))
    np.testing.assert_allclose(proba, want, rtol=0, atol=1e-5)


# --------------------------------------------------------------------------- config 1
@pytest.mark.parametrize("dtype,atol", [(np.float64, 1e-12), (np.float32, 1e-5)])
def test_config1_full_size_vs_live_sklearn(wrapped, golden, dtype, atol):
    """284,807 x 30 draw from the reference generator, reference call chain, both dtypes."""
    scaler, model, scaler_sk, model_sk = wrapped
    X = synth.reference_rows(int(golden["full_n"])).astype(dtype)
    Xs = scaler.transform(X)
    Xs_ref = scaler_sk.transform(X.astype(np.float64))
    np.testing.assert_allclose(Xs, Xs_ref, rtol=0, atol=1e-12 if dtype == np.float64 else 2e-5)
    proba = model.predict_proba(Xs)
    np.testing.assert_allclose(proba, model_sk.predict_proba(Xs.astype(np.float64)), rtol=0, atol=atol)
    pred, proba2 = model.score_rows(Xs)
    assert np.array_equal(proba2, proba)
    z_ref = model_sk.decision_function(Xs.astype(np.float64))
    decided = np.abs(z_ref) > 1e-4     # rows on the boundary may flip in float32
    assert np.array_equal(pred[decided], model_sk.predict(Xs.astype(np.float64))[decided])
    if dtype == np.float64:
        np.testing.assert_allclose(proba[:, 1].sum(), golden["full_proba1_sum"], rtol=1e-12)
        assert int(pred.sum()) == int(golden["full_pred_sum"])


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_linear_explainer_vs_oracle(wrapped, golden, dtype):
    from fraud_detection_b200 import B200LinearExplainer
    scaler, model, scaler_sk, model_sk = wrapped
    Xs = scaler_sk.transform(synth.reference_rows(20000)).astype(dtype)
    expl = B200LinearExplainer(model, Xs, feature_perturbation="interventional")
    mean, ev = linear_shap.explainer_state(model_sk.coef_[0], model_sk.intercept_[0], Xs)
    np.testing.assert_allclose(expl.mean, mean, rtol=0, atol=1e-15)
    assert abs(expl.expected_value - ev) < 1e-12
    phi = expl.shap_values(Xs)
    ref = linear_shap.shap_values(model_sk.coef_[0], mean, Xs.astype(np.float64))
    assert phi.shape == ref.shape and phi.dtype == np.float64
    if dtype == np.float64:
        np.testing.assert_allclose(phi, ref, rtol=0, atol=1e-12)
    else:
        assert np.all(np.abs(phi - ref) <= _phi_tol(ref))
    # additivity on the logit (quirk Q5)
    z = model_sk.decision_function(Xs.astype(np.float64))
    np.testing.assert_allclose(phi.sum(1) + expl.expected_value, z, rtol=0, atol=1e-10 if dtype == np.float64 else 2e-4)
    phi2, mean_abs, top3 = expl.explain_with_summary(Xs)
    assert np.array_equal(phi2, phi)
    ref_abs, ref_top3 = linear_shap.global_summary(ref)
    np.testing.assert_allclose(mean_abs, ref_abs, rtol=1e-12 if dtype == np.float64 else 1e-5)
    assert list(top3) == list(ref_top3)
    # single row (api/worker.py:70-75 passes a 1-row DataFrame; callers index [0])
    import pandas as pd
    one = B200LinearExplainer(model, pd.DataFrame([[0] * 30])).shap_values(pd.DataFrame([Xs[0].astype(np.float64)]))
    np.testing.assert_allclose(one[0], linear_shap.coef_times_x(model_sk.coef_, Xs[0].astype(np.float64)), atol=1e-12)


# --------------------------------------------------------------------------- edges
@pytest.mark.parametrize("n", [0, 1, 2, 3, 4, 5, 7, 8, 63, 64, 65, 1023, 4097, 16384, 16385, 70001])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_ragged_sizes(wrapped, n, dtype):
    """Tail rows (n % 4, n % 2), the serving/bulk switch at 16384 rows, empty input."""
    scaler, model, scaler_sk, model_sk = wrapped
    X = synth.rows(n, seed=n + 1).astype(dtype)
    Xs = scaler.transform(X)
    assert Xs.shape == (n, 30)
    ref = scaler_sk.transform(X.astype(np.float64)) if n else np.zeros((0, 30))
    np.testing.assert_allclose(Xs, ref, rtol=0, atol=1e-12 if dtype == np.float64 else 2e-5)
    proba = model.predict_proba(Xs)
    assert proba.shape == (n, 2)
    if n:
        np.testing.assert_allclose(proba, model_sk.predict_proba(Xs.astype(np.float64)), rtol=0,
                                   atol=1e-12 if dtype == np.float64 else 1e-5)
        np.testing.assert_allclose(proba.sum(1), 1.0, atol=1e-6)
    assert model.predict(Xs).shape == (n,)


def test_extreme_logits_saturate_cleanly(wrapped):
    _, model, _, model_sk = wrapped
    X = np.zeros((4, 30)); X[0, :] = 60.0; X[1, :] = -60.0; X[2, 28] = 1e4; X[3, 28] = -1e4
    for dt in (np.float32, np.float64):
        p = model.predict_proba(X.astype(dt))
        assert np.all(np.isfinite(p))
        np.testing.assert_allclose(p, model_sk.predict_proba(X), atol=1e-6)


def test_errors_are_python_exceptions(wrapped):
    scaler, model, _, _ = wrapped
    with pytest.raises(ValueError, match="features"):
        model.predict_proba(np.zeros((3, 29)))
    with pytest.raises(ValueError, match="NaN"):
        model.predict(np.full((2, 30), np.nan))
    big = np.zeros((40000, 30)); big[12345, 7] = np.inf          # above the host-scan threshold
    with pytest.raises(ValueError, match="infinity"):
        model.predict_proba(big)
    big[12345, 7] = np.nan
    with pytest.raises(ValueError, match="NaN"):
        model.predict_proba(big)


def test_serving_call_shapes_parse_like_the_reference(wrapped):
    """api/app.py:209-231 parses str(predict(X)) and the 2nd number of str(predict_proba(X))."""
    import re
    import pandas as pd
    scaler, model, scaler_sk, model_sk = wrapped
    raw = pd.DataFrame([[0.1] * 30])
    with pytest.warns(UserWarning):
        scaled = scaler.transform(raw)
    X = pd.DataFrame(scaled, columns=[f"feature_{i}" for i in range(30)])
    with pytest.warns(UserWarning):
        pred, proba = model.predict(X), model.predict_proba(X)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        pred_ref = model_sk.predict(X)
        proba_ref = model_sk.predict_proba(X)
    assert str(pred) == str(pred_ref)
    ours = re.findall(r"[0-9]+\.?[0-9]*", str(proba))
    theirs = re.findall(r"[0-9]+\.?[0-9]*", str(proba_ref))
    assert abs(float(ours[1]) - float(theirs[1])) < 1e-8


# --------------------------------------------------------------------------- device entry points
@pytest.mark.parametrize("n", [4, 1000, 1000003])
def test_device_fused_raw_pipeline(params, n):
    """FraudScorer: raw rows -> proba + phi + cls + |phi| sums in one launch (config 2 path)."""
    from fraud_detection_b200.device import FraudScorer, linear_pass, transform
    X = synth.rows(n, seed=3).astype(np.float32)
    Xs = scoring.transform(params, X.astype(np.float64))
    mean_bg, ev = linear_shap.explainer_state(params.coef, params.intercept, Xs[:5000])
    scorer = FraudScorer(params.coef, params.intercept, params.mean, params.scale, mean_bg)
    assert abs(scorer.expected_value - ev) < 1e-12
    x = torch.from_numpy(X).cuda()
    out = scorer(x, phi=True, cls=True, abs_sum=True)
    torch.cuda.synchronize()
    p_ref = scoring.predict_proba(params, Xs)[:, 1]
    phi_ref = linear_shap.shap_values(params.coef, mean_bg, Xs)
    np.testing.assert_allclose(out.proba.cpu().numpy(), p_ref, rtol=0, atol=1e-5)
    phi = out.phi.cpu().numpy().astype(np.float64)
    assert np.all(np.abs(phi - phi_ref) <= _phi_tol(phi_ref))
    z_ref = scoring.decision_function(params, Xs)
    decided = np.abs(z_ref) > 1e-3
    assert np.array_equal(out.cls.cpu().numpy()[decided], (z_ref > 0)[decided].astype(np.uint8))
    abs_sum = out.abs_sum.cpu().numpy()
    assert abs_sum[30] == n and abs_sum[31] == 0
    np.testing.assert_allclose(abs_sum[:30], np.abs(phi).sum(0), rtol=1e-6)
    np.testing.assert_allclose(abs_sum[:30], np.abs(phi_ref).sum(0), rtol=1e-5)
    # deterministic reduction: a second launch reproduces the sums bit for bit
    out2 = scorer(x, phi=False, abs_sum=True)
    torch.cuda.synchronize()
    assert np.array_equal(out2.abs_sum.cpu().numpy(), abs_sum)
    # the unfused chain (transform, then score scaled rows) agrees with the fused one
    xs = transform(x, params.mean, params.scale)
    o3 = linear_pass(xs, None, params.coef, params.intercept, proba=True, proba_cols=2, logit=True)
    torch.cuda.synchronize()
    np.testing.assert_allclose(o3.proba.cpu().numpy()[:, 1], p_ref, rtol=0, atol=1e-5)
    np.testing.assert_allclose(o3.proba.cpu().numpy().sum(1), 1.0, atol=1e-6)
    np.testing.assert_allclose(o3.logit.cpu().numpy(), z_ref, rtol=0, atol=2e-4)


def test_device_rejects_bad_tensors(params):
    from fraud_detection_b200.device import linear_pass
    x = torch.zeros((9, 30), device="cuda")
    with pytest.raises(ValueError, match="16-byte"):
        linear_pass(x[1:], None, params.coef, 0.0)          # 120-byte offset -> 8-byte aligned only
    with pytest.raises(ValueError):
        linear_pass(torch.zeros((4, 29), device="cuda"), None, params.coef, 0.0)
    with pytest.raises(ValueError):
        linear_pass(torch.zeros((4, 30)), None, params.coef, 0.0)


def test_host_entry_point_with_pinned_buffers(params):
    """The e2e path bench.py times: pinned host in/out through fdb_linear_pass_host."""
    from fraud_detection_b200 import PinnedBuffer
    from fraud_detection_b200.device import FraudScorer
    n = 300001
    X = synth.rows(n, seed=9).astype(np.float32)
    xin, pout, phiout = PinnedBuffer((n, 30), np.float32), PinnedBuffer((n,), np.float32), PinnedBuffer((n, 30), np.float32)
    xin.array[:] = X
    scorer = FraudScorer(params.coef, params.intercept, params.mean, params.scale)
    out = scorer.score_host(xin.array, phi=True, abs_sum=True, out={"proba": pout.array, "phi": phiout.array})
    Xs = scoring.transform(params, X.astype(np.float64))
    np.testing.assert_allclose(out["proba"], scoring.predict_proba(params, Xs)[:, 1], rtol=0, atol=1e-5)
    phi_ref = Xs * params.coef
    assert np.all(np.abs(out["phi"] - phi_ref) <= _phi_tol(phi_ref))
    np.testing.assert_allclose(out["abs_sum"][:30], np.abs(phi_ref).sum(0), rtol=1e-5)
    assert out["abs_sum"][30] == n


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("n", [16384, 16385, 3 * 262144 + 5])
def test_host_entry_point_every_output_at_once(params, dtype, n):
    """fdb_linear_pass_host straight through the C ABI with pageable buffers and every output requested: the
    copy-free serving path (<= 16,384 rows) and the chunked pipeline (more) against the oracle."""
    from fraud_detection_b200 import _lib
    Xs = scoring.transform(params, synth.rows(n, seed=17)).astype(dtype)
    mean_bg = Xs[:100].astype(np.float64).mean(0)
    ev = float(params.coef @ mean_bg + params.intercept)
    proba = np.empty((n, 2), dtype=dtype)
    logit = np.empty(n, dtype=dtype)
    phi = np.empty((n, 30), dtype=dtype)
    cls = np.empty(n, dtype=np.uint8)
    abs_sum = np.zeros(32)
    _lib.check(_lib.lib.fdb_linear_pass_host(_lib.get_ctx(None), _lib.dtype_code(np.dtype(dtype)), n, _lib.np_ptr(Xs),
                                             _lib.dvec(mean_bg), _lib.dvec(params.coef), ev, _lib.np_ptr(proba), 2,
                                             _lib.np_ptr(logit), _lib.np_ptr(phi), _lib.np_ptr(cls), _lib.np_ptr(abs_sum)))
    X64 = Xs.astype(np.float64)
    z_ref = scoring.decision_function(params, X64)
    p_ref = scoring.predict_proba(params, X64)
    phi_ref = linear_shap.shap_values(params.coef, mean_bg, X64)
    tol = 1e-12 if dtype == np.float64 else 1e-5
    np.testing.assert_allclose(proba, p_ref, rtol=0, atol=tol)
    np.testing.assert_allclose(logit, z_ref, rtol=0, atol=1e-11 if dtype == np.float64 else 2e-4)
    assert np.all(np.abs(phi.astype(np.float64) - phi_ref) <= (1e-12 if dtype == np.float64 else _phi_tol(phi_ref)))
    decided = np.abs(z_ref) > 1e-3
    assert np.array_equal(cls[decided], (z_ref > 0)[decided].astype(np.uint8))
    assert abs_sum[30] == n
    np.testing.assert_allclose(abs_sum[:30], np.abs(phi_ref).sum(0), rtol=1e-9 if dtype == np.float64 else 1e-5)


def test_serving_path_takes_unaligned_pinned_rows(params):
    """Page-locked caller memory is read in place by the copy-free path - unless it is not 16-byte aligned."""
    from fraud_detection_b200 import PinnedBuffer, _lib
    n = 37
    buf = PinnedBuffer((n * 30 + 2,), np.float64)
    X = scoring.transform(params, synth.rows(n, seed=18))
    for off in (0, 1):                              # 0: aligned, read in place; 1: 8-byte offset, staged
        view = buf.array[off:off + n * 30].reshape(n, 30)
        view[:] = X
        p = np.empty(n)
        _lib.check(_lib.lib.fdb_linear_pass_host(_lib.get_ctx(None), _lib.F64, n, _lib.np_ptr(view), None, _lib.dvec(params.coef),
                                                 float(params.intercept), _lib.np_ptr(p), 1, None, None, None, None))
        np.testing.assert_allclose(p, scoring.predict_proba(params, X)[:, 1], rtol=0, atol=1e-12)


def test_host_and_device_calls_on_one_context_do_not_share_scratch(params):
    """A _dev launch (torch stream) and a _host call of the same context run at the same time: each keeps its own
    partial-sum scratch and arrival ticket, so both |phi| summaries stay right however they interleave."""
    import threading

    from fraud_detection_b200.device import FraudScorer
    scorer = FraudScorer(params.coef, params.intercept, params.mean, params.scale)
    Xd = synth.rows(3_000_000, seed=5).astype(np.float32)
    Xh = synth.rows(9000, seed=6).astype(np.float32)
    xd = torch.from_numpy(Xd).cuda()
    want_d = scorer(xd, abs_sum=True).abs_sum.cpu().numpy()
    want_h = scorer.score_host(Xh, abs_sum=True)["abs_sum"].copy()
    stop = threading.Event()
    bad = []

    def host_loop():
        while not stop.is_set():
            got = scorer.score_host(Xh, abs_sum=True)["abs_sum"]
            if not np.array_equal(got, want_h):
                bad.append(got.copy())

    th = threading.Thread(target=host_loop)
    th.start()
    try:
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            for _ in range(60):
                got = scorer(xd, abs_sum=True).abs_sum.cpu().numpy()
                assert np.array_equal(got, want_d)
    finally:
        stop.set()
        th.join()
    assert not bad
