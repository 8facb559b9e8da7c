"""Micro-batcher host logic on CPU with stand-in estimators (the oracle math), no GPU needed."""

import threading
import time

import numpy as np
import pytest

from fraud_detection_b200.serving import MicroBatcher
from oracle import scoring, synth


class _Scaler:
    def __init__(self, p):
        self.p, self.calls = p, 0

    def transform(self, X):
        self.calls += 1
        return scoring.transform(self.p, X)


class _Model:
    n_features_in_ = 30

    def __init__(self, p):
        self.p, self.batch_sizes = p, []

    def score_rows(self, Xs):
        if not np.all(np.isfinite(Xs)):
            raise ValueError("Input X contains NaN.")
        self.batch_sizes.append(len(Xs))
        return scoring.predict(self.p, Xs), scoring.predict_proba(self.p, Xs)


def test_concurrent_requests_share_calls_and_get_their_own_rows(params):
    scaler, model = _Scaler(params), _Model(params)
    X = synth.rows(400, seed=1)
    want_pred = scoring.predict(params, scoring.transform(params, X))
    want_proba = scoring.predict_proba(params, scoring.transform(params, X))
    with MicroBatcher(scaler, model, max_batch=64, max_wait_us=20000) as mb:
        out = [None] * len(X)

        def client(i0, i1):
            for i in range(i0, i1):
                out[i] = mb.score(X[i], timeout=30)
        threads = [threading.Thread(target=client, args=(k * 50, (k + 1) * 50)) for k in range(8)]
        [t.start() for t in threads]
        [t.join() for t in threads]
        assert mb.rows == 400 and mb.batches < 400            # requests were coalesced
    assert max(model.batch_sizes) <= 64
    for i, (pred, proba) in enumerate(out):
        assert pred == want_pred[i]
        np.testing.assert_array_equal(proba, want_proba[i])


def test_bad_request_fails_alone(params):
    with MicroBatcher(_Scaler(params), _Model(params), max_batch=8, max_wait_us=50000) as mb:
        good = [mb.submit(r) for r in synth.rows(3, seed=2)]
        bad_nan = mb.submit(np.full(30, np.nan))
        bad_len = mb.submit(np.zeros(29))
        with pytest.raises(ValueError, match="must have 30 features"):
            bad_len.result(10)
        with pytest.raises(ValueError, match="NaN"):
            bad_nan.result(10)
        for f in good:
            assert f.result(10).proba.shape == (2,)


def test_single_request_latency_is_bounded_by_max_wait(params):
    with MicroBatcher(_Scaler(params), _Model(params), max_batch=4096, max_wait_us=2000) as mb:
        mb.score(np.zeros(30))
        best = float("inf")
        for _ in range(5):                      # best of five: a loaded CI box may stall any single call
            t0 = time.perf_counter()
            mb.score(np.ones(30))
            best = min(best, time.perf_counter() - t0)
        assert best < 0.2


def test_asyncio_entry_point(params):
    import asyncio

    async def go():
        with MicroBatcher(_Scaler(params), _Model(params), max_wait_us=1000) as mb:
            res = await asyncio.gather(*[mb.score_async(r) for r in synth.rows(20, seed=3)])
            return res
    res = asyncio.run(go())
    assert len(res) == 20 and all(r.proba.shape == (2,) for r in res)


def test_json_rows_is_byte_for_byte_json_dumps():
    """The result sink of the explanation task (api/worker.py:82-88) stores json.dumps(shap_values[0].tolist()); the
    library's formatter must write exactly that text (host code - runs without a GPU)."""
    import json
    from fraud_detection_b200.serving import explanation_records, json_rows
    rng = np.random.default_rng(0)
    A = rng.standard_normal((3000, 30)) * np.exp(rng.uniform(-40, 40, (3000, 30)))
    A[0, :14] = [0.0, -0.0, 1e-4, 1e-5, 9.999e-5, 1e15, 1e16, 123456789012345678.0, 5e-324, 1.7976931348623157e308,
                 np.nan, np.inf, -np.inf, 2.2250738585072014e-308]
    A[1, :8] = [1.0, -1.5, 100.0, 0.1, 1 / 3, 1e22, 1e23, 0.30000000000000004]
    A[2] = np.float32(rng.standard_normal(30)).astype(np.float64)          # float32 attributions widened, as tolist() does
    assert json_rows(A) == [json.dumps(r.tolist()) for r in A]
    assert json_rows(np.zeros((0, 30))) == [] and json_rows(np.zeros((2, 0))) == ["[]", "[]"]
    with pytest.raises(ValueError):
        json_rows(np.zeros(30))
    names = [f"feature_{i}" for i in range(30)]
    recs = explanation_records(["t0", "t1"], ["c0", "c1"], A[:2], names)
    assert recs[1] == {"transaction_id": "t1", "correlation_id": "c1", "shap_values": json.dumps(A[1].tolist()),
                       "feature_names": json.dumps(names)}
    one = explanation_records("t", "c", A[2], names)
    assert len(one) == 1 and json.loads(one[0]["shap_values"]) == A[2].tolist()
    with pytest.raises(ValueError):
        explanation_records(["t0"], ["c0", "c1"], A[:2], names)
