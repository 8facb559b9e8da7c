"""Serving micro-batcher for the /predict and XAI-task call sites.

The reference scores ONE row per HTTP request (api/app.py:178-240) and explains one row per
Celery task (xai_tasks.py:63-110, api/worker.py:65-99).  A GPU call costs ~25-60 us of launch
and copy latency whatever the batch size (profiles/r01/serving_latency.json), so concurrent
requests should share one call.  ``MicroBatcher`` does that above the drop-in classes, without
changing what a single caller sees:

    batcher = MicroBatcher(scaler, model, explainer=None, max_batch=4096, max_wait_us=200)
    pred, proba = batcher.score(features)            # blocking; same values as predict / predict_proba
    fut = batcher.submit(features)                    # concurrent.futures.Future -> ScoreResult
    res = await batcher.score_async(features)         # from the FastAPI event loop

One worker thread drains the queue: it takes the first waiting request, keeps collecting until
``max_batch`` rows are in hand or ``max_wait_us`` has passed, runs ONE fused pass
(``scaler.transform`` -> ``model.score_rows`` [+ ``explainer.shap_values``]) and resolves every
future with its own row.  Errors raised for a batch are re-checked row by row so that one
malformed request fails alone.

``explanation_records`` / ``json_rows`` produce what the explanation task stores per row
(api/worker.py:82-88: ``json.dumps(shap_values[0].tolist())``) for a whole batch, with the text
formatted in the library - byte for byte what ``json.dumps`` writes.

``attributions_for_task`` restates the row preparation of the deployed XAI task
(xai_tasks.py:86-107), including its quirks (missing keys read as 0.0, no scaler - Q2 in
SURVEY.md section 3); ``apply_scaler=True`` is the opt-in fix.
"""

from __future__ import annotations

import queue
import threading
import time
from concurrent.futures import Future
from dataclasses import dataclass

import numpy as np


@dataclass
class ScoreResult:
    prediction: int
    proba: np.ndarray                 # (2,) float64, [1 - p, p] like predict_proba(X)[0]
    shap_values: np.ndarray | None    # (30,) when an explainer is attached


class MicroBatcher:
    def __init__(self, scaler, model, explainer=None, max_batch: int = 4096, max_wait_us: float = 200.0,
                 explain_scaled: bool = True):
        self.scaler, self.model, self.explainer = scaler, model, explainer
        self.max_batch, self.max_wait = int(max_batch), float(max_wait_us) * 1e-6
        self.explain_scaled = explain_scaled
        self.n_features = int(model.n_features_in_)
        self._q: queue.Queue = queue.Queue()
        self._stop = threading.Event()
        self.batches = 0
        self.rows = 0
        self._thread = threading.Thread(target=self._run, name="fdb-microbatcher", daemon=True)
        self._thread.start()

    # ------------------------------------------------------------------ client side
    def submit(self, features) -> Future:
        row = np.asarray(features, dtype=np.float64).reshape(-1)
        fut: Future = Future()
        if row.shape[0] != self.n_features:
            # same wording as the 422 of api/app.py:187-192
            fut.set_exception(ValueError(f"Input data must have {self.n_features} features, but got {row.shape[0]}."))
            return fut
        self._q.put((row, fut))
        return fut

    def score(self, features, timeout: float | None = None):
        r = self.submit(features).result(timeout)
        return r.prediction, r.proba

    async def score_async(self, features) -> ScoreResult:
        import asyncio
        return await asyncio.wrap_future(self.submit(features))

    def close(self):
        self._stop.set()
        self._q.put(None)
        self._thread.join(timeout=5.0)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ------------------------------------------------------------------ worker side
    def _collect(self):
        first = self._q.get()
        if first is None:
            return None
        items = [first]
        deadline = time.perf_counter() + self.max_wait
        while len(items) < self.max_batch:
            left = deadline - time.perf_counter()
            try:
                nxt = self._q.get_nowait() if left <= 0 else self._q.get(timeout=left)
            except queue.Empty:
                break
            if nxt is None:
                self._stop.set()
                break
            items.append(nxt)
        return items

    def _score_block(self, X: np.ndarray):
        Xs = self.scaler.transform(X) if self.scaler is not None else X
        pred, proba = self.model.score_rows(Xs)
        phi = None
        if self.explainer is not None:
            phi = self.explainer.shap_values(Xs if self.explain_scaled else X)
        return pred, proba, phi

    def _run(self):
        import warnings
        warnings.filterwarnings("ignore", message="X does not have valid feature names")
        while not self._stop.is_set():
            items = self._collect()
            if not items:
                break
            X = np.stack([r for r, _ in items])
            try:
                pred, proba, phi = self._score_block(X)
                for i, (_, fut) in enumerate(items):
                    fut.set_result(ScoreResult(int(pred[i]), proba[i].copy(), None if phi is None else phi[i].copy()))
            except Exception:
                for row, fut in items:        # isolate the offending request(s)
                    try:
                        p1, pr1, ph1 = self._score_block(row[None, :])
                        fut.set_result(ScoreResult(int(p1[0]), pr1[0].copy(), None if ph1 is None else ph1[0].copy()))
                    except Exception as exc:  # noqa: BLE001 - handed to the caller
                        fut.set_exception(exc)
            self.batches += 1
            self.rows += len(items)


def attributions_for_task(model, feature_names, input_data: dict, scaler=None, apply_scaler: bool = False):
    """Row preparation + score + attribution of the deployed XAI task (xai_tasks.py:86-110).

    Returns ``(pred_proba, {feature: attribution})``.  Quirks kept on purpose: keys missing from
    ``input_data`` read as 0.0 (api/app.py:243 sends ``feature_i`` keys while the task looks up
    ``Time, V1, ...``), the scaler is NOT applied unless ``apply_scaler=True``, and the attribution is
    ``coef_ * x`` (LinearExplainer against a zero background).
    """
    if feature_names:
        x = [float(input_data.get(f, 0.0)) for f in feature_names]
    else:
        x = [float(v) for _, v in sorted(input_data.items())]
        feature_names = [k for k, _ in sorted(input_data.items())]
    x_arr = np.array([x])
    if apply_scaler and scaler is not None:
        x_arr = scaler.transform(x_arr)
    # one fused GPU pass: proba and coef_ * x (linear SHAP against a zero background) together
    from . import _lib
    x_arr = np.ascontiguousarray(x_arr, dtype=np.float64)
    proba, phi = np.empty((1,), dtype=np.float64), np.empty_like(x_arr)
    _lib.check(_lib.lib.fdb_linear_pass_host(
        _lib.get_ctx(getattr(model, "device", None)), _lib.F64, 1, _lib.np_ptr(x_arr), None,
        _lib.dvec(np.asarray(model.coef_).ravel()), float(np.asarray(model.intercept_).ravel()[0]),
        _lib.np_ptr(proba), 1, None, _lib.np_ptr(phi), None, None))
    return float(proba[0]), {f: float(a) for f, a in zip(feature_names, phi[0].tolist())}


def json_rows(values) -> list:
    """``[json.dumps(row.tolist()) for row in values]`` for a 2-D array, formatted by the library
    (``fdb_format_json_rows``: shortest round-trip digits laid out as ``float.__repr__`` does) -
    the text the explanation task stores (api/worker.py:86)."""
    from . import _lib
    A = np.ascontiguousarray(values, dtype=np.float64)
    if A.ndim != 2:
        raise ValueError(f"expected a 2-D array of attribution rows, got shape {A.shape}")
    n, c = A.shape
    buf = np.empty(max(n * (c * 28 + 2), 1), dtype=np.uint8)
    off = np.empty(n + 1, dtype=np.int64)
    _lib.check(_lib.lib.fdb_format_json_rows(_lib.np_ptr(A), n, c, _lib.np_ptr(buf), buf.size, _lib.np_ptr(off)))
    text = buf[:int(off[-1])].tobytes().decode("ascii")
    o = off.tolist()
    return [text[o[i]:o[i + 1]] for i in range(n)]


def explanation_records(transaction_ids, correlation_ids, shap_values, feature_names) -> list:
    """The rows the explanation task hands to its store (api/worker.py:82-88), for a whole batch:
    ``{'transaction_id', 'correlation_id', 'shap_values': json.dumps(row.tolist()),
    'feature_names': json.dumps(list(X.columns))}`` per explained row.  ``shap_values`` is the
    ``(N, 30)`` result of ``explainer.shap_values``; a single id / 1-D row gives one record."""
    import json
    phi = np.asarray(shap_values, dtype=np.float64)
    if phi.ndim == 1:
        phi = phi[None, :]
        transaction_ids, correlation_ids = [transaction_ids], [correlation_ids]
    if not (len(transaction_ids) == len(correlation_ids) == phi.shape[0]):
        raise ValueError("one transaction id and one correlation id per explained row")
    names = json.dumps(list(feature_names))
    return [{"transaction_id": t, "correlation_id": c, "shap_values": s, "feature_names": names}
            for t, c, s in zip(transaction_ids, correlation_ids, json_rows(phi))]
