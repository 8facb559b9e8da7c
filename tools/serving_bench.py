#!/usr/bin/env python
"""Config 5 - serving micro-batch latency through the drop-in classes (H2D/D2H included).

Times exactly the brackets the reference instruments with Prometheus histograms:
  * ``api_inference_duration_seconds``  api/app.py:208-240: MODEL.predict(X) + MODEL.predict_proba(X)
    on a pandas DataFrame of scaled features (after SCALER.transform at :194);
  * ``xai_task_duration_seconds``       api/worker.py:74-75: EXPL.shap_values(X) on a DataFrame;
for batch sizes 1..4096, p50/p99 over N calls, next to live sklearn (+ the numpy LinearExplainer
restatement) in the same loop on the same host.  HTTP, Redis and Postgres are not part of the
bracket (and not installable here).

    python tools/serving_bench.py [--calls 10000] [--out profiles/r01/serving_latency.json]
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import time
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _percentiles(fn, calls, warm=50):
    for _ in range(warm):
        fn()
    ts = np.empty(calls)
    for i in range(calls):
        t0 = time.perf_counter()
        fn()
        ts[i] = time.perf_counter() - t0
    return {"p50_ms": float(np.percentile(ts, 50) * 1e3), "p99_ms": float(np.percentile(ts, 99) * 1e3),
            "mean_ms": float(ts.mean() * 1e3)}


def run(batches=(1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096), calls=10000, device=None):
    import pandas as pd
    from fraud_detection_b200 import B200LinearExplainer, wrap
    from oracle import linear_shap, scoring, synth

    warnings.filterwarnings("ignore")
    scaler_sk, model_sk, origin = scoring.live_estimators()
    scaler, model = wrap(scaler_sk, device=device), wrap(model_sk, device=device)
    names = list(getattr(scaler_sk, "feature_names_in_", synth.FEATURE_NAMES))
    zero_bg = pd.DataFrame([[0] * 30], columns=names)
    expl = B200LinearExplainer(model, zero_bg, device=device)            # api/worker.py:52-53
    coef = model_sk.coef_[0]
    out = {"estimators": origin, "calls": calls, "host_cpus": os.cpu_count(), "rows": []}
    for bsz in batches:
        # every call gets a DIFFERENT request (a ring of distinct batches): the drop-in keeps the fused result of a
        # request-size batch for the paired predict / predict_proba calls, so repeating one request would time a
        # dictionary look-up instead of the device path
        ring = max(2, min(64, (1 << 16) // bsz))
        raws = [pd.DataFrame(synth.rows(bsz, seed=1000 * bsz + r)) for r in range(ring)]
        Xss = [pd.DataFrame(scaler_sk.transform(raw), columns=names) for raw in raws]              # api/app.py:194-197
        Xws = [pd.DataFrame(raw.values, columns=[f"feature_{i}" for i in range(30)]) for raw in raws]   # api/worker.py:70
        n = max(200, calls // max(1, bsz // 64)) if bsz > 64 else calls
        row = {"batch": bsz, "calls": n, "distinct_requests": ring}
        tick = [0]

        def nxt(seq):
            tick[0] += 1
            return seq[tick[0] % ring]

        def pair(m):
            Xs = nxt(Xss)
            return m.predict(Xs), m.predict_proba(Xs)

        row["b200_predict_bracket"] = _percentiles(lambda: pair(model), n)
        row["b200_score_rows"] = _percentiles(lambda: model.score_rows(nxt(Xss)), n)
        row["sklearn_predict_bracket"] = _percentiles(lambda: pair(model_sk), n)
        row["b200_shap_bracket"] = _percentiles(lambda: expl.shap_values(nxt(Xws)), n)
        row["cpu_shap_bracket"] = _percentiles(lambda: linear_shap.shap_values(coef, np.zeros(30), nxt(Xws)), n)
        row["b200_transform"] = _percentiles(lambda: scaler.transform(nxt(raws)), n)
        row["sklearn_transform"] = _percentiles(lambda: scaler_sk.transform(nxt(raws)), n)
        out["rows"].append(row)
    return out


def run_microbatch(n_requests=20000, clients=(1, 8, 64), device=None):
    """Requests/s for one-row requests: the reference pattern (one sklearn call chain per request), the drop-in
    called once per request, and the drop-in behind MicroBatcher with concurrent client threads."""
    import threading
    from fraud_detection_b200 import wrap
    from fraud_detection_b200.serving import MicroBatcher
    from oracle import scoring, synth

    warnings.filterwarnings("ignore")
    scaler_sk, model_sk, origin = scoring.live_estimators()
    scaler, model = wrap(scaler_sk, device=device), wrap(model_sk, device=device)
    X = synth.rows(n_requests, seed=7)
    out = {"requests": n_requests, "estimators": origin}
    n_seq = min(n_requests, 3000)
    t0 = time.perf_counter()
    for i in range(n_seq):
        xs = scaler_sk.transform(X[i:i + 1]); model_sk.predict(xs); model_sk.predict_proba(xs)
    out["sklearn_one_row_rps"] = n_seq / (time.perf_counter() - t0)
    t0 = time.perf_counter()
    for i in range(n_seq):
        model.score_rows(scaler.transform(X[i:i + 1]))
    out["b200_one_row_rps"] = n_seq / (time.perf_counter() - t0)
    out["microbatcher"] = []
    for nc in clients:
        with MicroBatcher(scaler, model, max_batch=4096, max_wait_us=200) as mb:
            mb.score(X[0])
            per = n_requests // nc
            lat = [None] * nc

            def client(c):
                ts = np.empty(per)
                for k in range(per):
                    t1 = time.perf_counter()
                    mb.score(X[c * per + k])
                    ts[k] = time.perf_counter() - t1
                lat[c] = ts
            th = [threading.Thread(target=client, args=(c,)) for c in range(nc)]
            t0 = time.perf_counter()
            [t.start() for t in th]
            [t.join() for t in th]
            dt = time.perf_counter() - t0
            allv = np.concatenate(lat)
            out["microbatcher"].append({"clients": nc, "rps": per * nc / dt, "p50_ms": float(np.percentile(allv, 50) * 1e3),
                                        "p99_ms": float(np.percentile(allv, 99) * 1e3), "mean_batch": mb.rows / max(mb.batches, 1)})
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--calls", type=int, default=10000)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    res = run(calls=a.calls)
    res["microbatch"] = run_microbatch()
    print(json.dumps(res["microbatch"]))
    txt = json.dumps(res, indent=1)
    if a.out:
        os.makedirs(os.path.dirname(a.out), exist_ok=True)
        with open(a.out, "w") as fh:
            fh.write(txt)
    for r in res["rows"]:
        print(f"batch {r['batch']:5d}: predict+proba  b200 p50 {r['b200_predict_bracket']['p50_ms']:.3f} p99 {r['b200_predict_bracket']['p99_ms']:.3f} ms"
              f" | one-pass {r['b200_score_rows']['p50_ms']:.3f} | sklearn p50 {r['sklearn_predict_bracket']['p50_ms']:.3f} p99 {r['sklearn_predict_bracket']['p99_ms']:.3f}"
              f" || shap b200 p50 {r['b200_shap_bracket']['p50_ms']:.3f} p99 {r['b200_shap_bracket']['p99_ms']:.3f} | cpu p50 {r['cpu_shap_bracket']['p50_ms']:.3f}")
