#!/usr/bin/env python
"""Latency of the serving-size host path against the zero-copy threshold: FDB_ZERO_COPY_ROWS=N python tools/zero_copy_sweep.py"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fraud_detection_b200 import B200LogisticRegression, B200StandardScaler  # noqa: E402
from oracle import scoring, synth  # noqa: E402

prm = scoring.load_params()
model = B200LogisticRegression(prm.coef, prm.intercept)
scaler = B200StandardScaler(prm.mean, prm.scale)
print("FDB_ZERO_COPY_ROWS =", os.environ.get("FDB_ZERO_COPY_ROWS", "(default)"))
for batch in [int(b) for b in os.environ.get("FDB_SWEEP_BATCHES", "1,8,32,64,128,256,1024,4096").split(",")]:
    ring = [synth.rows(batch, seed=3 + r) for r in range(4)]      # distinct requests: the paired-call cache never hits
    for i in range(200):
        model.predict_proba(ring[i % 4])
    res = {}
    for name, fn in (("transform", scaler.transform), ("predict_proba", model.predict_proba), ("score_rows", model.score_rows)):
        ts = []
        for i in range(2000):
            X = ring[i % 4]
            t0 = time.perf_counter()
            fn(X)
            ts.append(time.perf_counter() - t0)
        res[name] = np.percentile(ts, 50) * 1e6
    print(f"batch {batch:5d}: " + "  ".join(f"{k} p50 {v:7.1f} us" for k, v in res.items()))
