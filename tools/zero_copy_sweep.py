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
    X = synth.rows(batch, seed=3)
    for _ in range(200):
        model.predict_proba(X)
    res = {}
    for name, fn in (("transform", lambda: scaler.transform(X)), ("predict_proba", lambda: model.predict_proba(X)),
                     ("score_rows", lambda: model.score_rows(X))):
        ts = []
        for _ in range(2000):
            t0 = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t0)
        res[name] = np.percentile(ts, 50) * 1e6
    print(f"batch {batch:5d}: " + "  ".join(f"{k} p50 {v:7.1f} us" for k, v in res.items()))
