#!/usr/bin/env python
"""Input-path throughput: CSV -> pinned rows -> fused raw-row pass, next to pandas.read_csv + sklearn.
    python tools/ingest_bench.py [rows]"""
import os
import sys
import tempfile
import time
import warnings

import numpy as np
import pandas as pd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fraud_detection_b200 import io as fio  # noqa: E402
from fraud_detection_b200.device import FraudScorer  # noqa: E402
from oracle import linear_shap, scoring, synth  # noqa: E402

warnings.filterwarnings("ignore")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 284807
prm = scoring.load_params()
X = synth.reference_rows(n)
with tempfile.TemporaryDirectory() as d:
    path = os.path.join(d, "creditcard.csv")
    pd.DataFrame(np.column_stack([X, np.zeros(n)]), columns=synth.FEATURE_NAMES + ["Class"]).to_csv(path, index=False)
    size = os.path.getsize(path)
    scorer = FraudScorer(prm.coef, prm.intercept, prm.mean, prm.scale)
    fio.score_file(path, scorer, dtype=np.float64)
    t0 = time.perf_counter(); rows = fio.read_rows(path, dtype=np.float64); t_read = time.perf_counter() - t0
    t0 = time.perf_counter(); out = fio.score_file(path, scorer, dtype=np.float64); t_all = time.perf_counter() - t0
    scaler, model, _ = scoring.live_estimators()
    t0 = time.perf_counter()
    df = pd.read_csv(path)
    t_pd = time.perf_counter() - t0
    Xs = scaler.transform(df[synth.FEATURE_NAMES]); p = model.predict_proba(Xs); phi = linear_shap.shap_values(model.coef_[0], np.zeros(30), Xs)
    ma = np.abs(phi).mean(0)
    t_ref = time.perf_counter() - t0
    print(f"{n} rows, {size / 1e6:.1f} MB CSV | b200: read {t_read * 1e3:.0f} ms, read+score+SHAP {t_all * 1e3:.0f} ms "
          f"({n / t_all / 1e6:.2f} M rows/s, {size / t_all / 1e6:.0f} MB/s) | reference: pandas.read_csv {t_pd * 1e3:.0f} ms, "
          f"read+sklearn+numpy SHAP {t_ref * 1e3:.0f} ms ({n / t_ref / 1e6:.2f} M rows/s)")
