#!/usr/bin/env python
"""Where config 1's end-to-end time goes, call by call (host numpy arrays in and out): python tools/config1_profile.py"""
import os
import sys
import time
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from fraud_detection_b200 import B200LinearExplainer  # noqa: E402

warnings.filterwarnings("ignore")
model, scaler, origin = bench.load_drop_ins(0)
X = bench.synth_rows_host(bench.CONFIG1_ROWS, seed=42)


def t(fn, n=5):
    fn()
    best = 1e9
    for _ in range(n):
        t0 = time.perf_counter()
        out = fn()
        best = min(best, time.perf_counter() - t0)
    return best * 1e3, out


ms, Xs = t(lambda: scaler.transform(X)); print(f"transform            {ms:7.2f} ms")
ms, p = t(lambda: model.predict_proba(Xs)); print(f"predict_proba        {ms:7.2f} ms")
ms, c = t(lambda: model.predict(Xs)); print(f"predict              {ms:7.2f} ms")
ms, ex = t(lambda: B200LinearExplainer(model, Xs)); print(f"LinearExplainer()    {ms:7.2f} ms")
ms, phi = t(lambda: ex.shap_values(Xs)); print(f"shap_values          {ms:7.2f} ms")
ms, _ = t(lambda: np.abs(phi).mean(axis=0)); print(f"np.abs(phi).mean(0)  {ms:7.2f} ms")
ms, _ = t(lambda: ex.explain_with_summary(Xs)); print(f"explain_with_summary {ms:7.2f} ms")
ms, _ = t(lambda: model.score_rows(Xs)); print(f"score_rows           {ms:7.2f} ms")
ms, _ = t(lambda: np.empty_like(X).fill(0)); print(f"touch 68 MB          {ms:7.2f} ms")
ms, _ = t(lambda: X.copy()); print(f"copy 68 MB           {ms:7.2f} ms")
from fraud_detection_b200 import PinnedBuffer  # noqa: E402
pin = PinnedBuffer(X.shape, np.float64); pin.array[:] = X
pout = PinnedBuffer(X.shape, np.float64)
from fraud_detection_b200 import _lib  # noqa: E402
def pinned_transform():
    _lib.check(_lib.lib.fdb_transform_host(_lib.get_ctx(0), _lib.F64, X.shape[0], _lib.np_ptr(pin.array), scaler._mean_c, scaler._scale_c, _lib.np_ptr(pout.array)))
ms, _ = t(pinned_transform); print(f"transform, pinned in/out {ms:7.2f} ms")
