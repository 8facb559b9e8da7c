#!/usr/bin/env python
"""Per-row coalition sets as a stream: python tools/kshap_stream_perf.py [rows per batch] [nsamples] [n_bg] [batches] [l1_reg]
Prints the device time of the redraw (sampler + per-set Cholesky), of evaluate + solve, and the stream rate of
KernelShapStream (two plans alternating on two CUDA streams)."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fraud_detection_b200 import KernelShapStream  # noqa: E402
from oracle import scoring, synth  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
nsamples = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
n_bg = int(sys.argv[3]) if len(sys.argv) > 3 else 100
batches = int(sys.argv[4]) if len(sys.argv) > 4 else 8
l1_reg = sys.argv[5] if len(sys.argv) > 5 else None          # e.g. "num_features(10)" (shap's default), "aic"
prm = scoring.load_params()
bg = scoring.transform(prm, synth.rows(n_bg, seed=43))
X = scoring.transform(prm, synth.rows(rows, seed=1)).astype(np.float32)
x = torch.from_numpy(X).cuda()
phis = [torch.empty_like(x) for _ in range(2)]
ks = KernelShapStream(prm.coef, prm.intercept, bg, batch_rows=rows, nsamples=nsamples, seed=3, l1_reg=l1_reg)
plan = ks.plans[0]


def timed(fn, n=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


ms_draw = timed(lambda: plan.resample(7, stream="current"))
ms_eval = timed(lambda: plan.explain_dev(x, phi=phis[0]))
for i in range(4):
    ks.explain(x, phi=phis[i & 1])
ks.synchronize()
t0 = time.perf_counter()
for i in range(batches):
    ks.explain(x, phi=phis[i & 1])
ks.synchronize()
dt = (time.perf_counter() - t0) / batches
print(f"per-row sets rows={rows} M={nsamples} B={n_bg} l1_reg={l1_reg}: redraw {ms_draw:.3f} ms, evaluate+solve {ms_eval:.3f} ms, "
      f"stream {dt * 1e3:.3f} ms per batch = {rows / dt / 1e6:.2f} M rows/s")
