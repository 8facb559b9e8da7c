#!/usr/bin/env python
"""Time Kernel SHAP (config 3/4 shapes) on the device: python tools/kshap_perf.py [rows] [nsamples] [n_bg] [reps]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fraud_detection_b200 import KernelShapPlan, sampler  # noqa: E402
from oracle import scoring, synth  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
nsamples = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
n_bg = int(sys.argv[3]) if len(sys.argv) > 3 else 100
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
num_features = int(sys.argv[5]) if len(sys.argv) > 5 else 0
link = sys.argv[6] if len(sys.argv) > 6 else "identity"
prm = scoring.load_params()
bg = scoring.transform(prm, synth.rows(n_bg, seed=43))
X = scoring.transform(prm, synth.rows(rows, seed=1)).astype(np.float32)
bits, kw = sampler.sample_coalitions(30, nsamples, np.random.RandomState(0))
plan = KernelShapPlan(prm.coef, prm.intercept, bg, bits, kw, link)
if num_features:
    plan.set_num_features(num_features)
x = torch.from_numpy(X).cuda()
phi = torch.empty_like(x)
abs_sum = torch.empty(32, dtype=torch.float64, device="cuda")
for ev, name in ((2, "factored"), (0, "tcgen05"), (1, "simt")):
    if ev == 1 and rows > 4096:
        xx, pp = x[:4096], phi[:4096]
    else:
        xx, pp = x, phi
    plan.explain_dev(xx, evaluator=ev, phi=pp, abs_sum=abs_sum)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        plan.explain_dev(xx, evaluator=ev, phi=pp, abs_sum=abs_sum)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    n = xx.shape[0]
    sig = n * len(bits) * n_bg
    print(f"{name:8s} l1={num_features} rows={n} M={len(bits)} B={n_bg}: {ms:9.3f} ms  {n / ms * 1e3:12.0f} rows/s  {sig / ms / 1e6:8.2f} G sigmoid/s")
