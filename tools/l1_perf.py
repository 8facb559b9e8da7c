#!/usr/bin/env python
"""Time per 100,000 explained rows of every l1_reg mode of the Kernel SHAP solve (shared coalition set, config-3 shape):
none | num_features(10) | aic | bic | alpha.     python tools/l1_perf.py [rows] [tc] [modes...]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fraud_detection_b200 import KernelShapPlan, sampler  # noqa: E402
from oracle import scoring, synth  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
want = set(a for a in sys.argv[2:] if a != "tc")
evaluator = 0 if "tc" in sys.argv[2:] else -1          # "tc": the general tcgen05 evaluator (expectations through HBM)
prm = scoring.load_params()
bg = scoring.transform(prm, synth.rows(100, seed=43))
X = torch.from_numpy(scoring.transform(prm, synth.rows(rows, seed=7)).astype(np.float32)).cuda()
bits, kw = sampler.sample_coalitions(30, 2048, np.random.RandomState(0))
plan = KernelShapPlan(prm.coef, prm.intercept, bg, bits, kw)
for mode, par, name in ((0, 0, "none"), (1, 10, "num_features(10)"), (2, 0, "aic"), (3, 0, "bic"), (4, 1e-4, "alpha=1e-4")):
    if want and name.split("(")[0].split("=")[0] not in want:
        continue
    plan.set_l1(mode, par)
    best = 1e9
    for rep in range(4):
        torch.cuda.synchronize()
        t = time.perf_counter()
        phi = plan.explain_dev(X, evaluator=evaluator)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t)
    print(f"{name:18s} {best * 1e3:8.2f} ms per {rows} rows   {rows / best / 1e6:7.2f} M rows/s   "
          f"kept features per row {float((phi != 0).sum(1).float().mean()):.2f}", flush=True)
