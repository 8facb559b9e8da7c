#!/usr/bin/env python
"""Per-row coalition sets at scale: python tools/kshap_perrow_perf.py [rows] [n_sets] (row r uses set r % n_sets)."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fraud_detection_b200 import KernelShapPlan, sampler  # noqa: E402
from oracle import scoring, synth  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
n_sets = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
prm = scoring.load_params()
bg = scoring.transform(prm, synth.rows(100, seed=43))
X = scoring.transform(prm, synth.rows(rows, seed=1)).astype(np.float32)
rs = np.random.RandomState(0)
t0 = time.perf_counter()
sets = [sampler.sample_coalitions(30, 2048, rs) for _ in range(n_sets)]
t_s = time.perf_counter() - t0
t0 = time.perf_counter()
plan = KernelShapPlan(prm.coef, prm.intercept, bg, np.stack([b for b, _ in sets]), np.stack([w for _, w in sets]))
t_p = time.perf_counter() - t0
x = torch.from_numpy(X).cuda()
phi = torch.empty_like(x)
plan.explain_dev(x, phi=phi)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    plan.explain_dev(x, phi=phi)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
print(f"per-row sets: rows={rows} sets={n_sets}: host sampler {t_s / n_sets * 1e3:.2f} ms/set, plan {t_p * 1e3:.0f} ms, "
      f"explain {ms:.3f} ms = {rows / ms * 1e3:.0f} rows/s")
