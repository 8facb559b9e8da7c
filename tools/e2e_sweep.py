#!/usr/bin/env python
"""e2e K1 through fdb_linear_pass_host against the number of pipeline chunks: python tools/e2e_sweep.py [rows]."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fraud_detection_b200 import PinnedBuffer, host_near_gpu  # noqa: E402
from fraud_detection_b200.device import FraudScorer  # noqa: E402
from oracle import scoring, synth  # noqa: E402

ne = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
prm = scoring.load_params()
scorer = FraudScorer(prm.coef, prm.intercept, prm.mean, prm.scale, np.zeros(30))
with host_near_gpu(0) as near:
    hx, hp, hphi = PinnedBuffer((ne, 30), np.float32), PinnedBuffer((ne,), np.float32), PinnedBuffer((ne, 30), np.float32)
print("cpus near gpu:", None if near.cpus is None else len(near.cpus))
base = synth.rows(1 << 20, seed=1).astype(np.float32)
for i in range(0, ne, 1 << 20):
    hx.array[i:i + (1 << 20)] = base[: min(1 << 20, ne - i)]
out = {"proba": hp.array, "phi": hphi.array, "abs_sum": np.zeros(32)}
for chunks in (12, 16, 24, 32, 48, 64, 96, 128):
    os.environ["FDB_PIPE_CHUNKS"] = str(chunks)
    for _ in range(2):
        scorer.score_host(hx.array, phi=True, abs_sum=True, out=out, device=0)
    t0 = time.perf_counter()
    for _ in range(5):
        scorer.score_host(hx.array, phi=True, abs_sum=True, out=out, device=0)
    dt = (time.perf_counter() - t0) / 5
    print(f"chunks={chunks:4d}  {dt * 1e3:7.2f} ms  {ne / dt / 1e6:7.1f} M rows/s  h2d {ne * 120 / dt / 1e9:5.1f} GB/s  d2h {ne * 124 / dt / 1e9:5.1f} GB/s")
