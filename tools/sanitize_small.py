#!/usr/bin/env python
"""Tiny run of every kernel for compute-sanitizer (memcheck / racecheck), one tool per gpurun call:
    compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fraud_detection_b200 import KernelShapPlan, sampler, wrap  # noqa: E402
from fraud_detection_b200.device import FraudScorer  # noqa: E402
from oracle import scoring, synth  # noqa: E402

prm = scoring.load_params()
scaler_sk, model_sk, _ = scoring.live_estimators(prm)
scaler, model = wrap(scaler_sk), wrap(model_sk)
X = synth.rows(45001, seed=1)
for dt in (np.float32, np.float64):
    Xs = scaler.transform(X.astype(dt))                      # TMA kernel (>= 37,888 rows) + tail rows
    model.score_rows(Xs)
    model.predict_proba(Xs[:1001])                           # grid-stride kernel
x = torch.from_numpy(X.astype(np.float32)).cuda()
FraudScorer(prm.coef, prm.intercept, prm.mean, prm.scale)(x, phi=True, cls=True, abs_sum=True)
bg = scoring.transform(prm, synth.rows(130, seed=43))
bits, kw = sampler.sample_coalitions(30, 300, np.random.RandomState(0))
plan = KernelShapPlan(prm.coef, prm.intercept, bg, bits, kw)
xr = torch.from_numpy(scoring.transform(prm, synth.rows(300, seed=2)).astype(np.float32)).cuda()
for ev in (0, 1, 2):
    plan.explain_dev(xr, evaluator=ev)
plan.set_num_features(10)
plan.explain_dev(xr, evaluator=2)
plan.close()
plan = KernelShapPlan.sampled(prm.coef, prm.intercept, bg, 8, 256, seed=1)
plan.explain_dev(xr[:8])
plan.close()
torch.cuda.synchronize()
print("sanitize_small: done")
# stream-ordered redraw and the multi-chunk factored path (n_bg > 128)
from fraud_detection_b200 import KernelShapStream  # noqa: E402
st = KernelShapStream(prm.coef, prm.intercept, bg, batch_rows=16, nsamples=256, seed=2)
for i in range(3):
    st.explain(xr[:16])
st.synchronize()
st.close()
bits, kw = sampler.sample_coalitions(30, 200, np.random.RandomState(1))
plan = KernelShapPlan(prm.coef, prm.intercept, scoring.transform(prm, synth.rows(40, seed=44)), bits, kw, link="logit")
plan.explain_dev(xr[:150], evaluator=2)
plan.close()
torch.cuda.synchronize()
print("sanitize_small: stream + logit done")
