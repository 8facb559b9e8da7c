import sys, time, numpy as np
sys.path.insert(0, ".")
from fraud_detection_b200 import KernelShapPlan
from oracle import scoring, synth
import torch
prm = scoring.load_params(); bg = scoring.transform(prm, synth.rows(100, seed=43))
KernelShapPlan.sampled(prm.coef, prm.intercept, bg, 64, 2048, seed=0).close()
torch.cuda.synchronize()
t0=time.perf_counter(); plan = KernelShapPlan.sampled(prm.coef, prm.intercept, bg, 32768, 2048, seed=1); torch.cuda.synchronize(); print("plan ms", (time.perf_counter()-t0)*1e3)
plan.close()
