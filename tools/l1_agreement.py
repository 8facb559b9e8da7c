#!/usr/bin/env python
"""How often does the device's feature selection pick exactly the features sklearn picks?  For every l1_reg mode: N rows,
2048 coalitions (shared set), the oracle (real sklearn estimators on shap's augmented system) fed the GPU's own coalition
expectations and float32 rows.     python tools/l1_agreement.py [rows]   -> one JSON document"""
import json
import os
import sys
import warnings

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fraud_detection_b200 import KernelShapPlan, sampler  # noqa: E402
from oracle import kernel_shap as ks  # noqa: E402
from oracle import scoring, synth  # noqa: E402

warnings.filterwarnings("ignore")
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
prm = scoring.load_params()
bg = scoring.transform(prm, synth.rows(100, seed=43))
X = scoring.transform(prm, synth.rows(rows, seed=11)).astype(np.float32)
f = lambda A: scoring.predict_proba(prm, A)[:, 1]  # noqa: E731
fnull = float(f(bg).mean())
Z, kw, _ = ks.sample_coalitions(30, 2048, np.random.RandomState(0))
plan = KernelShapPlan(prm.coef, prm.intercept, bg, sampler.masks_to_bits(Z), kw)
x = torch.from_numpy(X).cuda()
ey = torch.empty((rows, plan.n_masks), dtype=torch.float32, device="cuda")
out = {"rows": rows, "coalitions": int(len(kw)), "background": 100, "modes": {}}
for name, mode, par, l1 in (("num_features(10)", 1, 10, "num_features(10)"), ("aic", 2, 0, "aic"), ("bic", 3, 0, "bic"), ("alpha=1e-4", 4, 1e-4, 1e-4)):
    plan.set_l1(mode, par)
    phi = plan.explain_dev(x, evaluator=0, ey=ey).cpu().numpy().astype(np.float64)
    e = ey.cpu().numpy().astype(np.float64)
    same, worst, kept = 0, 0.0, 0
    for i in range(rows):
        xi = X[i].astype(np.float64)
        ref = ks.solve(Z, kw, e[i], float(f(xi[None])[0]), fnull, l1_reg=l1)
        kept += np.count_nonzero(ref)
        if np.array_equal(ref != 0, phi[i] != 0):
            same += 1
            worst = max(worst, float(np.abs(phi[i] - ref).max() / max(np.abs(ref).max(), 1e-300)))
    out["modes"][name] = {"identical_support_rows": same, "of": rows, "worst_rel_dphi_on_those": worst, "features_kept_per_row": kept / rows}
    print(name, out["modes"][name], file=sys.stderr, flush=True)
print(json.dumps(out, indent=1))
