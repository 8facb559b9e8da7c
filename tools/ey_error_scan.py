import sys; sys.path.insert(0, ".")
import numpy as np, torch
from fraud_detection_b200 import KernelShapPlan, sampler
from oracle import scoring, synth, kernel_shap as ks
prm = scoring.load_params()
f = lambda A: scoring.predict_proba(prm, A)[:, 1]
for n_bg in (256, 512, 896, 1000, 1024):
    bg = scoring.transform(prm, synth.rows(n_bg, seed=43)); X = scoring.transform(prm, synth.rows(6, seed=100))
    Z, kw, _ = ks.sample_coalitions(30, 512, np.random.RandomState(0))
    ey_ref = np.array([ks.evaluate(f, x, bg, Z) for x in X])
    plan = KernelShapPlan(prm.coef, prm.intercept, bg, sampler.masks_to_bits(Z), kw)
    x = torch.from_numpy(X.astype(np.float32)).cuda()
    for ev, name in ((2, "FX"), (0, "TC"), (1, "SIMT")):
        ey = torch.empty((6, 512), device="cuda")
        plan.explain_dev(x, evaluator=ev, ey=ey); torch.cuda.synchronize()
        d = np.abs(ey.cpu().numpy() - ey_ref)
        print(n_bg, name, "max abs ey err %.2e" % d.max(), "at ey=%.3f" % ey_ref.flat[d.argmax()])
    plan.close()
