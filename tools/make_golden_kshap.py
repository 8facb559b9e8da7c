#!/usr/bin/env python
"""Generate tests/golden/kernel_shap_golden.npz (run in the build container).

The shap package is not installable here, so the Kernel SHAP ALGORITHM in these vectors is the oracle's
restatement (oracle/kernel_shap.py) - parity with shap itself stays unpinned.  What IS the reference's: the model
function.  Every coalition expectation ``ey`` in the file was computed by calling the reference's own unpickled
sklearn ``LogisticRegression.predict_proba`` (models/logistic_model.joblib) on the synthetic coalition rows, and the
explained / background rows come from the reference's data generator pushed through its ``scaler.joblib``.

    python tools/make_golden_kshap.py [/root/reference]

Pins: the coalition set (host sampler, seed 0, nsamples 2048), ey, f(x), E[f], and phi for the identity link, the
logit link and l1_reg="num_features(10)"; plus a row where only 7 features vary (single-row background).
"""

from __future__ import annotations

import os
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main(ref: str) -> None:
    import joblib
    from oracle import kernel_shap as ks
    from oracle import synth
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        model = joblib.load(os.path.join(ref, "models", "logistic_model.joblib"))
        scaler = joblib.load(os.path.join(ref, "models", "scaler.joblib"))

    def f(A):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            return model.predict_proba(np.asarray(A, dtype=np.float64))[:, 1]

    def scaled(n, seed):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            return scaler.transform(synth.rows(n, seed=seed))

    bg = scaled(100, 43)
    X = scaled(8, 2024)
    Z, kw, info = ks.sample_coalitions(30, 2048, np.random.RandomState(0))
    fnull = float(f(bg).mean())
    fx = f(X)
    ey = np.array([ks.evaluate(f, x, bg, Z) for x in X])
    phi_id = np.array([ks.solve(Z, kw, ey[i], float(fx[i]), fnull, link="identity") for i in range(len(X))])
    phi_logit = np.array([ks.solve(Z, kw, ey[i], float(fx[i]), fnull, link="logit") for i in range(len(X))])
    phi_nf10 = np.array([ks.solve(Z, kw, ey[i], float(fx[i]), fnull, l1_reg="num_features(10)") for i in range(len(X))])
    bits = (Z.astype(np.uint32) << np.arange(30, dtype=np.uint32)[None, :]).sum(1).astype(np.uint32)

    # a row where only 7 features differ from a one-row background (shap step 1, varyingInds)
    bg1 = scaled(1, 44)
    xv = scaled(1, 45)[0]
    vary = np.array([0, 3, 4, 9, 17, 22, 26])
    keep = np.ones(30, dtype=bool)
    keep[vary] = False
    xv[keep] = bg1[0, keep]
    phi_vary, parts = ks.explain_row(f, xv, bg1, nsamples="auto", random=np.random.RandomState(1), return_parts=True)
    assert parts["varying"].tolist() == vary.tolist()

    out = os.path.join(ROOT, "tests", "golden", "kernel_shap_golden.npz")
    np.savez_compressed(out, background=bg, X=X, bits=bits, kw=kw, ey=ey, fx=fx, fnull=np.float64(fnull),
                        phi_identity=phi_id, phi_logit=phi_logit, phi_num_features_10=phi_nf10,
                        weight_left=np.float64(info["weight_left"]), nfixed=np.int64(info["nfixed_samples"]),
                        vary_background=bg1, vary_x=xv, vary_features=vary, vary_phi=phi_vary,
                        vary_fx=np.float64(parts["fx"]), vary_fnull=np.float64(parts["fnull"]))
    print("wrote", out, os.path.getsize(out), "bytes | max |phi|", float(np.abs(phi_id).max()),
          "| additivity", float(np.abs(phi_id.sum(1) + fnull - fx).max()))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
