#!/usr/bin/env python
"""Export golden vectors from the REAL shap package, the moment one is importable.

    python tools/make_golden_shap.py [/root/reference] [--out tests/golden]

shap==0.49.1 (the reference's requirements.txt:21) is not installed in the build image and cannot be installed
there (no network), so every SHAP parity test in this repository compares the CUDA path with CPU restatements
(oracle/linear_shap.py, oracle/kernel_shap.py): "parity unpinned".  This script is what pins it.  Run it on any
machine that has shap (ideally 0.49.1), scikit-learn and the reference's models/*.joblib; it writes

    tests/golden/shap_linear.npz   shap.LinearExplainer(model, X_s, feature_perturbation="interventional") on a
                                   background of MORE than 100 rows (explain_model.py:24-27): the subsampled
                                   background's mean, expected_value, phi of 64 rows; and the zero-background
                                   form of api/worker.py:52-53,75
    tests/golden/shap_kernel.npz   shap.KernelExplainer(f, background) for f = predict_proba[:, 1]:
                                   (a) shap_values(X, nsamples=2048, l1_reg=False) -> phi, plus maskMatrix,
                                       kernelWeights and ey of the LAST explained row (they survive on the
                                       explainer), fx, fnull;
                                   (b) the default call shap_values(X) (nsamples="auto", default l1_reg);
                                   (c) np.random state bookkeeping so the host sampler can replay the draws.
tests/test_golden_shap.py picks the files up automatically (and skips with the reason while they are absent).
Without shap the script exits 2 and writes nothing.
"""

from __future__ import annotations

import argparse
import os
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("ref", nargs="?", default="/root/reference")
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden"))
    args = ap.parse_args()
    try:
        import shap
    except Exception as exc:  # noqa: BLE001
        print(f"shap is not importable here ({exc!r}); nothing written. Reference KernelExplainer export: unavailable on this box.")
        return 2
    import joblib
    from oracle import synth
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        model = joblib.load(os.path.join(args.ref, "models", "logistic_model.joblib"))
        scaler = joblib.load(os.path.join(args.ref, "models", "scaler.joblib"))
        X = synth.reference_rows(2000, seed=42)
        Xs = scaler.transform(X)
    os.makedirs(args.out, exist_ok=True)

    # ---- LinearExplainer, as explain_model.py:24-27 builds and calls it (background > 100 rows -> subsampled)
    expl = shap.LinearExplainer(model, Xs, feature_perturbation="interventional")
    rows = Xs[:64]
    phi = np.asarray(expl.shap_values(rows))
    zero = shap.LinearExplainer(model, np.zeros((1, 30)))                     # api/worker.py:52-53
    np.savez_compressed(
        os.path.join(args.out, "shap_linear.npz"), shap_version=str(shap.__version__), background=Xs, rows=rows,
        mean=np.asarray(getattr(expl, "mean", np.nan)), expected_value=np.asarray(expl.expected_value), phi=phi,
        zero_bg_expected_value=np.asarray(zero.expected_value), zero_bg_phi=np.asarray(zero.shap_values(rows[:4])))

    # ---- KernelExplainer (no call site in the reference; BASELINE.json configs 3-4)
    def f(A):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            return model.predict_proba(np.asarray(A, dtype=np.float64))[:, 1]

    bg = Xs[100:200]
    krows = Xs[300:308]
    out = {"shap_version": str(shap.__version__), "background": bg, "rows": krows}
    for link in ("identity", "logit"):
        ke = shap.KernelExplainer(f, bg, link=link)
        np.random.seed(0)
        phi = np.asarray(ke.shap_values(krows, nsamples=2048, l1_reg=False, silent=True))
        out[f"{link}_phi"] = phi
        out[f"{link}_expected_value"] = np.asarray(ke.expected_value)
        # state of the LAST explained row
        out[f"{link}_maskMatrix"] = np.asarray(ke.maskMatrix)
        out[f"{link}_kernelWeights"] = np.asarray(ke.kernelWeights)
        out[f"{link}_ey"] = np.asarray(ke.ey).reshape(-1)
        out[f"{link}_fx"] = np.asarray(ke.fx).reshape(-1)
        out[f"{link}_fnull"] = np.asarray(ke.fnull).reshape(-1)
        out[f"{link}_nsamplesAdded"] = np.asarray(getattr(ke, "nsamplesAdded", -1))
    ke = shap.KernelExplainer(f, bg)
    np.random.seed(0)
    out["default_phi"] = np.asarray(ke.shap_values(krows, silent=True))          # nsamples="auto", default l1_reg
    out["default_nsamples"] = np.asarray(getattr(ke, "nsamples", -1))
    out["default_nonzero_per_row"] = (np.abs(out["default_phi"]) > 0).sum(1)
    np.savez_compressed(os.path.join(args.out, "shap_kernel.npz"), **out)
    print(f"wrote shap_linear.npz and shap_kernel.npz to {args.out} (shap {shap.__version__})")
    return 0


if __name__ == "__main__":
    sys.exit(main())
