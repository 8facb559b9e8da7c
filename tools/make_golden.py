#!/usr/bin/env python
"""Generate tests/golden/* from the REAL reference artefacts (run in the build container).

Needs ``/root/reference`` (read-only there; absent on the GPU box), so the outputs are
committed.  What it pins:

* ``model_params.json``   - coef_/intercept_/mean_/scale_/var_ of the reference's
  ``models/logistic_model.joblib`` + ``models/scaler.joblib`` as float.hex (bit exact).
* ``scoring_golden.npz``  - 512 rows of the reference's own synthetic generator
  (``scripts/generate_synthetic_data.py``, imported, seed 42) pushed through the
  unpickled sklearn objects: transform, decision_function, predict_proba, predict; plus
  float64 checksums of the same outputs over the full 284,807-row draw (config 1).

    python tools/make_golden.py [/root/reference]
"""

from __future__ import annotations

import importlib.util
import json
import os
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main(ref: str) -> None:
    import joblib
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        model = joblib.load(os.path.join(ref, "models", "logistic_model.joblib"))
        scaler = joblib.load(os.path.join(ref, "models", "scaler.joblib"))
    with open(os.path.join(ref, "models", "feature_names.json"), "r", encoding="utf-8") as fh:
        names = json.load(fh)
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)

    params = {
        "source": "models/logistic_model.joblib + models/scaler.joblib (sklearn pickles, written by 1.6.1)",
        "feature_names": names,
        "coef_hex": [float(v).hex() for v in model.coef_[0]],
        "intercept_hex": float(model.intercept_[0]).hex(),
        "mean_hex": [float(v).hex() for v in scaler.mean_],
        "scale_hex": [float(v).hex() for v in scaler.scale_],
        "var_hex": [float(v).hex() for v in scaler.var_],
        "classes": [int(c) for c in model.classes_],
        "n_samples_seen": int(scaler.n_samples_seen_),
    }
    with open(os.path.join(out_dir, "model_params.json"), "w", encoding="utf-8") as fh:
        json.dump(params, fh, indent=1)

    spec = importlib.util.spec_from_file_location(
        "ref_synth", os.path.join(ref, "scripts", "generate_synthetic_data.py"))
    ref_synth = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref_synth)
    cols = names

    def run(n):
        df = ref_synth.generate_synthetic_data(n_samples=n)
        X = df[cols].to_numpy(dtype=np.float64)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            Xs = scaler.transform(X)
            return X, Xs, model.decision_function(Xs), model.predict_proba(Xs), model.predict(Xs)

    from oracle import synth
    X, Xs, z, proba, pred = run(512)
    assert np.array_equal(X, synth.reference_rows(512)), "oracle.synth drifted from the reference generator"
    Xf, Xsf, zf, probaf, predf = run(284807)
    assert np.array_equal(Xf, synth.reference_rows(284807))
    np.savez_compressed(
        os.path.join(out_dir, "scoring_golden.npz"),
        X=X, X_scaled=Xs, decision=z, proba=proba, pred=pred,
        full_n=np.int64(284807),
        full_scaled_colsum=Xsf.sum(0), full_decision_sum=zf.sum(), full_proba1_sum=probaf[:, 1].sum(),
        full_pred_sum=np.int64(predf.sum()),
        full_decision_head=zf[:8], full_decision_tail=zf[-8:],
        full_decision_minmax=np.array([zf.min(), zf.max()]),
    )
    print("wrote", out_dir, "| pred positives in full draw:", int(predf.sum()),
          "| logit range", float(zf.min()), float(zf.max()))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
