#!/usr/bin/env python
"""Stage the reference's own pickled estimators under oracle/_ref/ (test infrastructure).

The reference has no compiled code: its CPU implementation of the scoring path IS
scikit-learn (present in this image) applied to ``models/scaler.joblib`` and
``models/logistic_model.joblib``.  When ``/root/reference`` is mounted (the build
container) this copies those two binary artefacts - not source - into the git-ignored
``oracle/_ref/`` so they travel to the GPU box, where ``oracle.scoring.live_estimators``
unpickles them for the timed CPU baseline (``cpu_baseline.kind == "reference"``).
Without them the baseline falls back to sklearn estimators rebuilt from
``tests/golden/model_params.json`` (same arithmetic, checked bit-for-bit in
tests/test_oracle_scoring.py).
"""

from __future__ import annotations

import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = ("scaler.joblib", "logistic_model.joblib")


def build(ref: str = "/root/reference") -> bool:
    src = os.path.join(ref, "models")
    if not all(os.path.exists(os.path.join(src, f)) for f in FILES):
        return False
    dst = os.path.join(HERE, "_ref")
    os.makedirs(dst, exist_ok=True)
    for f in FILES:
        shutil.copyfile(os.path.join(src, f), os.path.join(dst, f))
    return True


if __name__ == "__main__":
    ok = build(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
    print("oracle/_ref staged" if ok else "reference not mounted; nothing staged")
