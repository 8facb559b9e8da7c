#!/usr/bin/env python
"""Instruction-pipe rates of the GPU (the Kernel SHAP roofline denominators): python tools/microbench/pipes.py
The kernels live in fraud_detection_b200/csrc/pipe_probe.cu and run through the C ABI (fdb_probe_pipes), the same call
bench.py makes; membench.cu next to this file is the HBM counterpart used for K1."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from fraud_detection_b200 import _lib  # noqa: E402

p = _lib.probe_pipes(0)
sms, ghz = p["sm_count"], 1.965
print(json.dumps(p, indent=1))
print(f"per clock per SM at {ghz} GHz: ex2 {p['ex2_per_s'] / sms / ghz / 1e9:.2f}, rcp {p['rcp_per_s'] / sms / ghz / 1e9:.2f}, "
      f"ffma {p['ffma_per_s'] / sms / ghz / 1e9:.1f}; K2f issue bound {p['ffma_per_s'] / 3.5 / 1e12:.2f} T sigmoid/s, "
      f"quad loop alone {p['k2f_quad_lds_sigmoids_per_s'] / 1e12:.2f} T (64 warps/SM), "
      f"{p['k2f_quad_16warps_sigmoids_per_s'] / 1e12:.2f} T (16 warps/SM)")
