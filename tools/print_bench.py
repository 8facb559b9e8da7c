#!/usr/bin/env python
"""Print the headline numbers of a bench.py JSON line: python tools/print_bench.py gpurun_out/x_bench.json"""
import json
import sys

d = json.load(open(sys.argv[1]))
print(f"n_gpus {d['n_gpus']}  K1 {d['value'] / 1e9:.2f} G rows/s  roofline {d['roofline']['frac']:.3f}  "
      f"e2e {d['e2e']['value'] / 1e6:.1f} M rows/s ({d['e2e'].get('frac_of_link', float('nan')):.2f} of link)")
k = d.get("kernel_shap")
if k:
    pr = k["per_row_masks"]
    print(f"Kernel SHAP cfg3 {k['value'] / 1e6:.2f} M rows/s (general {k['general_evaluator']['value'] / 1e6:.2f})  "
          f"cfg4 {k['config4']['value'] / 1e6:.3f} M rows/s ({k['config4']['ms_per_step']:.1f} ms)  "
          f"per-row {pr['value'] / 1e6:.2f} / stream {pr.get('value_stream', 0) / 1e6:.2f} M rows/s  e2e {k['e2e']['value'] / 1e6:.2f}")
