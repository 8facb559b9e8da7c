#!/usr/bin/env python
"""Print the headline numbers of a bench.py JSON line: python tools/print_bench.py gpurun_out/x_bench.json"""
import json
import sys

d = json.load(open(sys.argv[1]))
print(f"n_gpus {d['n_gpus']}  K1 {d['value'] / 1e9:.2f} G rows/s  roofline {d['roofline']['frac']:.3f}  "
      f"e2e {d['e2e']['value'] / 1e6:.1f} M rows/s ({d['e2e'].get('frac_of_link', float('nan')):.2f} of link)  "
      f"summary_ok {d.get('summary', {}).get('summary_ok')}  clocks {d.get('clocks')}")
k = d.get("kernel_shap")
if k:
    pr = k["per_row_sets"]
    print(f"Kernel SHAP cfg3 {k['value'] / 1e6:.2f} M rows/s (general {k['general_evaluator']['value'] / 1e6:.2f})  "
          f"cfg4 {k['config4']['value'] / 1e6:.3f} M rows/s ({k['config4']['ms_per_step']:.1f} ms)  "
          f"per-row stream {pr['value'] / 1e6:.2f} (evaluate only {pr['evaluate_only_value'] / 1e6:.2f}, "
          f"with num_features(10) {pr.get('num_features_10', {}).get('value', float('nan')) / 1e6:.2f}) M rows/s  e2e {k['e2e']['value'] / 1e6:.2f}")
    if "l1_reg_modes" in k:
        print("l1_reg modes (shared set)", {n: f"{v['value'] / 1e6:.2f} M rows/s" for n, v in k["l1_reg_modes"].items()})
    r = k["roofline"]
    print("probes", {a: f"{b:.4g}" for a, b in r["probes"].items()})
    print(f"K2  {r['general']['achieved']:.2f} of {r['general']['peak']:.2f} {r['general']['unit']} = {r['general']['frac']:.3f}")
    f = r["factored"]
    print(f"K2f {f['achieved']:.2f} of {f['peak']:.2f} {f['unit']} = {f['frac']:.3f}; inner loop alone {f['inner_loop_alone']:.2f} "
          f"({f['frac_of_inner_loop_alone']:.3f}); cfg4 {f['config4_achieved']} = {f['config4_frac']}")
    if "cpu_baseline" in k:
        print("kshap cpu", json.dumps({a: b for a, b in k["cpu_baseline"].items() if a != "sample"}))
if "config1" in d:
    c = d["config1"]
    print("config1", {a: c[a] for a in c if a != "workload"})
if "config5" in d:
    for key, row in d["config5"].items():
        if key.startswith("batch_"):
            print("config5", key, json.dumps(row))
if "cpu_baseline" in d:
    print("cpu_baseline", json.dumps({a: b for a, b in d["cpu_baseline"].items() if a != "sample"}))
