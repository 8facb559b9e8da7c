#!/usr/bin/env python
"""bench.py - throughput of the fraud-detection scoring + SHAP hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Headline (BASELINE.json configs[1]): synthetic credit-card-shaped rows, float32, fused
StandardScaler + LogisticRegression.predict_proba + exact linear SHAP in ONE pass (K1).
A step is one pass over the resident batch (1e8 rows per GPU by default; weak scaling, rows
are independent so ranks shard them with no data-path collective - the only exchange is the
31-double all-reduce of the |SHAP| column sums, explain_model.py:39).

The JSON line carries
  value     rows/s with inputs resident in HBM (CUDA events, max over ranks)
  e2e       rows/s through the C-ABI host entry point with pinned HOST buffers
            (H2D + kernel + D2H inside the timed region)
  roofline  the kernel against the measured HBM copy peak (244 algorithmic bytes per row)
  cpu_baseline  the reference's CPU path (live sklearn + numpy linear SHAP) on this box
`--impl reference` times that CPU path alone with all host threads.
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

_OUT = sys.stdout
METRIC = "scored_rows_per_sec"
UNIT = "rows/s"
BYTES_PER_ROW = 244           # 120 B x in + 4 B proba + 120 B phi
FALLBACK_HBM_GBS = 6650.0     # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent
# dram__bytes_read.sum + dram__bytes_write.sum of linear_pass_tma_kernel<float,0,1>, one `ncu --set full` capture at
# 2e7 rows (profiles/r01/ncu_k1_tma.txt: 2.400650 GB + 2.436800 GB), per row; scaled to the launch's row count
NCU_TRAFFIC_BYTES_PER_ROW = (2.400650e9 + 2.436800e9) / 2.0e7


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=int(os.environ.get("FDB_BENCH_ROWS", 100_000_000)),
                    help="rows per GPU resident in HBM (config 2: 1e8)")
    ap.add_argument("--e2e-rows", type=int, default=int(os.environ.get("FDB_BENCH_E2E_ROWS", 32_000_000)))
    ap.add_argument("--kshap-rows", type=int, default=int(os.environ.get("FDB_BENCH_KSHAP_ROWS", 100_000)))
    ap.add_argument("--kshap-per-row", type=int, default=32768)
    ap.add_argument("--kshap4-rows", type=int, default=125_000)
    ap.add_argument("--kshap-cpu-rows", type=int, default=256)
    ap.add_argument("--no-kshap", action="store_true")
    ap.add_argument("--cpu-rows", type=int, default=int(os.environ.get("FDB_BENCH_CPU_ROWS", 4_000_000)))
    return ap.parse_args()


def workload_config(args, world):
    return {
        "workload": "config2: synthetic credit-card-shaped rows (Time, V1..V28, Amount), fp32, fused "
                    "scale+logistic score+exact linear SHAP+|SHAP| column sums, one pass",
        "rows_per_gpu": args.rows, "global_rows": args.rows * world, "features": 30,
        "bytes_per_row": BYTES_PER_ROW, "sharding": f"rows x{world}, all-reduce of 31 doubles per step",
        "l2": "inputs exceed L2 (24.4 GB touched per step vs 126 MB)",
        "e2e_rows_per_step": args.e2e_rows,
    }


# ------------------------------------------------------------------------------- CPU path
def cpu_reference_pass(scaler, model, mu_bg, X):
    """The reference's own CPU work per batch: SCALER.transform -> predict_proba -> predict
    (api/app.py:194,209,223) + LinearExplainer.shap_values and mean|SHAP|
    (explain_model.py:27,39).  sklearn executes the scoring; shap is not installed, so the
    two explainer lines are the numpy restatement from oracle/linear_shap.py."""
    from oracle import linear_shap
    Xs = scaler.transform(X)
    proba = model.predict_proba(Xs)
    pred = model.predict(Xs)
    phi = linear_shap.shap_values(model.coef_[0], mu_bg, Xs)
    return proba, pred, phi, np.abs(phi).sum(axis=0)


def time_cpu(rows: int, threads: int, repeats: int = 3):
    import warnings
    from concurrent.futures import ThreadPoolExecutor
    from oracle import scoring, synth
    warnings.filterwarnings("ignore")
    scaler, model, origin = scoring.live_estimators()
    X = synth.rows(rows, seed=42)                      # float64, like the reference
    mu_bg = scaler.transform(X[:100]).mean(0)
    chunks = np.array_split(np.arange(rows), max(threads * 4, 1))
    bounds = [(int(c[0]), int(c[-1]) + 1) for c in chunks if len(c)]

    def run_all():
        if threads <= 1:
            cpu_reference_pass(scaler, model, mu_bg, X)
            return
        with ThreadPoolExecutor(threads) as ex:        # numpy releases the GIL in these kernels
            list(ex.map(lambda b: cpu_reference_pass(scaler, model, mu_bg, X[b[0]:b[1]]), bounds))

    run_all()
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        run_all()
        best = min(best, time.perf_counter() - t0)
    return rows / best, best, origin


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    import warnings
    from concurrent.futures import ThreadPoolExecutor
    from oracle import scoring, synth
    warnings.filterwarnings("ignore")
    scaler, model, origin = scoring.live_estimators()
    rows = args.cpu_rows
    X = synth.rows(rows, seed=42)
    mu_bg = scaler.transform(X[:100]).mean(0)
    chunks = np.array_split(np.arange(rows), threads * 4)
    bounds = [(int(c[0]), int(c[-1]) + 1) for c in chunks if len(c)]
    pool = ThreadPoolExecutor(threads)

    def step():
        list(pool.map(lambda b: cpu_reference_pass(scaler, model, mu_bg, X[b[0]:b[1]]), bounds))

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = rows * args.steps / dt
    sample = (f"{rows} float64 rows per step from the same generator; sklearn {origin} estimators "
              f"(transform+predict_proba+predict) + numpy linear SHAP; row-chunked over {threads} threads")
    _OUT.write(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads,
                         "kind": "reference" if origin == "joblib" else "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }) + "\n")
    _OUT.flush()


# ------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nv = None

    def _loop(self):
        nv = self._nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                try:
                    mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                except Exception:
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def __enter__(self):
        if self._nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=1.0)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------- B200 arm
def synth_rows_device(torch, n, seed, device, chunk=10_000_000):
    """Config 2 input: the reference generator's three distributions, drawn on the device."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    x = torch.empty((n, 30), dtype=torch.float32, device=device)
    for r0 in range(0, n, chunk):
        r1 = min(n, r0 + chunk)
        blk = x[r0:r1]
        blk.normal_(generator=g)
        blk[:, 0].uniform_(0.0, 172800.0, generator=g)
        blk[:, 29] = torch.exp(torch.empty(r1 - r0, device=device).normal_(3.0, 1.0, generator=g))
    return x



# ------------------------------------------------------------------------------- Kernel SHAP (configs 3/4)
def bench_kernel_shap(args, torch, dist, device, rank, world, prm, sm_max_mhz):
    """Config 3: nsamples=2048, 100-row background, 100k explained rows per GPU, shared coalition set
    (the throughput mode), plus a per-row-mask subset (the reference's sampler).  Returns a dict for
    the JSON line; timing is CUDA events on the launching stream, max over ranks."""
    from fraud_detection_b200 import KernelShapPlan, PinnedBuffer, host_near_gpu, sampler
    from oracle import scoring, synth

    rows, nsamples, n_bg = args.kshap_rows, 2048, 100
    bg = scoring.transform(prm, synth.rows(n_bg, seed=43))
    X = scoring.transform(prm, synth.rows(rows, seed=1000 + rank)).astype(np.float32)
    bits, kw = sampler.sample_coalitions(30, nsamples, np.random.RandomState(0))
    plan = KernelShapPlan(prm.coef, prm.intercept, bg, bits, kw, device=device.index)
    x = torch.from_numpy(X).to(device)
    phi = torch.empty_like(x)
    abs_sum = torch.empty(32, dtype=torch.float64, device=device)
    steps = max(3, min(args.steps, 10))

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        for _ in range(3):
            fn()
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        sync()
        t = torch.tensor([e0.elapsed_time(e1) / n], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    from fraud_detection_b200.distributed import allreduce_abs_sum

    def step_auto():
        plan.explain_dev(x, phi=phi, abs_sum=abs_sum)            # AUTO -> factored (K2f) + K3
        allreduce_abs_sum(abs_sum)                                # global mean|SHAP| summary (no-op at world 1)

    ms = timed(step_auto, steps)
    ms_tc = timed(lambda: plan.explain_dev(x, evaluator=0, phi=phi, abs_sum=abs_sum), steps)   # general tcgen05 evaluator (K2) + K3
    sig = rows * len(bits) * n_bg
    out = {
        "workload": f"config3: Kernel SHAP nsamples={nsamples}, {n_bg}-row background, {rows} explained rows per GPU, "
                    "shared coalition set, l1_reg=False, identity link, f=P(fraud)",
        "value": rows * world / (ms * 1e-3), "unit": "explained rows/s", "ms_per_step": ms, "steps": steps,
        "sigmoids_per_s": sig * world / (ms * 1e-3),
        "evaluator": "K2f factored (single coalition set: one exponential per (row, coalition), E[m][b] precomputed) + K3",
        "gpu_launches_per_step": 2 * -(-rows // 32768),
        "general_evaluator": {"evaluator": "K2 tcgen05 per-element sigmoid (any plan, the per-row-mask path) + K3",
                              "value": rows * world / (ms_tc * 1e-3), "unit": "explained rows/s", "ms_per_step": ms_tc,
                              "sigmoids_per_s": sig * world / (ms_tc * 1e-3)},
    }
    # e2e through the host entry point (H2D of rows, D2H of phi inside the timed region)
    with host_near_gpu(device.index):
        hx, hphi = PinnedBuffer((rows, 30), np.float32), PinnedBuffer((rows, 30), np.float32)
    hx.array[:] = X
    import ctypes as C  # noqa: F401
    from fraud_detection_b200 import _lib
    habs = np.zeros(32)

    def host_call():
        _lib.check(_lib.lib.fdb_kshap_explain_host(plan._h, rows, _lib.np_ptr(hx.array), None, -1, _lib.np_ptr(hphi.array),
                                                   _lib.np_ptr(habs)))
    host_call()
    sync()
    t0 = time.perf_counter()
    for _ in range(3):
        host_call()
    dt = (time.perf_counter() - t0) / 3
    t = torch.tensor([dt], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out["e2e"] = {"value": rows * world / float(t.item()), "unit": "explained rows/s", "h2d_bytes_per_step": rows * 120,
                  "d2h_bytes_per_step": rows * 120 + 256, "api": "fdb_kshap_explain_host (pinned host rows in, phi out)"}
    plan.close()

    # config 4: nsamples=4096, 1000-row background, 125,000 explained rows per GPU (1M rows over 8 GPUs), one all-reduce
    rows4, ns4, nb4 = args.kshap4_rows, 4096, 1000
    if rows4 > 0:
        bg4 = scoring.transform(prm, synth.rows(nb4, seed=44))
        b4, w4 = sampler.sample_coalitions(30, ns4, np.random.RandomState(4))
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            plan4 = KernelShapPlan(prm.coef, prm.intercept, bg4, b4, w4, device=device.index)
        x4 = torch.from_numpy(scoring.transform(prm, synth.rows(rows4, seed=2000 + rank)).astype(np.float32)).to(device)
        phi4 = torch.empty_like(x4)

        def step4():
            plan4.explain_dev(x4, phi=phi4, abs_sum=abs_sum)
            allreduce_abs_sum(abs_sum)

        ms4 = timed(step4, 3)
        ms4_tc = timed(lambda: plan4.explain_dev(x4, evaluator=0, phi=phi4), 2)
        sig4 = rows4 * len(b4) * nb4
        out["config4"] = {"workload": f"config4: Kernel SHAP nsamples={ns4}, {nb4}-row background, {rows4} explained rows per GPU, "
                                      "shared coalition set, all-reduce of the |SHAP| summary",
                          "value": rows4 * world / (ms4 * 1e-3), "unit": "explained rows/s", "ms_per_step": ms4,
                          "sigmoids_per_s": sig4 * world / (ms4 * 1e-3),
                          "general_evaluator_value": rows4 * world / (ms4_tc * 1e-3)}
        plan4.close()
        del x4, phi4

    # the reference's behaviour: a different coalition set for every explained row, drawn on the device
    pr = args.kshap_per_row
    x2, phi2 = x[:pr].contiguous(), torch.empty((pr, 30), dtype=torch.float32, device=device)
    KernelShapPlan.sampled(prm.coef, prm.intercept, bg, 64, nsamples, seed=0, device=device.index).close()   # warm the module
    sync()
    t0 = time.perf_counter()
    plan2 = KernelShapPlan.sampled(prm.coef, prm.intercept, bg, pr, nsamples, seed=1 + rank, device=device.index)
    torch.cuda.synchronize()
    t_plan = time.perf_counter() - t0
    ms2 = timed(lambda: plan2.explain_dev(x2, phi=phi2), steps)
    # steady state of a row stream: the plan is kept, its sets are redrawn in place for every new batch of rows
    sync()
    t0 = time.perf_counter()
    for i in range(3):
        plan2.resample(100 + 3 * rank + i)
        plan2.explain_dev(x2, phi=phi2)
    torch.cuda.synchronize()
    t_stream = (time.perf_counter() - t0) / 3
    out["per_row_masks"] = {"rows": pr, "sets": pr, "sampler": "device (Philox, reference schedule)",
                            "value": pr * world / (ms2 * 1e-3), "unit": "explained rows/s", "ms_per_step": ms2,
                            "plan_ms": t_plan * 1e3,
                            "value_including_sampling": pr * world / (ms2 * 1e-3 + t_plan),
                            "stream_ms_per_batch": t_stream * 1e3, "value_stream": pr * world / t_stream,
                            "note": "plan_ms = drawing the sets + operand images + per-set (X'WX)^-1, paid once per batch of rows; "
                                    "value_stream = fdb_kshap_plan_resample + explain per batch (wall clock, fresh sets every batch)"}
    plan2.close()

    sms = torch.cuda.get_device_properties(device).multi_processor_count
    clock_hz = (sm_max_mhz or 1965) * 1e6
    peak = 16.0 * sms * clock_hz
    # SFU roofline (north-star form: one exp per sigmoid against SFU exp throughput).  K2 issues 1.25 MUFU ops per
    # sigmoid (ex2 + a reciprocal shared by four); K2f issues 0.26 (reciprocal shared by four + one ex2 per 100), so it
    # is bound by issue slots (3.5 per sigmoid in the clamp-free loop), not by the SFU, and may exceed the one-exp-per-sigmoid line.
    out["roofline"] = {
        "bound": "sfu", "unit": "T sigmoid/s", "peak": peak / 1e12,
        "achieved": sig / (ms_tc * 1e-3) / 1e12, "frac": sig / (ms_tc * 1e-3) / peak, "kernel": "kshap_eval_tc_kernel (+K3 in the step)",
        "mufu_ops_per_sigmoid": 1.25, "mufu_frac": 1.25 * sig / (ms_tc * 1e-3) / peak,
        "factored": {"achieved": sig / (ms * 1e-3) / 1e12, "frac": sig / (ms * 1e-3) / peak, "kernel": "kshap_eval_fx_kernel (+K3 in the step)",
                     "mufu_ops_per_sigmoid": 0.26, "issue_slots_per_sigmoid": 3.5},
        "peak_source": f"nominal 16 MUFU/clk/SM x {sms} SMs x {clock_hz / 1e9:.3f} GHz = one exp per sigmoid (no measured SFU peak on file)",
    }
    if rank == 0 and world == 1:
        # CPU: the oracle restatement of shap.KernelExplainer on a handful of rows, extrapolated
        from oracle import kernel_shap as ks
        f = lambda A: scoring.predict_proba(prm, A)[:, 1]  # noqa: E731
        n_cpu = args.kshap_cpu_rows
        t0 = time.perf_counter()
        ks.shap_values(f, X[:n_cpu].astype(np.float64), bg, nsamples=nsamples, l1_reg=False, random=np.random.RandomState(0))
        dtc = time.perf_counter() - t0
        # all host cores: the same oracle over row chunks in worker processes (not single-thread-vs-GPU)
        all_cores = None
        try:
            from joblib import Parallel, delayed
            n_jobs = len(os.sched_getaffinity(0))
            n_all = n_cpu * min(n_jobs, 8)
            Xa = np.ascontiguousarray(np.resize(X[:n_cpu].astype(np.float64), (n_all, 30)))
            chunks = np.array_split(np.arange(n_all), n_jobs)
            with Parallel(n_jobs=n_jobs) as pool:
                pool(delayed(_kshap_cpu_chunk)(Xa[:2], bg, nsamples, 0) for _ in range(n_jobs))     # start the workers
                t0 = time.perf_counter()
                pool(delayed(_kshap_cpu_chunk)(Xa[c], bg, nsamples, i) for i, c in enumerate(chunks))
                dta = time.perf_counter() - t0
            all_cores = {"value": n_all / dta, "cores": n_jobs, "sample": f"{n_all} rows over {n_jobs} worker processes in {dta:.1f}s"}
        except Exception as exc:  # noqa: BLE001
            all_cores = {"unavailable": repr(exc)}
        out["cpu_baseline"] = {"value": n_cpu / dtc, "unit": "explained rows/s", "cores": 1, "kind": "port", "all_cores": all_cores,
                               "sample": f"{n_cpu} rows through oracle/kernel_shap.py (numpy restatement of shap.KernelExplainer, "
                                         f"per-row sampling, one process; shap itself is not installable here) in {dtc:.1f}s; extrapolates "
                                         f"to {rows / (n_cpu / dtc) / 3600:.1f} h for {rows} rows"}
    return out


def _kshap_cpu_chunk(Xc, bg, nsamples, seed):
    """Worker of the all-cores Kernel SHAP CPU baseline: the oracle on a chunk of rows."""
    from oracle import kernel_shap as ks
    from oracle import scoring
    prm = scoring.load_params()
    f = lambda A: scoring.predict_proba(prm, A)[:, 1]  # noqa: E731
    return ks.shap_values(f, Xc, bg, nsamples=nsamples, l1_reg=False, random=np.random.RandomState(seed)).shape[0]


def run_b200(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=device)

    from fraud_detection_b200 import PinnedBuffer, host_near_gpu
    from fraud_detection_b200.device import FraudScorer, LinearPassOut
    from fraud_detection_b200.distributed import allreduce_abs_sum
    from oracle import scoring

    prm = scoring.load_params()          # parameters only (tests/golden/model_params.json)
    n = args.rows
    x = synth_rows_device(torch, n, 42 + rank, device)
    xs_head = (x[:100].double().cpu().numpy() - prm.mean) / prm.scale
    scorer = FraudScorer(prm.coef, prm.intercept, prm.mean, prm.scale, xs_head.mean(0))
    out = LinearPassOut(proba=torch.empty(n, dtype=torch.float32, device=device), phi=torch.empty_like(x),
                        abs_sum=torch.empty(32, dtype=torch.float64, device=device))
    summary = torch.empty(32, dtype=torch.float64, device=device)

    def step():
        scorer(x, phi=True, abs_sum=True, out=out)
        summary.copy_(out.abs_sum)
        allreduce_abs_sum(summary)       # no-op at world 1; NCCL all-reduce of 31 doubles otherwise

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    # ---- timed region: K steps, CUDA events on the launching stream, max over ranks
    k_events = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        e0.record()
        for a, b in k_events:
            a.record()
            scorer(x, phi=True, abs_sum=True, out=out)
            b.record()
            summary.copy_(out.abs_sum)
            allreduce_abs_sum(summary)
        e1.record()
        barrier()
    total_ms = e0.elapsed_time(e1)
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in k_events]))
    t = torch.tensor([total_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = n * world * args.steps / (total_ms * 1e-3)

    # ---- e2e: pinned host buffers through fdb_linear_pass_host (H2D + kernel + D2H timed)
    ne = args.e2e_rows
    with host_near_gpu(local) as near:     # page-locked buffers on the socket this GPU's PCIe link hangs off
        hx, hp, hphi = PinnedBuffer((ne, 30), np.float32), PinnedBuffer((ne,), np.float32), PinnedBuffer((ne, 30), np.float32)
    hx.array[:] = x[:ne].cpu().numpy() if ne <= n else 0
    host_out = {"proba": hp.array, "phi": hphi.array, "abs_sum": np.zeros(32)}
    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        scorer.score_host(hx.array, phi=True, abs_sum=True, out=host_out, device=local)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        scorer.score_host(hx.array, phi=True, abs_sum=True, out=host_out, device=local)
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = ne * world * e2e_steps / float(t.item())

    # what the link itself carries: plain pinned copies of the same sizes, both directions at once (the e2e roofline)
    s_in, s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
    t_in, t_out = torch.from_numpy(hx.array), torch.from_numpy(hphi.array)
    d_in, d_out = torch.empty_like(x[:ne]), torch.empty_like(x[:ne])
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    for rep in range(2):
        barrier()
        with torch.cuda.stream(s_in):
            ev[0].record()
            d_in.copy_(t_in, non_blocking=True)
            ev[1].record()
        with torch.cuda.stream(s_out):
            ev[2].record()
            t_out.copy_(d_out, non_blocking=True)
            ev[3].record()
        barrier()
    pcie = {"h2d_GBps": ne * 120 / (ev[0].elapsed_time(ev[1]) * 1e-3) / 1e9, "d2h_GBps": ne * 120 / (ev[2].elapsed_time(ev[3]) * 1e-3) / 1e9,
            "what": "plain pinned copies of one step's input and phi, both directions concurrently, this rank"}
    pcie["rows_per_s_at_link_speed"] = min(pcie["h2d_GBps"] * 1e9 / 120, pcie["d2h_GBps"] * 1e9 / 124)
    del hx, hp, hphi, t_in, t_out, d_in, d_out
    kshap = None if args.no_kshap else bench_kernel_shap(args, torch, dist, device, rank, world, prm, clocks.max_mhz)

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            with open(peaks_path) as fh:
                peak, peak_src = float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"
        achieved = n * BYTES_PER_ROW / (kernel_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, world),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": ne * 120,
                    "d2h_bytes_per_step": ne * 124 + 256, "rows_per_step": ne, "steps": e2e_steps,
                    "api": "fdb_linear_pass_host (pinned host in/out, 3-stream chunked pipeline)",
                    "pcie": pcie, "frac_of_link": (e2e_value / world) / pcie["rows_per_s_at_link_speed"],
                    "host_buffers_numa_local_cpus": None if near.cpus is None else len(near.cpus)},
            "gpu_launches": args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_TRAFFIC_BYTES_PER_ROW * n, "kernel": "linear_pass_tma_kernel<float,false,true>", "kernel_ms": kernel_ms,
                         "algorithmic_bytes_per_launch": n * BYTES_PER_ROW, "peak_source": peak_src},
            "clocks": clocks.summary(),
        }
        if kshap is not None:
            line["kernel_shap"] = kshap
        if world == 1:
            cpu_threads = os.cpu_count() or 1
            v, secs, origin = time_cpu(args.cpu_rows, cpu_threads)
            v1, secs1, _ = time_cpu(max(args.cpu_rows // 8, 1), 1, repeats=2)
            line["cpu_baseline"] = {
                "value": v, "unit": UNIT, "cores": cpu_threads, "kind": "reference" if origin == "joblib" else "port",
                "sample": f"{args.cpu_rows} float64 rows, best of 3 ({secs:.2f}s): sklearn {origin} estimators "
                          f"transform+predict_proba+predict + numpy linear SHAP, row-chunked over {cpu_threads} threads",
                "single_thread_value": v1,
            }
        _OUT.write(json.dumps(line) + "\n")
        _OUT.flush()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _claim_stdout():
    """Keep stdout for the ONE JSON line: everything else a library prints there (NCCL's version banner, warnings)
    is sent to stderr at the file-descriptor level; returns a writer for the real stdout."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    return os.fdopen(real, "w")


def main():
    args = parse_args()
    global _OUT
    _OUT = _claim_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
