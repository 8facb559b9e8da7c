#!/usr/bin/env python
"""bench.py - throughput of the fraud-detection scoring + SHAP hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Headline (BASELINE.json configs[1]): synthetic credit-card-shaped rows, float32, fused
StandardScaler + LogisticRegression.predict_proba + exact linear SHAP in ONE pass (K1).
A step is one pass over the resident batch (1e8 rows per GPU by default; weak scaling, rows
are independent so ranks shard them with no data-path collective - the only exchange is the
32-double all-reduce of the |SHAP| column sums, explain_model.py:39, issued on a side stream).

The JSON line carries
  value         rows/s with inputs resident in HBM (CUDA events, max over ranks)
  e2e           rows/s through the C-ABI host entry point with pinned HOST buffers
                (H2D + kernel + D2H inside the timed region)
  roofline      K1 against the measured HBM copy peak (244 algorithmic bytes per row)
  cpu_baseline  the reference's CPU path (live sklearn + numpy linear SHAP) on this box
  kernel_shap   configs 3/4: Kernel SHAP explained rows/s, rooflines against pipe rates MEASURED in this process
  config1       284,807 x 30 float64 rows through the drop-in transform / predict_proba / predict / shap_values
  config5       serving brackets (api/app.py:208-240, api/worker.py:74-75): p50/p99 at batch 1, 64, 4096
`--impl reference` times the CPU path alone with all host cores (scoring as the line's value; Kernel SHAP oracle
port, config 1 and config 5 as keyed extras).
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

_OUT = sys.stdout
METRIC = "scored_rows_per_sec"
UNIT = "rows/s"
BYTES_PER_ROW = 244           # 120 B x in + 4 B proba + 120 B phi
FALLBACK_HBM_GBS = 6650.0     # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent
# dram__bytes_read.sum + dram__bytes_write.sum of linear_pass_tma_kernel<float,0,1> from ONE `ncu --set full` capture
# (not re-measured by this run): 2e7 rows, 2.400175 GB read + 2.436786 GB written; scaled to the launch's row count
NCU_TRAFFIC_SOURCE = "profiles/r02/ncu_k1_tma.txt"
NCU_TRAFFIC_BYTES_PER_ROW = (2.400175e9 + 2.436786e9) / 2.0e7
CONFIG1_ROWS = 284_807
CONFIG5_BATCHES = (1, 64, 4096)
FEATURE_NAMES = ["Time"] + [f"V{i}" for i in range(1, 29)] + ["Amount"]


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
    ap.add_argument("--kshap-cpu-rows", type=int, default=int(os.environ.get("FDB_BENCH_KSHAP_CPU_ROWS", 1024)))
    ap.add_argument("--no-kshap", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip config 1 / config 5")
    ap.add_argument("--serving-calls", type=int, default=300)
    ap.add_argument("--cpu-rows", type=int, default=int(os.environ.get("FDB_BENCH_CPU_ROWS", 4_000_000)))
    return ap.parse_args()


def workload_config(args, world):
    return {
        "workload": "config2: synthetic credit-card-shaped rows (Time, V1..V28, Amount), fp32, fused "
                    "scale+logistic score+exact linear SHAP+|SHAP| column sums, one pass",
        "rows_per_gpu": args.rows, "global_rows": args.rows * world, "features": 30,
        "bytes_per_row": BYTES_PER_ROW, "sharding": f"rows x{world}, all-reduce of 32 doubles per step (side stream)",
        "l2": "inputs exceed L2 (24.4 GB touched per step vs 126 MB)",
        "e2e_rows_per_step": args.e2e_rows,
    }


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def synth_rows_host(n: int, seed: int) -> np.ndarray:
    """Credit-card-shaped float64 rows: the three distributions of the reference's generator
    (scripts/generate_synthetic_data.py:6-27): Time ~ U(0, 172800), V1..V28 ~ N(0, 1), Amount ~ exp(N(3, 1))."""
    rng = np.random.default_rng(seed)
    out = np.empty((n, 30), dtype=np.float64)
    out[:, 1:29] = rng.standard_normal((n, 28))
    out[:, 0] = rng.uniform(0.0, 172800.0, n)
    out[:, 29] = np.exp(rng.normal(3.0, 1.0, n))
    return out


# ------------------------------------------------------------------------------- CPU path (the checker's side)
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


def blas_info() -> dict:
    try:
        import threadpoolctl
        pools = threadpoolctl.threadpool_info()
        return {"default_threads": max([p.get("num_threads", 1) for p in pools] or [1]),
                "libraries": sorted({str(p.get("internal_api")) for p in pools})}
    except Exception as exc:  # noqa: BLE001
        return {"default_threads": None, "error": repr(exc)}


class CpuScoringArm:
    """The scoring + linear SHAP CPU arm in the two forms BASELINE.md section 3 names:
      library_default : ONE call chain over the whole sample, BLAS at its default thread count (what a user of the
                        reference gets from explain_model.py / evaluate_model.py);
      all_cores       : the sample cut into row chunks over a pool of `cores` Python threads, every BLAS / OpenMP pool
                        limited to ONE thread (threadpoolctl), so threads x BLAS threads = cores instead of cores^2.
    `best()` runs both and keeps the faster; that is the figure the GPU is compared with."""

    def __init__(self, rows: int):
        from concurrent.futures import ThreadPoolExecutor
        from oracle import scoring
        warnings.filterwarnings("ignore")
        self.scaler, self.model, self.origin = scoring.live_estimators()
        self.rows = rows
        self.cores = host_cores()
        self.X = synth_rows_host(rows, seed=42)                      # float64, like the reference
        self.mu_bg = self.scaler.transform(self.X[:100]).mean(0)
        chunks = np.array_split(np.arange(rows), max(self.cores * 4, 1))
        self.bounds = [(int(c[0]), int(c[-1]) + 1) for c in chunks if len(c)]
        self.pool = ThreadPoolExecutor(self.cores)
        self.blas = blas_info()

    def library_default(self):
        cpu_reference_pass(self.scaler, self.model, self.mu_bg, self.X)

    def all_cores(self):
        import threadpoolctl
        with threadpoolctl.threadpool_limits(limits=1):
            list(self.pool.map(lambda b: cpu_reference_pass(self.scaler, self.model, self.mu_bg, self.X[b[0]:b[1]]), self.bounds))

    def time_variant(self, fn, repeats: int) -> float:
        fn()
        best = float("inf")
        for _ in range(repeats):
            t0 = time.perf_counter()
            fn()
            best = min(best, time.perf_counter() - t0)
        return best

    def best(self, repeats: int = 3) -> dict:
        t_def = self.time_variant(self.library_default, max(1, repeats - 1))
        t_all = self.time_variant(self.all_cores, repeats)
        variants = {"library_default": {"value": self.rows / t_def, "seconds": t_def, "blas_threads": self.blas.get("default_threads"),
                                        "python_threads": 1},
                    "all_cores": {"value": self.rows / t_all, "seconds": t_all, "blas_threads": 1, "python_threads": self.cores}}
        pick = "all_cores" if t_all <= t_def else "library_default"
        return {"pick": pick, "variants": variants, "value": variants[pick]["value"]}

    def sample_text(self, pick: str) -> str:
        how = (f"row-chunked over {self.cores} Python threads with every BLAS/OpenMP pool limited to 1 thread" if pick == "all_cores"
               else f"one call chain, BLAS default ({self.blas.get('default_threads')} threads)")
        return (f"{self.rows} float64 rows per pass from the reference generator's distributions; sklearn {self.origin} estimators "
                f"(transform + predict_proba + predict) + numpy linear SHAP + |phi| column sums; {how}")


def _kshap_cpu_chunk(Xc, bg, nsamples, seed, shared):
    """Worker of the Kernel SHAP CPU baseline: the oracle port on a chunk of rows (one BLAS thread per worker)."""
    import threadpoolctl
    from oracle import kernel_shap as ks
    from oracle import scoring
    prm = scoring.load_params()
    f = lambda A: scoring.predict_proba(prm, A)[:, 1]  # noqa: E731
    with threadpoolctl.threadpool_limits(limits=1):
        if shared:        # ONE coalition set for every row (the GPU throughput mode): sampling paid once
            Z, kw, _ = ks.sample_coalitions(30, nsamples, np.random.RandomState(0))
            fnull = float(f(bg).mean())
            for r in Xc:
                ks.solve(Z, kw, ks.evaluate(f, r, bg, Z), float(f(r[None])[0]), fnull)
            return Xc.shape[0]
        return ks.shap_values(f, Xc, bg, nsamples=nsamples, l1_reg=False, random=np.random.RandomState(seed)).shape[0]


def cpu_kernel_shap(n_rows_per_core: int, nsamples: int = 2048, n_bg: int = 100) -> dict:
    """Oracle port of shap.KernelExplainer (per-row coalition sampling, l1_reg=False) over ALL host cores (worker
    processes), and the same with one shared coalition set; a bounded sample, extrapolated by the caller."""
    from joblib import Parallel, delayed
    from oracle import scoring
    prm = scoring.load_params()
    cores = host_cores()
    bg = scoring.transform(prm, synth_rows_host(n_bg, seed=43))
    n_all = max(cores * n_rows_per_core, cores)
    Xa = scoring.transform(prm, synth_rows_host(n_all, seed=1000))
    chunks = [c for c in np.array_split(np.arange(n_all), cores) if len(c)]
    out = {"cores": cores, "kind": "port", "unit": "explained rows/s"}
    with Parallel(n_jobs=cores) as pool:
        pool(delayed(_kshap_cpu_chunk)(Xa[:2], bg, nsamples, 0, False) for _ in range(cores))      # start the workers
        for key, shared in (("per_row_sets", False), ("shared_set", True)):
            t0 = time.perf_counter()
            pool(delayed(_kshap_cpu_chunk)(Xa[c], bg, nsamples, i, shared) for i, c in enumerate(chunks))
            dt = time.perf_counter() - t0
            out[key] = {"value": n_all / dt, "seconds": dt, "rows": n_all}
    out["value"] = out["per_row_sets"]["value"]
    out["sample"] = (f"{n_all} rows over {cores} worker processes through oracle/kernel_shap.py (numpy restatement of "
                     f"shap.KernelExplainer, nsamples={nsamples}, {n_bg}-row background, l1_reg=False; shap itself is not installable "
                     f"here): {out['per_row_sets']['seconds']:.1f}s with a fresh coalition sample per row (what shap does), "
                     f"{out['shared_set']['seconds']:.1f}s with one shared set")
    return out


def _percentiles(fn, calls, warm=20):
    for _ in range(warm):
        fn()
    ts = np.empty(calls)
    for i in range(calls):
        t0 = time.perf_counter()
        fn()
        ts[i] = time.perf_counter() - t0
    return {"p50_ms": float(np.percentile(ts, 50) * 1e3), "p99_ms": float(np.percentile(ts, 99) * 1e3)}


def serving_brackets(model, expl_fn, scaler_transform, calls: int) -> dict:
    """Config 5: the two brackets the reference instruments (api/app.py:208-240 predict + predict_proba on a DataFrame of
    scaled features; api/worker.py:74-75 shap_values on a DataFrame), a different request on every call."""
    import pandas as pd
    out = {}
    for bsz in CONFIG5_BATCHES:
        ring = max(2, min(32, (1 << 15) // bsz))
        raws = [synth_rows_host(bsz, seed=1000 * bsz + r) for r in range(ring)]
        Xss = [pd.DataFrame(scaler_transform(raw), columns=FEATURE_NAMES) for raw in raws]                    # api/app.py:194-197
        Xws = [pd.DataFrame(raw, columns=[f"feature_{i}" for i in range(30)]) for raw in raws]                # api/worker.py:70
        n = calls if bsz <= 64 else max(50, calls // 4)
        tick = [0]

        def nxt(seq):
            tick[0] += 1
            return seq[tick[0] % ring]

        def pair():
            Xs = nxt(Xss)
            return model.predict(Xs), model.predict_proba(Xs)

        out[f"batch_{bsz}"] = {"calls": n, "predict_bracket": _percentiles(pair, n),
                               "shap_bracket": _percentiles(lambda: expl_fn(nxt(Xws)), n)}
    return out


def cpu_config1(repeats: int = 3) -> dict:
    """Config 1 on the CPU: the reference's call chain (api/app.py:194-223, explain_model.py:19-40) on 284,807 float64 rows."""
    from oracle import linear_shap, scoring
    scaler, model, origin = scoring.live_estimators()
    X = synth_rows_host(CONFIG1_ROWS, seed=42)
    best = float("inf")
    for _ in range(repeats + 1):
        t0 = time.perf_counter()
        Xs = scaler.transform(X)
        proba = model.predict_proba(Xs)
        pred = model.predict(Xs)
        mean, _ev = linear_shap.explainer_state(model.coef_[0], float(model.intercept_[0]), Xs)
        phi = linear_shap.shap_values(model.coef_[0], mean, Xs)
        mean_abs, top3 = linear_shap.global_summary(phi)
        best = min(best, time.perf_counter() - t0)
    return {"value": CONFIG1_ROWS / best, "unit": UNIT, "seconds": best, "estimators": origin, "blas": blas_info(),
            "top3": [int(i) for i in top3], "_proba": proba, "_pred": pred, "_phi": phi, "_mean_abs": mean_abs}


def cpu_config5(calls: int) -> dict:
    from oracle import linear_shap, scoring
    scaler, model, _ = scoring.live_estimators()
    coef = model.coef_[0]
    zero = np.zeros(30)
    return serving_brackets(model, lambda Xw: linear_shap.shap_values(coef, zero, Xw), scaler.transform, calls)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    warnings.filterwarnings("ignore")
    arm = CpuScoringArm(args.cpu_rows)
    probe = arm.best(repeats=max(1, min(args.warmup, 3)))        # warm-up: both variants, keep the faster
    step = arm.all_cores if probe["pick"] == "all_cores" else arm.library_default
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = arm.rows * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.cores, "kind": "reference" if arm.origin == "joblib" else "port",
                         "sample": arm.sample_text(probe["pick"]), "variant": probe["pick"], "variants": probe["variants"],
                         "blas": arm.blas, "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS")},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    del arm
    if not args.no_kshap:
        try:
            line["kernel_shap"] = cpu_kernel_shap(max(2, args.kshap_cpu_rows // max(host_cores(), 1)))
        except Exception as exc:  # noqa: BLE001
            line["kernel_shap"] = {"unavailable": repr(exc)}
    if not args.no_extras:
        c1 = cpu_config1()
        line["config1"] = {k: v for k, v in c1.items() if not k.startswith("_")}
        line["config5"] = cpu_config5(args.serving_calls)
    _OUT.write(json.dumps(line) + "\n")
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
def load_drop_ins(device):
    """(MODEL, SCALER, origin): the documented integration - fraud_detection_b200.from_joblib on the reference's pickles
    (staged under oracle/_ref by oracle/build_ref.py, data only); without them the same classes are built from the
    exported parameters in tests/golden/model_params.json."""
    from fraud_detection_b200 import B200LogisticRegression, B200StandardScaler, from_joblib
    ref = os.path.join(ROOT, "oracle", "_ref")
    mp, sp = os.path.join(ref, "logistic_model.joblib"), os.path.join(ref, "scaler.joblib")
    if os.path.exists(mp) and os.path.exists(sp):
        model, scaler = from_joblib(mp, sp, device=device)
        return model, scaler, "from_joblib(oracle/_ref/*.joblib)"
    with open(os.path.join(ROOT, "tests", "golden", "model_params.json"), "r", encoding="utf-8") as fh:
        raw = json.load(fh)
    arr = lambda key: np.array([float.fromhex(h) for h in raw[key]], dtype=np.float64)  # noqa: E731
    model = B200LogisticRegression(arr("coef_hex"), float.fromhex(raw["intercept_hex"]), device=device)
    scaler = B200StandardScaler(arr("mean_hex"), arr("scale_hex"), arr("var_hex"), 284807, device=device)
    return model, scaler, "tests/golden/model_params.json"


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


def bench_config1(model, scaler, device_index) -> dict:
    """Config 1 through the drop-in surface, host arrays in and out (what an unmodified explain_model.py /
    api/app.py caller does): SCALER.transform -> MODEL.predict_proba -> MODEL.predict -> LinearExplainer(model, Xs)
    .shap_values(Xs) -> abs().mean(0) -> top-3.  `fused` is the same work in the drop-in's one-pass calls."""
    from fraud_detection_b200 import B200LinearExplainer
    X = synth_rows_host(CONFIG1_ROWS, seed=42)
    res = {}

    def chain():
        Xs = scaler.transform(X)
        proba = model.predict_proba(Xs)
        pred = model.predict(Xs)
        expl = B200LinearExplainer(model, Xs, feature_perturbation="interventional", device=device_index)
        phi = expl.shap_values(Xs)
        mean_abs = np.abs(phi).mean(axis=0)
        res.update(proba=proba, pred=pred, phi=phi, mean_abs=mean_abs, top3=mean_abs.argsort()[-3:][::-1])

    def fused():
        Xs = scaler.transform(X)
        pred, proba = model.score_rows(Xs)
        expl = B200LinearExplainer(model, Xs, feature_perturbation="interventional", device=device_index)
        phi, mean_abs, top3 = expl.explain_with_summary(Xs)
        res.update(f_top3=top3)

    def best_of(fn, n=4):
        fn()
        best = float("inf")
        for _ in range(n):
            t0 = time.perf_counter()
            fn()
            best = min(best, time.perf_counter() - t0)
        return best

    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        t_chain, t_fused = best_of(chain), best_of(fused)
    out = {"workload": f"config1: {CONFIG1_ROWS} x 30 float64 host rows (reference generator's distributions; the literal "
                       "data/credit_card.csv is a different dataset, SURVEY.md 0.4), transform -> predict_proba -> predict -> "
                       "LinearExplainer.shap_values -> mean|SHAP| -> top-3, host arrays in and out",
           "value": CONFIG1_ROWS / t_chain, "unit": UNIT, "seconds": t_chain, "fused_calls_value": CONFIG1_ROWS / t_fused,
           "top3": [int(i) for i in res["top3"]], "dtype": "f64"}
    return out, res


def bench_kernel_shap(args, torch, dist, device, rank, world, model, scaler, probes):
    """Config 3: nsamples=2048, 100-row background, 100k explained rows per GPU, shared coalition set (the throughput
    mode), the general tcgen05 evaluator, and the reference's own semantics (a fresh device-drawn coalition set per
    row).  Config 4 shape on top.  Timing is CUDA events on the launching stream, max over ranks."""
    from fraud_detection_b200 import KernelShapPlan, PinnedBuffer, _lib, host_near_gpu, sampler
    from fraud_detection_b200.distributed import allreduce_abs_sum

    coef, intercept = model.coef_[0], float(model.intercept_[0])
    rows, nsamples, n_bg = args.kshap_rows, 2048, 100
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        bg = scaler.transform(synth_rows_host(n_bg, seed=43))
        X = scaler.transform(synth_rows_host(rows, seed=1000 + rank)).astype(np.float32)
    bits, kw = sampler.sample_coalitions(30, nsamples, np.random.RandomState(0))
    plan = KernelShapPlan(coef, intercept, bg, bits, kw, device=device.index)
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

    def step_auto():
        plan.explain_dev(x, phi=phi, abs_sum=abs_sum)            # AUTO -> factored (K2f) + K3
        allreduce_abs_sum(abs_sum)                                # global mean|SHAP| summary (no-op at world 1)

    ms = timed(step_auto, steps)
    ms_tc = timed(lambda: plan.explain_dev(x, evaluator=0, phi=phi, abs_sum=abs_sum), steps)   # general tcgen05 evaluator (K2) + K3
    # the other l1_reg values of KernelExplainer.shap_values (shap's own default is "num_features(10)" [recalled]):
    # evaluate (K2f, expectations through HBM) + selection + reduced WLS, all on the device
    l1_modes = {}
    for name, mode, param in (("num_features(10)", 1, 10.0), ("aic", 2, 0.0), ("bic", 3, 0.0), ("alpha=1e-4", 4, 1e-4)):
        plan.set_l1(mode, param)
        ms_l1 = timed(step_auto, 3)
        l1_modes[name] = {"value": rows * world / (ms_l1 * 1e-3), "unit": "explained rows/s", "ms_per_step": ms_l1,
                          "features_kept_per_row": float((phi != 0).sum(1).float().mean().item())}
    plan.set_l1(0, 0.0)
    sig = rows * len(bits) * n_bg
    out = {
        "workload": f"config3: Kernel SHAP nsamples={nsamples}, {n_bg}-row background, {rows} explained rows per GPU, "
                    "l1_reg=False, identity link, f=P(fraud)",
        "value": rows * world / (ms * 1e-3), "unit": "explained rows/s", "ms_per_step": ms, "steps": steps,
        "mode": "ONE shared coalition set for all rows (throughput mode; compare with cpu_baseline.shared_set)",
        "sigmoids_per_s": sig * world / (ms * 1e-3),
        "evaluator": "K2f factored (single coalition set: one exponential per (row, coalition), E[m][b] precomputed) + K3",
        "gpu_launches_per_step": 2 * -(-rows // 262144),      # K2f + K3 per round of up to 262,144 rows
        "l1_reg_modes": l1_modes,
        "general_evaluator": {"evaluator": "K2 tcgen05 per-element sigmoid (any plan, the per-row-set path) + K3",
                              "value": rows * world / (ms_tc * 1e-3), "unit": "explained rows/s", "ms_per_step": ms_tc,
                              "sigmoids_per_s": sig * world / (ms_tc * 1e-3)},
    }
    # e2e through the host entry point (H2D of rows, D2H of phi inside the timed region)
    with host_near_gpu(device.index):
        hx, hphi = PinnedBuffer((rows, 30), np.float32), PinnedBuffer((rows, 30), np.float32)
    hx.array[:] = X
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
    sig4_rate = None
    if rows4 > 0:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            bg4 = scaler.transform(synth_rows_host(nb4, seed=44))
            X4 = scaler.transform(synth_rows_host(rows4, seed=2000 + rank)).astype(np.float32)
            b4, w4 = sampler.sample_coalitions(30, ns4, np.random.RandomState(4))
            plan4 = KernelShapPlan(coef, intercept, bg4, b4, w4, device=device.index)
        x4 = torch.from_numpy(X4).to(device)
        phi4 = torch.empty_like(x4)

        def step4():
            plan4.explain_dev(x4, phi=phi4, abs_sum=abs_sum)
            allreduce_abs_sum(abs_sum)

        ms4 = timed(step4, 3)
        ms4_tc = timed(lambda: plan4.explain_dev(x4, evaluator=0, phi=phi4), 2)
        sig4 = rows4 * len(b4) * nb4
        sig4_rate = sig4 / (ms4 * 1e-3)
        out["config4"] = {"workload": f"config4: Kernel SHAP nsamples={ns4}, {nb4}-row background, {rows4} explained rows per GPU, "
                                      "shared coalition set, all-reduce of the |SHAP| summary",
                          "value": rows4 * world / (ms4 * 1e-3), "unit": "explained rows/s", "ms_per_step": ms4,
                          "sigmoids_per_s": sig4 * world / (ms4 * 1e-3),
                          "general_evaluator_value": rows4 * world / (ms4_tc * 1e-3)}
        plan4.close()
        del x4, phi4

    # the reference's behaviour: a different coalition set for every explained row, drawn on the device.  Steady state of
    # a row stream: two plans alternate on two CUDA streams (KernelShapStream), the sets of batch i + 1 are drawn and
    # their X'WX inverted while batch i is evaluated and solved.
    from fraud_detection_b200 import KernelShapStream
    pr = args.kshap_per_row
    x2 = x[:pr].contiguous()
    phis = [torch.empty((pr, 30), dtype=torch.float32, device=device) for _ in range(2)]
    sync()
    t0 = time.perf_counter()
    stream = KernelShapStream(coef, intercept, bg, batch_rows=pr, nsamples=nsamples, seed=1 + 1000 * rank, device=device.index)
    torch.cuda.synchronize()
    t_plan = (time.perf_counter() - t0) / 2
    ms2 = timed(lambda: stream.plans[0].explain_dev(x2, phi=phis[0]), steps)       # evaluate + solve only, sets kept
    n_stream = 8
    for i in range(4):
        stream.explain(x2, phi=phis[i & 1])
    stream.synchronize()
    sync()
    t0 = time.perf_counter()
    for i in range(n_stream):
        stream.explain(x2, phi=phis[i & 1])
    stream.synchronize()
    t_stream = (time.perf_counter() - t0) / n_stream
    ts = torch.tensor([t_stream], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ts, op=dist.ReduceOp.MAX)
    t_stream = float(ts.item())
    out["per_row_sets"] = {"rows": pr, "sets": pr, "sampler": "device (Philox, reference schedule)",
                           "mode": "a FRESH coalition set per explained row (what shap does; compare with cpu_baseline.per_row_sets)",
                           "value": pr * world / t_stream, "unit": "explained rows/s",
                           "stream_ms_per_batch": t_stream * 1e3,
                           "evaluate_only_value": pr * world / (ms2 * 1e-3), "evaluate_only_ms": ms2,
                           "plan_ms": t_plan * 1e3,
                           "note": "value = per batch: draw the sets + per-set (X'WX)^-1 + evaluate + solve, two plans alternating on two "
                                   "streams (KernelShapStream), wall clock over 8 batches, fresh sets every batch; evaluate_only keeps the "
                                   "sets; plan_ms = creating one plan incl. allocation"}
    # ... and with shap's own default in front of the solve, l1_reg="num_features(10)" [recalled]: the same stream, the
    # plans switched to the LAR selection (their sets are redrawn every batch, the Z'WZ products with them)
    for pl in stream.plans:
        pl.set_l1(1, 10.0)
    for i in range(4):
        stream.explain(x2, phi=phis[i & 1])
    stream.synchronize()
    sync()
    t0 = time.perf_counter()
    for i in range(n_stream):
        stream.explain(x2, phi=phis[i & 1])
    stream.synchronize()
    ts = torch.tensor([(time.perf_counter() - t0) / n_stream], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ts, op=dist.ReduceOp.MAX)
    out["per_row_sets"]["num_features_10"] = {"value": pr * world / float(ts.item()), "unit": "explained rows/s",
                                              "stream_ms_per_batch": float(ts.item()) * 1e3,
                                              "note": "fresh set per row AND num_features(10): what an unmodified "
                                                      "KernelExplainer.shap_values(X) call does"}
    stream.close()

    # Rooflines against pipe rates measured in this process (fdb_probe_pipes), nominal figures beside them.
    sms = torch.cuda.get_device_properties(device).multi_processor_count
    nominal_hz = 1.965e9
    mufu_meas = min(probes["ex2_per_s"], probes["rcp_per_s"])
    issue_meas = probes["ffma_per_s"]                     # FP32 lane-results/s = warp-instruction issue rate x 32
    k2_rate, k2f_rate = sig / (ms_tc * 1e-3), sig / (ms * 1e-3)
    out["roofline"] = {
        "probes": probes,
        "general": {"kernel": "kshap_eval_tc_kernel (+K3 in the step)", "bound": "sfu", "unit": "T MUFU op/s",
                    "mufu_ops_per_sigmoid": 1.25, "achieved": 1.25 * k2_rate / 1e12, "peak": mufu_meas / 1e12,
                    "frac": 1.25 * k2_rate / mufu_meas, "peak_source": "measured: min(ex2.approx, rcp.approx) results/s, fdb_probe_pipes",
                    "peak_nominal": 16.0 * sms * nominal_hz / 1e12, "sigmoids_per_s": k2_rate},
        "factored": {"kernel": "kshap_eval_fx_kernel (+K3 in the step)", "bound": "fp32 issue", "unit": "T sigmoid/s",
                     "issue_slots_per_sigmoid": 3.5, "fma_pipe_instr_per_sigmoid": 3.0, "mufu_ops_per_sigmoid": 0.25,
                     "achieved": k2f_rate / 1e12, "peak": issue_meas / 3.5 / 1e12, "frac": k2f_rate / (issue_meas / 3.5),
                     "peak_source": "measured: independent-FFMA results/s (= issue rate x 32 lanes), / 3.5 issue slots per sigmoid "
                                    "(12 FP32 + 1 MUFU + 1 LDS.128 per 4 sigmoids)",
                     "peak_nominal": 128.0 * sms * nominal_hz / 3.5 / 1e12,
                     "inner_loop_alone": probes["k2f_quad_lds_sigmoids_per_s"] / 1e12,
                     "frac_of_inner_loop_alone": k2f_rate / probes["k2f_quad_lds_sigmoids_per_s"],
                     "inner_loop_alone_16_warps": probes["k2f_quad_16warps_sigmoids_per_s"] / 1e12,
                     "config4_achieved": None if sig4_rate is None else sig4_rate / 1e12,
                     "config4_frac": None if sig4_rate is None else sig4_rate / (issue_meas / 3.5)},
    }
    return out


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

    from fraud_detection_b200 import PinnedBuffer, _lib, host_near_gpu
    from fraud_detection_b200.device import FraudScorer, LinearPassOut
    from fraud_detection_b200.distributed import allreduce_abs_sum

    warnings.filterwarnings("ignore")
    model, scaler, params_origin = load_drop_ins(local)
    n = args.rows
    x = synth_rows_device(torch, n, 42 + rank, device)
    xs_head = (x[:100].double().cpu().numpy() - scaler.mean_) / scaler.scale_
    scorer = FraudScorer.from_estimators(model, scaler, xs_head.mean(0))
    proba, phi = torch.empty(n, dtype=torch.float32, device=device), torch.empty_like(x)
    # two |phi| summary buffers: the all-reduce of step k runs on a side stream while step k + 1's pass is in flight
    outs = [LinearPassOut(proba=proba, phi=phi, abs_sum=torch.empty(32, dtype=torch.float64, device=device)) for _ in range(2)]
    summaries = [torch.zeros(32, dtype=torch.float64, device=device) for _ in range(2)]
    side = torch.cuda.Stream(device)
    main = torch.cuda.current_stream(device)
    done = [None, None]

    def step(i, ev_a=None, ev_b=None):
        b = i & 1
        if done[b] is not None:
            main.wait_event(done[b])               # summary buffer b is free again (its all-reduce finished long ago)
        if ev_a is not None:
            ev_a.record()
        scorer(x, phi=True, abs_sum=True, out=outs[b])
        if ev_b is not None:
            ev_b.record()
        ready = torch.cuda.Event()
        ready.record(main)
        with torch.cuda.stream(side):
            side.wait_event(ready)
            summaries[b].copy_(outs[b].abs_sum)
            allreduce_abs_sum(summaries[b])        # no-op at world 1; NCCL all-reduce of 32 doubles otherwise
            d = torch.cuda.Event()
            d.record(side)
            done[b] = d

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(max(args.warmup, 3)):
        step(i)
    barrier()
    # ---- timed region: K steps, CUDA events on the launching stream, max over ranks
    k_events = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        e0.record()
        for i, (a, b) in enumerate(k_events):
            step(i, a, b)
        main.wait_stream(side)                     # the last all-reduce belongs to the timed region
        e1.record()
        barrier()
    total_ms = e0.elapsed_time(e1)
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in k_events]))
    t = torch.tensor([total_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = n * world * args.steps / (total_ms * 1e-3)

    # ---- the collective's result, checked (outside the timed region): all-gather every rank's own 32-vector and compare
    last = (args.steps - 1) & 1
    local_vec = outs[last].abs_sum.clone()
    if world > 1:
        gathered = torch.empty((world, 32), dtype=torch.float64, device=device)
        dist.all_gather_into_tensor(gathered, local_vec)
        expect = gathered.sum(0)
    else:
        expect = local_vec
    got = summaries[last]
    summary_ok = bool(torch.allclose(got[:30], expect[:30], rtol=1e-12, atol=0.0)) and float(got[30].item()) == float(n * world)
    from fraud_detection_b200.distributed import mean_abs_and_top3
    mean_abs, top3 = mean_abs_and_top3(got)

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
    del hx, hp, hphi, t_in, t_out, d_in, d_out, host_out

    probes = None
    if not args.no_kshap:
        barrier()
        probes = _lib.probe_pipes(local)           # every rank at once: the rates under the load the bench itself creates
    kshap = None if args.no_kshap else bench_kernel_shap(args, torch, dist, device, rank, world, model, scaler, probes)

    extras = {}
    if not args.no_extras and rank == 0:
        c1, c1_res = bench_config1(model, scaler, local)
        from fraud_detection_b200 import B200LinearExplainer
        import pandas as pd
        expl = B200LinearExplainer(model, pd.DataFrame([[0] * 30], columns=[f"feature_{i}" for i in range(30)]), device=local)   # api/worker.py:52-53
        c5 = serving_brackets(model, expl.shap_values, scaler.transform, args.serving_calls)
        extras = {"config1": c1, "config5": {"workload": "config5: serving brackets, host DataFrame in / numpy out, H2D + D2H inside; "
                                                         "predict_bracket = MODEL.predict + MODEL.predict_proba (api/app.py:208-240), "
                                                         "shap_bracket = EXPL.shap_values (api/worker.py:74-75); HTTP/Redis/Postgres excluded",
                                             **c5}}
        if world == 1:
            # the checker's side: the same calls on sklearn + the numpy LinearExplainer restatement, and the parity of config 1
            cpu1 = cpu_config1()
            c1["cpu"] = {k: v for k, v in cpu1.items() if not k.startswith("_")}
            c1["speedup_vs_cpu"] = c1["value"] / cpu1["value"]
            c1["parity"] = {"max_abs_dproba": float(np.abs(c1_res["proba"] - cpu1["_proba"]).max()),
                            "pred_equal": bool(np.array_equal(c1_res["pred"], cpu1["_pred"])),
                            "max_abs_dphi": float(np.abs(c1_res["phi"] - cpu1["_phi"]).max()),
                            "top3_equal": [int(i) for i in c1_res["top3"]] == cpu1["top3"]}
            cpu5 = cpu_config5(args.serving_calls)
            for key, row in cpu5.items():
                extras["config5"][key]["cpu"] = row
        del c1_res

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
            "parameters": params_origin,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": ne * 120,
                    "d2h_bytes_per_step": ne * 124 + 256, "rows_per_step": ne, "steps": e2e_steps,
                    "api": "fdb_linear_pass_host (pinned host in/out, 3-stream chunked pipeline)",
                    "pcie": pcie, "frac_of_link": (e2e_value / world) / pcie["rows_per_s_at_link_speed"],
                    "host_buffers_numa_local_cpus": None if near.cpus is None else len(near.cpus)},
            "gpu_launches": args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_TRAFFIC_BYTES_PER_ROW * n, "traffic_source": f"one ncu --set full capture at 2e7 rows ({NCU_TRAFFIC_SOURCE}), "
                                                                                      "scaled by rows; not re-measured by this run",
                         "kernel": "linear_pass_tma_kernel<float,false,true>", "kernel_ms": kernel_ms,
                         "algorithmic_bytes_per_launch": n * BYTES_PER_ROW, "peak_source": peak_src},
            "summary": {"summary_ok": summary_ok, "rows_counted": float(got[30].item()), "top3": [int(i) for i in top3],
                        "what": "all-reduced |SHAP| column sums == sum of the all-gathered per-rank vectors, [30] == global rows",
                        "collective": "ncclAllReduce of 32 doubles on a side stream, overlapped with the next pass" if world > 1 else "none (1 GPU)"},
            "clocks": clocks.summary(),
        }
        if kshap is not None:
            line["kernel_shap"] = kshap
        line.update(extras)
        if world == 1:
            arm = CpuScoringArm(args.cpu_rows)
            res = arm.best()
            line["cpu_baseline"] = {
                "value": res["value"], "unit": UNIT, "cores": arm.cores, "kind": "reference" if arm.origin == "joblib" else "port",
                "sample": arm.sample_text(res["pick"]) + ", best of 3", "variant": res["pick"], "variants": res["variants"], "blas": arm.blas,
            }
            del arm
            if kshap is not None:
                try:
                    kshap["cpu_baseline"] = cpu_kernel_shap(max(2, args.kshap_cpu_rows // max(host_cores(), 1)))
                except Exception as exc:  # noqa: BLE001
                    kshap["cpu_baseline"] = {"unavailable": repr(exc)}
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
