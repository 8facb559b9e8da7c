set -e
python tools/plan_prof.py
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/plan_launches.csv python tools/plan_prof.py > gpurun_out/plan_ncu.log 2>&1 || true
python - <<'PY'
import csv, collections
rows = list(csv.reader(l for l in open("gpurun_out/plan_launches.csv") if l.startswith('"')))
h = rows[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value"); ui = h.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    k = r[ki][:60]; v = float(r[vi].replace(",", "")); u = r[ui]
    v = v / 1e3 if u in ("ns", "nsecond") else (v * 1e3 if u in ("ms", "msecond") else v)
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
for k, (n, t) in agg.items(): print(f"{k:60s} n={n:4d} total_us={t:10.1f}")
PY
