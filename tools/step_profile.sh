#!/bin/bash
# Per-kernel device time of one training step (eager, no graph) — run on the GPU box.
# Plain run first (must exit 0), then the same command under ncu (B200_PROFILING.md).
set -e
export HSTD_BENCH_NO_COOLDOWN=1
CMD="python bench.py --steps 1 --warmup 3 --no_graph ${BENCH_ARGS}"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 340 -c 520 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu.log 2>&1
python - <<'PY'
import csv, collections, re
rows = []
with open("gpurun_out/launches.csv") as f:
    lines = [l for l in f if l.startswith('"')]
r = csv.DictReader(lines)
for row in r:
    try:
        rows.append((row["Kernel Name"], float(row["Metric Value"].replace(",", "")), row.get("Metric Unit", "")))
    except Exception:
        pass
agg = collections.defaultdict(lambda: [0, 0.0])
for name, v, unit in rows:
    scale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3}.get(unit, 1e-3)
    short = re.sub(r"\(.*", "", name)[:90]
    agg[short][0] += 1
    agg[short][1] += v * scale
tot = sum(v[1] for v in agg.values())
with open("gpurun_out/step_profile.txt", "w") as out:
    out.write(f"total {tot:.1f} us over {len(rows)} launches\n")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.write(f"{t:10.1f} us {100*t/tot:5.1f}% n={n:4d} avg={t/n:8.1f}  {k}\n")
print(open("gpurun_out/step_profile.txt").read()[:6000])
PY
