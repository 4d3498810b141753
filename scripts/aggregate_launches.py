"""Aggregate an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel name.
usage: python scripts/aggregate_launches.py gpurun_out/rXX_launches_step.csv "<command line>" > profiles/rXX_launches_step_aggregated.txt"""
import csv, sys
from collections import defaultdict

path, cmd = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
rows = []
with open(path, newline="") as f:
    lines = [ln for ln in f if not ln.startswith("==")]
rd = csv.DictReader(lines)
tot, by = 0.0, defaultdict(lambda: [0, 0.0])
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
    by[r["Kernel Name"]][0] += 1
    by[r["Kernel Name"]][1] += us
    tot += us
n = sum(c for c, _ in by.values())
print(cmd)
print(f"(one step between cudaProfilerStart/Stop; {n} kernel launches, {tot / 1e3:.1f} ms serialised, cold-cache per-launch times)")
print(" share  launches  time_us  kernel")
for k, (c, t) in sorted(by.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{100 * t / tot:6.2f}%  {c:5d}  {t:11.1f}  {k[:120]}")
