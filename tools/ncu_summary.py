#!/usr/bin/env python
"""Summarise ncu captures into text for profiles/ (run here, no GPU needed).
  python tools/ncu_summary.py <tag>   reads gpurun_out/<tag>_*.ncu-rep and gpurun_out/<tag>_launches.csv"""
import csv, glob, io, os, subprocess, sys, collections

tag = sys.argv[1]
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "sm__maximum_warps_per_active_cycle_pct", "smsp__cycles_active.avg",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]
out = []
for rep in sorted(glob.glob(f"gpurun_out/{tag}_k_*.ncu-rep")):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    if len(rows) < 3:
        continue
    h, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(h, r)); u = dict(zip(h, units))
        out.append(f"== {d.get('Kernel Name','?')[:90]}  grid {d.get('Grid Size')} block {d.get('Block Size')}")
        for k in KEYS:
            if k in d:
                out.append(f"   {k:62s} {d[k]:>14s} {u[k]}")
# launch list -> share per kernel
p = f"gpurun_out/{tag}_launches.csv"
if os.path.exists(p):
    lines = [l for l in open(p) if l.startswith('"')]
    rows = list(csv.DictReader(io.StringIO("".join(lines))))
    tot = collections.defaultdict(float); cnt = collections.Counter()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = r["Kernel Name"].split("(")[0]
        v = float(r["Metric Value"].replace(",", ""))
        if r["Metric Unit"] in ("us", "usecond"): v *= 1e3
        if r["Metric Unit"] in ("ms", "msecond"): v *= 1e6
        tot[name] += v; cnt[name] += 1
    s = sum(tot.values())
    out.append(f"\n== launch list {p}: {sum(cnt.values())} launches, {s/1e6:.2f} ms total (cold-cache, serialised)")
    for n, v in sorted(tot.items(), key=lambda x: -x[1]):
        out.append(f"   {n:28s} {cnt[n]:6d} launches {v/1e6:9.3f} ms {100*v/s:6.2f} %  avg {v/cnt[n]/1e3:8.1f} us")
print("\n".join(out))
