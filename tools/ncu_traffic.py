#!/usr/bin/env python
"""profiles/traffic.json from `ncu --set full` captures (run here, no GPU needed).

  python tools/ncu_traffic.py <tag> [summary.txt]     reads gpurun_out/<tag>_k_*.ncu-rep

For every kernel name the capture with the LARGEST grid x duration is kept (the level-0 launch of the multigrid kernels;
the step-sized launch of the particle kernels): DRAM bytes read / written, duration, DRAM and SM utilisation, occupancy.
bench.py attaches the table to its JSON line (`roofline.ncu_traffic`) and reads the k_spmv entry for `roofline.traffic`."""
import csv, glob, io, json, os, re, subprocess, sys

tag = sys.argv[1]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
M = {"gpu__time_duration.sum": "duration", "dram__bytes_read.sum": "dram_bytes_read", "dram__bytes_write.sum": "dram_bytes_write",
     "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct",
     "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct", "launch__registers_per_thread": "registers",
     "l1tex__t_sector_hit_rate.pct": "l1_hit_pct", "lts__t_sector_hit_rate.pct": "l2_hit_pct"}
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "nsecond": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3}


def num(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


best = {}
reps = sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"{tag}_k_*.ncu-rep")))
if not reps:          # the reports stayed on the GPU box (too large): their raw pages came back as CSV
    reps = sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"{tag}_k_*_raw.csv")))


def grid_size(e):
    g = [int(x) for x in re.findall(r"\d+", e.get("grid") or "0")]
    n = 1
    for x in g:
        n *= x
    return n


for rep in reps:
    txt = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    if len(rows) < 3:
        continue
    h, units = rows[0], rows[1]
    for r in rows[2:]:
        d, u = dict(zip(h, r)), dict(zip(h, units))
        name = re.sub(r"^(void )?(<unnamed>::|\(anonymous namespace\)::)?", "", d.get("Kernel Name", "?")).split("(")[0].split("<")[0].strip()
        e = {"grid": d.get("Grid Size"), "block": d.get("Block Size"), "capture": os.path.basename(rep)}
        for k, nm in M.items():
            if k in d and num(d[k]) is not None:
                e[nm] = num(d[k]) * SCALE.get(u[k], 1)
        if "duration" not in e:
            continue
        e["duration_us"] = e.pop("duration")
        # the LAST capture at the largest grid: level 0 of the multigrid kernels, the steady (already sorted) step of the sort kernels
        if name not in best or grid_size(e) >= grid_size(best[name]):
            best[name] = e
for e in best.values():
    if "dram_bytes_read" in e and "dram_bytes_write" in e:
        e["dram_gbs"] = (e["dram_bytes_read"] + e["dram_bytes_write"]) / e["duration_us"] * 1e-3
out = {"_source": f"ncu --set full --clock-control none (tools/gpu_ncu.sh {tag}), 256^3 full tank, 127 M particles, default parameters; "
                  "per kernel the last launch captured at its largest grid (level 0 of the multigrid kernels); cold-cache, serialised"}
out["_grid"] = [256, 256, 256]
out.update(dict(sorted(best.items())))
if "k_spmv4f" in best:
    out["k_spmv"] = dict(best["k_spmv4f"], alias_of="k_spmv4f")
# on the GPU box the result goes to gpurun_out/ (the .ncu-rep files are too large to travel back); copy it to profiles/ here
dst = os.path.join(ROOT, "gpurun_out", f"{tag}_traffic.json") if os.environ.get("GRAFT_REPO_ROOT") or "--box" in sys.argv else os.path.join(ROOT, "profiles", "traffic.json")
json.dump(out, open(dst, "w"), indent=1)
lines = [f"{'kernel':28s} {'us':>8s} {'rd MB':>8s} {'wr MB':>8s} {'GB/s':>7s} {'dram%':>6s} {'sm%':>6s} {'warps%':>6s} {'regs':>5s}  grid"]
for n, e in sorted(best.items(), key=lambda x: -x[1]["duration_us"]):
    lines.append(f"{n:28s} {e['duration_us']:8.1f} {e.get('dram_bytes_read', 0)/1e6:8.1f} {e.get('dram_bytes_write', 0)/1e6:8.1f} "
                 f"{e.get('dram_gbs', 0):7.0f} {e.get('dram_pct', 0):6.1f} {e.get('sm_pct', 0):6.1f} {e.get('warps_active_pct', 0):6.1f} "
                 f"{int(e.get('registers', 0)):5d}  {e['grid']} x {e['block']}")
txt = "\n".join(lines)
print(txt)
if len(sys.argv) > 2 and not sys.argv[2].startswith('--'):
    open(sys.argv[2], "w").write(out["_source"] + "\n\n" + txt + "\n")
