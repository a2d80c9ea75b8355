#!/usr/bin/env python
"""bench.py -- FLIP particle-steps/s (BASELINE.json metric) on N GPUs of one node.

  python bench.py --gpus N --steps K --warmup W            # CUDA path (this repo)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU implementation (oracle/_ref)

A "step" is one pass of the hot path (fluid.cc:1378-1490: sort, P2G, classify, relaxation loop with the
pressure solves, G2P, advection) over the synthetic dam-break full-tank scene (fill (i)):
n^3 grid, exactly 8 particles per fluid voxel, zero initial velocity.  Default n = 256 (BASELINE configs[2],
the configuration the metric is quoted on; ~127 M particles, fits one B200).

Prints ONE JSON line (rank 0):
  value        particles x K / device time, inputs resident in HBM
  e2e          the same through flipb200_step_host with pinned HOST buffers (H2D + step + D2H per step)
  roofline     the PCG stencil kernel (17 algorithmic bytes per swept node) against the measured HBM peak,
               .by_kernel = every kernel of the step with its algorithmic bytes, .dominant = largest share
  cpu_baseline the CPU oracle on a bounded sample + the CUDA path on that SAME sample (same_config_gpu)
  check        correctness of THIS run, computed on all N GPUs before the line is printed; a failed check
               makes the process exit non-zero: hydrostatic pressure of the tank at rest against rho|g|depth on
               every rank's planes (through the z-slab halo / all-reduce / gather path), closed-form cell
               count, particle conservation over ranks, pass and iteration counts identical on every rank
  pcg          BASELINE configs[4]: pressure projection only, PCG iterations/s, weak scaling
  cube_drop_512 BASELINE configs[3]: cube drop in a 512^3 box cut into N z-slabs, STRONG scaling, with a
               3-step agreement check of the slab run against a single-GPU run of the same scene (N >= 2)
  dam_break_column  SURVEY 8d fill (ii): the true dam break (column x < 0.4n, y < 0.6n) at the same n
  fp32_mode    (N = 1) params.precision = 32 on the headline scene, with its distance from the precision = 64 run
  grid_128, default_scene (N = 1)  BASELINE configs[1] and configs[0] (the reference's own 121^3 cube drop)

Single workloads: --workload flip | dam-break-column | cube-drop | pcg (with --grid-n).  --no-extra keeps
only the headline workload.

N > 1 (torchrun, one rank per GPU): ONE tank of n x n x (n*N) nodes cut into N z-slabs --
weak scaling, every GPU owns an n^3 block and ~the same particles; ghost planes, PCG dot products and
particle migration go over the library's own NCCL communicator / NVLink peer stores.  `value` counts the
particles of the whole tank; time = max over ranks of the device time of the K collective steps.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "flip_particle_steps_per_s"
UNIT = "particle-steps/s"
CPU_SAMPLE_N = 48          # grid of the bounded CPU sample (both arms time the SAME scene: full tank, n = 48)
HYDRO_DT = 0.125           # exact in fp32: the reference's float-stored matrix coefficients carry no rounding
PCG_BLOCK = 200            # iterations per "step" of the pcg workload

# algorithmic bytes per swept node of the level-0 stages of one MG-PCG iteration (fp64 CG vectors, fp32 V-cycle,
# 1-byte flags; SURVEY 8(d) counting rule: every value that must be read or written once counts once)
STAGE_BYTES = {
    "k_spmv": 17.0,                    # read p 8, flag 1; write Ap 8
    "k_update_xr": 33.0,               # read r, Ap 16, flag 1; write r 8, b0 4, x1 4
    "k_update_p": 37.0,                # read p, x 16, z 4, flag 1; write p, x 16
    "L0 halo+smooth": 13.0,            # read b, x 8, flag 1; write x' 4
    "L0 halo+resid": 14.0,             # + normaliser byte
    "L0 halo+restrict": 4.5,           # read residual 4; write coarse b, x (1/8 of the nodes)
    "L0 halo+prolong": 10.5,
    "L0 halo+smooth (post 1)": 13.0,
    "L0 halo+smooth (post 2)": 13.0,   # also accumulates r.z from the values it already holds
}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.samples = []
        self._stop = threading.Event()
        self._t = None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.25)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0])); mx = max(mx, float(s[1]))
                for nm, v in zip(names, s[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def full_tank_box(n, nz=None):
    """SURVEY 8d fill (i): every interior voxel below a 2-cell air gap (wall = 2)."""
    nz = n if nz is None else nz
    return (2, 2, 2), (n - 3, n - 5, nz - 3)


def column_box(n, nz=None):
    """SURVEY 8d fill (ii): the true dam break, column x < 0.4 n, y < 0.6 n, full z."""
    nz = n if nz is None else nz
    return (2, 2, 2), (int(0.4 * n) - 1, int(0.6 * n) - 1, nz - 3)


def cube_box(n):
    """the reference's cube drop (41/121 of the box, fluid.cc:1176) scaled to n^3: centred cube of side n/3"""
    side = n // 3
    lo = tuple((n - side) // 2 for _ in range(3))
    return lo, tuple(l + side - 1 for l in lo)


def box_count(lo, hi):
    return (hi[0] - lo[0] + 1) * (hi[1] - lo[1] + 1) * (hi[2] - lo[2] + 1)


# -------------------------------------------------------------------------------------------------
# CPU arms (the ONLY code in this file that touches oracle/)
# -------------------------------------------------------------------------------------------------
def cpu_sample_config():
    n = CPU_SAMPLE_N
    lo, hi = full_tank_box(n)
    return {"workload": f"BOUNDED SAMPLE of the headline workload: dam-break FLIP full-tank fill, {n}x{n}x{n} grid, 8 particles/cell "
                        f"({box_count(lo, hi) * 8} particles), fp64",
            "grid": [n, n, n], "particles_total": box_count(lo, hi) * 8,
            "sample_of": "dam-break FLIP full-tank fill, 256x256x256 grid, 8 particles/cell (127008000 particles), BASELINE configs[2]"}


def run_cpu_sample(steps, warmup, threads):
    """The CPU arm: oracle/_ref = the restatement of fluid.cc (pinned bit for bit to the reference's own code,
    tests/test_ref_pin.py) + the reference's own Eigen IC-PCG, on the full-tank scene at n = CPU_SAMPLE_N."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py
    from fluid_simulation_b200 import scenes
    n = CPU_SAMPLE_N
    sc = scenes.dam_break(n=n, fill="full")
    o = oracle_py.Oracle(nx=n, ny=n, nz=n, ox=n // 2, oy=n // 2, oz=n // 2, threads=threads)
    o.set_particles(sc.pos, sc.vel)
    dt = 0.1
    for _ in range(warmup):
        dt, _st = o.step(dt)
    t0 = time.perf_counter()
    iters = passes = 0
    for _ in range(steps):
        dt, st = o.step(dt)
        iters += st.pcg_iters_total; passes += st.passes
    el = time.perf_counter() - t0
    kind = "reference" if o.has_eigen else "port"
    desc = (f"dam-break full-tank fill at {n}^3 ({sc.num_particles} particles), {steps} full steps after {warmup} warm-up: "
            f"oracle/_ref = restatement of fluid.cc pinned bit for bit to the reference's own compiled fluid.cc (tests/test_ref_pin.py) "
            f"with the reference's vendored Eigen IC-PCG (single-threaded, as run.sh builds it) and OpenMP particle loops on {threads} threads "
            f"(standing in for tbb::parallel_for); {passes} relaxation passes, {iters} PCG iterations; generous to the CPU: the dense-array "
            f"restatement drops the reference's ValueAccessor / VDB-tree overhead, and IC-PCG iterations grow ~linearly with n")
    return {"value": sc.num_particles * steps / el, "ms_per_step": el / steps * 1e3, "kind": kind, "sample": desc,
            "eigen_from_reference": bool(o.has_eigen), "grid_n": n, "particles": sc.num_particles, "steps": steps, "warmup": warmup,
            "passes": passes, "pcg_iters": iters}


def run_reference_main(steps):
    """The reference ITSELF: /root/reference/fluid.cc main() compiled unmodified over the dense-array OpenVDB/TBB
    stand-in (oracle/_ref/libfluid_ref.so), run for `steps` steps of its own loop on its own scene (BASELINE configs[0])."""
    try:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import ref_py
        if not ref_py.available():
            return {"unavailable": "oracle/_ref/libfluid_ref.so not built (needs /root/reference at build time)"}
        t0 = time.perf_counter()
        m = ref_py.MainRun(steps)
        el = time.perf_counter() - t0
        n = int(m.pos.shape[0])
        return {"value": n * steps / el, "unit": UNIT, "ms_per_step": el / steps * 1e3, "steps": steps, "particles": n,
                "grid": [121, 121, 121], "passes": m.passes, "threads": 1,
                "what": "fluid.cc main() unmodified (scene build + its own step loop) over oracle/shim; tbb::parallel_for stand-in serial "
                        "(a threaded stand-in measured slower), VDB containers replaced by dense arrays (faster than the real tree); "
                        "time includes the one-off scene construction"}
    except Exception as e:     # never let the side measurement break the arm
        return {"unavailable": f"{type(e).__name__}: {e}"}


def reference_arm(args):
    threads = os.cpu_count() or 1
    r = run_cpu_sample(args.steps, max(args.warmup, 0), threads)
    cfg = cpu_sample_config()
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": threads, "kind": r["kind"], "sample": r["sample"],
                             "eigen_from_reference": r["eigen_from_reference"], "grid_n": r["grid_n"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "the reference hard-codes a 121^3 box (fluid.cc:1159) and solves with a serial IC-PCG: it cannot run the 256^3 headline "
                    "workload, so this arm times a bounded sample (config.grid) of the same scene family; the CUDA arm reports its own "
                    "throughput on this exact sample as cpu_baseline.same_config_gpu, and the reference's own scene as default_scene"}
    if not args.no_extra:
        line["reference_main"] = run_reference_main(3)
    print(json.dumps(line))
    return 0


# -------------------------------------------------------------------------------------------------
# CUDA arm
# -------------------------------------------------------------------------------------------------
class Job:
    """rank / world plumbing shared by the workloads"""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        import fluid_simulation_b200 as fb
        from fluid_simulation_b200 import dist as fdist
        self.torch, self.dist, self.fb, self.fdist, self.args = torch, dist, fb, fdist, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (the FLIP path has no CPU fallback)")
        torch.cuda.set_device(self.local_rank)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
        self.peak, self.peak_src = measured_peak_gbs()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, x, op="max"):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op={"max": self.dist.ReduceOp.MAX, "min": self.dist.ReduceOp.MIN, "sum": self.dist.ReduceOp.SUM}[op])
        return float(t.item())

    def gather_ints(self, vals):
        """list of python ints on this rank -> list (per rank) of lists, on every rank"""
        if self.world == 1:
            return [list(vals)]
        t = self.torch.tensor(list(vals), dtype=self.torch.int64, device="cuda")
        out = [self.torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [o.cpu().tolist() for o in out]

    def make(self, n, nzg, cap, world=None, rank=None, origin=None, zsplit=None, **kw):
        world = self.world if world is None else world
        rank = self.rank if rank is None else rank
        o = origin if origin is not None else (n // 2, n // 2, nzg // 2)
        if world == 1:
            return self.fb.Flip(nx=n, ny=n, nz=nzg, ox=o[0], oy=o[1], oz=o[2], max_particles=cap, precond=self.args.precond,
                                device=self.local_rank, **kw)
        return self.fdist.create_slab_flip(self.fb, rank, world, device=self.local_rank, nx=n, ny=n, nz=nzg, ox=o[0], oy=o[1], oz=o[2],
                                           max_particles=cap, precond=self.args.precond, zsplit=zsplit, **kw)

    def finish(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def hydrostatic_check(job, f, n, nzg, top):
    """Tank at rest (seeded, v = 0): p2g, classify, one pressure solve with dt exact in fp32.  The converged pressure must be
    rho*|g|*dx*depth on EVERY owned plane of every rank (rtol 1e-9, the bound tests/test_gpu_parity.py uses on one GPU), which on
    N GPUs goes through the fused ghost-plane pushes, the mailbox all-reduce and the replicated-level gather.  Returns a dict."""
    fb = job.fb
    f.p2g()
    n_fluid = f.classify()
    f.rhs_div(HYDRO_DT)
    iters, rel = f.pcg_solve(HYDRO_DT)
    p = f.get_grid(fb.GRID_P)                           # owned planes (z1-z0, ny, nx)
    expect = 10.0 * (top - np.arange(2, top))           # rho |g| dx (top - y), fluid.cc:1357-1358
    zg = np.arange(f.z0, f.z1)
    zsel = np.nonzero((zg >= 2) & (zg <= nzg - 3))[0]   # planes inside the shell
    worst = 0.0
    if zsel.size:
        cols = p[zsel][:, 2:top, 2:n - 2]               # every interior column of every owned plane
        worst = float(np.max(np.abs(cols - expect[None, :, None]) / expect[None, :, None]))
    worst = job.reduce(worst, "max")
    expected_fluid = (n - 4) * (n - 5) * (nzg - 4)      # every interior node up to the layer the splat reaches above the surface
    its = job.gather_ints([iters, n_fluid])
    same = all(v == its[0] for v in its)
    ok = bool(worst <= 1e-9 and rel < 2.220446049250313e-16 and n_fluid == expected_fluid and same)
    return {"ok": ok, "hydrostatic_max_rel_err": worst, "rtol": 1e-9, "columns": "all interior columns of all owned planes of all ranks",
            "pcg_iters": iters, "pcg_rel_resid": rel, "n_fluid": n_fluid, "n_fluid_closed_form": expected_fluid,
            "same_on_all_ranks": same, "dt": HYDRO_DT}


def run_steps(job, f, steps, warmup, dt=0.1, sample_clocks=True):
    """W untimed + K timed collective steps; returns (dev_s max over ranks, wall, stats list, clocks, launches, dt)."""
    for _ in range(warmup):
        dt, st = f.step(dt)
    sampler = ClockSampler(job.local_rank)
    job.barrier()
    if sample_clocks and job.rank == 0 and not os.environ.get("FLIPB200_BENCH_NOSMI"):
        sampler.start()
    t0 = time.perf_counter()
    dev_ms = 0.0
    launches = 0
    stats = []
    for _ in range(steps):
        dt, st = f.step(dt)
        dev_ms += st.ms_total; launches += st.launches
        stats.append(st.as_dict())
    job.barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    return job.reduce(dev_ms * 1e-3), job.reduce(wall), stats, clocks, launches, dt


def step_consistency(job, f, stats, total_particles):
    """after the timed steps: particles conserved over ranks, pass / iteration / cell counts identical on every rank"""
    n_sum = int(round(job.reduce(float(f.num_particles), "sum")))
    seq = []
    for s in stats:
        seq += [s["passes"], s["pcg_iters_total"], s["n_fluid"]]
    per_rank = job.gather_ints(seq)
    same = all(v == per_rank[0] for v in per_rank)
    finite = all(np.isfinite(s["last_error"]) or s["passes"] == 1 for s in stats) and all(np.isfinite(s["dt_out"]) and s["dt_out"] > 0 for s in stats)
    conv = all(s["pcg_rel_resid"] < 2.220446049250313e-16 or s["pcg_iters_last"] == 0 for s in stats)
    return {"ok": bool(n_sum == total_particles and same and finite and conv), "particles_after": n_sum, "particles_expected": total_particles,
            "passes_iters_cells_same_on_all_ranks": same, "pcg_converged_to_2.2e-16_every_pass": conv}


def per_step(stats):
    keys = ("passes", "pcg_iters_total", "n_fluid", "ms_sort", "ms_p2g", "ms_grid", "ms_pcg", "ms_g2p", "ms_total")
    return [{k: (round(s[k], 3) if isinstance(s[k], float) else s[k]) for k in keys} for s in stats]


def stage_bytes(mixed):
    """STAGE_BYTES with the storage choices of params.mixed: bit 0 = CG search direction as fp32, bit 1 = level-0 b0 and residual as fp16"""
    b = dict(STAGE_BYTES)
    if mixed & 1:
        b["k_spmv"] = 13.0          # read p 4, flag 1; write Ap 8
        b["k_update_p"] = 29.0      # read p 4, x 8, z 4, flag 1; write p 4, x 8
    if mixed & 2:
        b["k_update_xr"] = 31.0     # b0 as fp16
        for k in ("L0 halo+smooth", "L0 halo+smooth (post 1)", "L0 halo+smooth (post 2)"):
            b[k] = 11.0             # read b0 2, x 4, flag 1; write x' 4
        b["L0 halo+resid"] = 10.0   # read b0 2, x 4, flag + normaliser 2; write residual 2
        b["L0 halo+restrict"] = 2.5
    return b


def iteration_profile(job, f, nodes, mixed=0):
    """stage profile of one PCG iteration -> by_kernel entries + the list of small-level stages"""
    SB = stage_bytes(mixed)
    prof = f.profile_iteration(10)
    by = {}
    agg = {}
    for name, us in prof:
        a = agg.setdefault(name, [0.0, 0])
        a[0] += us; a[1] += 1
    total = sum(us for _n, us in prof)
    for name, (us, cnt) in agg.items():
        e = {"us_per_iteration": us, "launches_per_iteration": cnt}
        if name in SB and us > 0:
            bpn = SB[name]
            e["bytes_per_node"] = bpn
            e["gbs"] = bpn * nodes * cnt / (us * 1e-6) / 1e9
            e["frac"] = e["gbs"] / job.peak
        by[name] = e
    return by, total


def bench_flip(job, n, fill, steps, warmup, headline, tag):
    """dam-break FLIP (full tank = headline, or the column) on an n x n x (n*world) tank, weak scaling"""
    args, fb, world, rank = job.args, job.fb, job.world, job.rank
    nzg = n * world
    lo, hi = full_tank_box(n, nzg) if fill == "full" else column_box(n, nzg)
    total_particles = box_count(lo, hi) * 8
    particles = total_particles // world
    cap = int(particles * 1.10) + 65536
    cfg_idx = {256: 2, 128: 1}.get(n, "-")
    config = {"workload": f"dam-break FLIP {'full-tank' if fill == 'full' else 'column (x < 0.4n, y < 0.6n)'} fill, {n}x{n}x{nzg} grid, 8 particles/cell "
                          f"({total_particles} particles, ~{particles}/GPU), fp64, {'MG-' if args.precond == 1 else 'Jacobi-'}PCG to 2.2e-16"
                          + (f" (BASELINE configs[{cfg_idx}]" + (" per GPU" if world > 1 else "") + ")" if fill == "full" else " (SURVEY 8d fill (ii))"),
              "grid": [n, n, nzg], "particles_per_gpu": particles, "particles_total": total_particles,
              "parallelism": "single GPU" if world == 1 else f"{world} z-slabs of {n} planes, NVLink peer-store halo exchange + PCG all-reduces + particle migration",
              "l2": "inputs larger than L2 (particle arrays 48 B x particles, grid vectors 8 B x n^3)" if n >= 256
                    else "working set may fit L2 at this n; no flush between steps"}
    f = job.make(n, nzg, cap)
    f.seed_box(lo, hi, seed=1234)
    check = {}
    if fill == "full":
        check["hydrostatic"] = hydrostatic_check(job, f, n, nzg, top=2 + (n - 5))
    dev_s, wall, stats, clocks, launches, dt = run_steps(job, f, steps, warmup, sample_clocks=headline)
    value = total_particles * steps / dev_s
    check["steps"] = step_consistency(job, f, stats, total_particles)
    check["ok"] = all(v["ok"] for v in check.values() if isinstance(v, dict))
    config["engine"] = f.debug_info()
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": dev_s / steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config, "gpu_launches": int(launches), "check": check,
            "wall_ms_per_step": wall / steps * 1e3, "value_basis": "device time (CUDA events on the library stream, max over ranks); "
            "wall_ms_per_step is the host clock around the same K steps"}
    if headline:
        line["clocks"] = clocks
    # ---- kernel rooflines ----
    peak = job.peak
    f.p2g(); f.classify(); f.rhs_div(dt)
    ms_spmv, nodes = f.time_kernel(fb.KERNEL_SPMV, args.kernel_iters)
    ms_iter, _ = f.time_kernel(fb.KERNEL_PCG_ITER, args.kernel_iters)
    ms_p2g, npart = f.time_kernel(fb.KERNEL_P2G, 5)
    ms_sort, _ = f.time_kernel(fb.KERNEL_SORT, 5)
    # params.mixed (rows >= 128 nodes, also on z-slabs): the product's stencil reads the search direction as fp32
    mixed = int(getattr(f.params, "mixed", 0)) if (n >= 128 and args.precond == 1) else 0
    if os.environ.get("FLIPB200_MIXED") is not None and mixed:
        mixed = int(os.environ["FLIPB200_MIXED"])
    SB = stage_bytes(mixed)
    spmv_bytes = SB["k_spmv"]
    ach = spmv_bytes * nodes / (ms_spmv * 1e-3) / 1e9
    traffic = None          # dram__bytes_read.sum + dram__bytes_write.sum of one launch, from the committed ncu capture
    tj_all = {}
    try:
        tj_all = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        tj = tj_all["k_spmv4f" if mixed & 1 else "k_spmv"]
        if tj_all.get("_grid", tj.get("grid")) == [n, n, n] and world == 1:
            traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
    except Exception:
        pass
    by = {}
    ms_g2p_step = float(np.mean([s["ms_g2p"] for s in stats]))
    ms_sort_step = float(np.mean([s["ms_sort"] for s in stats]))
    ms_part_step = float(np.mean([s["ms_sort"] + s["ms_p2g"] + s["ms_g2p"] for s in stats]))
    ms_total_step = float(np.mean([s["ms_total"] for s in stats]))
    iters_step = float(np.mean([s["pcg_iters_total"] for s in stats]))

    def pk(bytes_per, units, ms, **kw):
        g = bytes_per * units / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        return dict({"bytes_per_unit": bytes_per, "ms": ms, "gbs": g, "frac": g / peak, "share_of_step": ms / ms_total_step if ms_total_step > 0 else None}, **kw)

    by["p2g"] = pk(52.0, npart, ms_p2g, unit="particle", what="memsets + P2G kernel, flipb200_time_kernel(P2G)")
    by["g2p_advect"] = pk(99.1, npart, ms_g2p_step, unit="particle", what="flipb200_stats.ms_g2p (cell delta + G2P + advection), mean of the timed steps")
    by["sort"] = pk(0.0, npart, ms_sort_step, unit="particle", what="cell sort: overhead, no algorithmic bytes (SURVEY 8d)")
    by["particle_phase"] = pk(151.1, npart, ms_part_step, unit="particle", what="(52 + 99.1) B x particles / (ms_sort + ms_p2g + ms_g2p)")
    try:
        stage, us_total = iteration_profile(job, f, nodes, mixed)
    except Exception as e:          # Jacobi mode has no stage marks beyond the three CG kernels
        stage, us_total = {}, 0.0
    coarse_us = 0.0
    for name, e in stage.items():
        e["share_of_step"] = e["us_per_iteration"] * 1e-3 * iters_step / ms_total_step if ms_total_step > 0 else None
        by["pcg/" + name] = e
        if not (name.startswith("L0") or name.startswith("k_") or name.startswith("allreduce") or name.startswith("halo") or name == "start"):
            coarse_us += e["us_per_iteration"]
    by["pcg_iteration"] = {"us": ms_iter * 1e3, "iters_per_s": 1e3 / ms_iter, "share_of_step": ms_iter * iters_step / ms_total_step if ms_total_step > 0 else None,
                           "us_plain_launch_sum": us_total, "us_levels_below_0": coarse_us,
                           "jacobi_equivalent_bytes_per_node": 91, "jacobi_equivalent_frac": 91.0 * nodes / (ms_iter * 1e-3) / 1e9 / peak,
                           "level0_algorithmic_bytes_per_node": sum(SB[k] * stage[k]["launches_per_iteration"] for k in SB if k in stage),
                           "what": "CUDA-graph replay of the whole iteration, flipb200_time_kernel(PCG_ITER)"}
    # dominant = the kernel family with the largest share of the step's device time
    fam = {"k_mg_smooth (level 0, pre + 2 post sweeps)": sum(stage[k]["us_per_iteration"] for k in stage if k.startswith("L0 halo+smooth")),
           "k_mg_resid (level 0)": stage.get("L0 halo+resid", {}).get("us_per_iteration", 0.0),
           "k_mg_restrict (level 0)": stage.get("L0 halo+restrict", {}).get("us_per_iteration", 0.0),
           "k_mg_prolong (level 0)": stage.get("L0 halo+prolong", {}).get("us_per_iteration", 0.0),
           "k_spmv": stage.get("k_spmv", {}).get("us_per_iteration", 0.0),
           "k_update_xr": stage.get("k_update_xr", {}).get("us_per_iteration", 0.0),
           "k_update_p": stage.get("k_update_p", {}).get("us_per_iteration", 0.0),
           "multigrid levels >= 1 (latency-bound launches)": coarse_us}
    fam_ms = {k: v * 1e-3 * iters_step for k, v in fam.items()}
    fam_ms["p2g"] = float(np.mean([s["ms_p2g"] for s in stats])); fam_ms["g2p + advect"] = ms_g2p_step; fam_ms["cell sort"] = ms_sort_step
    dom = max(fam_ms, key=lambda k: fam_ms[k])
    dominant = {"kernel": dom, "ms_per_step": fam_ms[dom], "share_of_step": fam_ms[dom] / ms_total_step if ms_total_step > 0 else None}
    if dom.startswith("k_mg_smooth"):
        us = fam[dom]; cnt = sum(stage[k]["launches_per_iteration"] for k in stage if k.startswith("L0 halo+smooth"))
        g = SB["L0 halo+smooth"] * nodes * cnt / (us * 1e-6) / 1e9
        dominant.update({"bytes_per_node": SB["L0 halo+smooth"], "launches_per_iteration": cnt, "gbs": g, "frac": g / peak})
    line["roofline"] = {"bound": "hbm", "kernel": "k_spmv (7-pt matrix-free A.p + p.Ap" + ("; search direction stored as fp32, A.p and p.Ap evaluated in fp64)" if mixed & 1 else ")"),
                        "achieved": ach, "peak": peak, "unit": "GB/s",
                        "frac": ach / peak, "traffic": traffic, "traffic_source": "profiles/traffic.json (ncu, per launch)" if traffic else None,
                        "algorithmic_bytes": spmv_bytes * nodes, "peak_source": job.peak_src, "algorithmic_bytes_per_node": spmv_bytes, "swept_nodes": nodes,
                        "mixed": mixed,
                        "us_per_launch": ms_spmv * 1e3, "dominant": dominant, "by_kernel": by,
                        "ncu_traffic": {k: v for k, v in tj_all.items() if isinstance(v, dict)} if world == 1 and n == 256 else None}
    line["per_step"] = per_step(stats)

    # ---- e2e: the same step through the host-buffer entry point ----
    if headline and not args.no_e2e:
        import ctypes as C
        nloc = f.num_particles
        capn = nloc + nloc // 16 + 4096 if world > 1 else nloc
        arrs, ptrs = [], []
        for q in range(6):
            a, p = fb.pinned_empty(capn)
            arrs.append(a); ptrs.append(p)
        m = C.c_int64(0)
        fb._check(f.lib.flipb200_get_particles(f.h, *[fb._ptr(a) for a in arrs], capn, C.byref(m)), "get")
        out, outp = fb.pinned_empty(f.N, np.float32)          # this rank's planes of outputGrid
        dte = dt

        def host_step(nl, d):
            if world == 1:
                d, _st = f.step_host(arrs, d, output_grid=out)
                return nl, d
            d, _st, nl = f.step_host_slab(arrs, nl, d, output_grid=out)
            return nl, d

        nloc, dte = host_step(nloc, dte)                       # warm-up
        job.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            nloc, dte = host_step(nloc, dte)
        job.barrier()
        el = job.reduce(time.perf_counter() - t0)
        line["e2e"] = {"value": total_particles * steps / el, "unit": UNIT,
                       "h2d_bytes_per_step": 48 * total_particles, "d2h_bytes_per_step": 48 * total_particles + 4 * n * n * nzg,
                       "ms_per_step": el / steps * 1e3,
                       "api": ("flipb200_step_host" if world == 1 else "flipb200_step_host_slab per rank") + " (pinned host SoA in/out + outputGrid)"}
        for p in ptrs + [outp]:
            fb.pinned_free(p)
    elif headline:
        line["e2e"] = None
    f.close()
    return line


def bench_small_scene(job, n, lo, hi, origin, steps, warmup, workload, single=True):
    """a single-GPU scene stepped on rank 0's GPU (configs[0], configs[1], the CPU sample's twin)"""
    fb = job.fb
    total = box_count(lo, hi) * 8
    f = job.fb.Flip(nx=n, ny=n, nz=n, ox=origin, oy=origin, oz=origin, max_particles=int(total * 1.1) + 65536, precond=job.args.precond,
                    device=job.local_rank)
    f.seed_box(lo, hi, seed=1234)
    dt = 0.1
    for _ in range(warmup):
        dt, _st = f.step(dt)
    job.torch.cuda.synchronize()
    t0 = time.perf_counter()
    stats = []
    for _ in range(steps):
        dt, st = f.step(dt)
        stats.append(st.as_dict())
    wall = time.perf_counter() - t0
    dev_s = sum(s["ms_total"] for s in stats) * 1e-3
    f.close()
    return {"workload": workload, "grid": [n, n, n], "particles": total, "steps": steps, "warmup": warmup, "value": total * steps / dev_s, "unit": UNIT,
            "ms_per_step": dev_s / steps * 1e3, "wall_ms_per_step": wall / steps * 1e3, "value_wall": total * steps / wall,
            "passes": [s["passes"] for s in stats], "pcg_iters": [s["pcg_iters_total"] for s in stats]}


def bench_fp32_mode(job, n, steps, warmup):
    """north_star "fp32 mode reported separately": params.precision = 32 (pressure solves stop at FLT_EPSILON) on the headline
    scene, single GPU.  Reports its throughput and the rel-L2 distance of its grid velocity / pressure after ONE step from the
    precision = 64 run of the same particles (the cost of the mode in accuracy, next to its gain in time)."""
    fb = job.fb
    lo, hi = full_tank_box(n)
    total = box_count(lo, hi) * 8
    cap = int(total * 1.06) + 65536
    ref = {}
    for prec in (64, 32):
        f = job.make(n, n, cap, precision=prec)
        f.seed_box(lo, hi, seed=1234)
        dt, st = f.step(0.1)
        g = {"u": np.stack([f.get_grid(fb.GRID_U + a) for a in range(3)]), "p": f.get_grid(fb.GRID_P), "iters": st.pcg_iters_total}
        if prec == 64:
            ref = g
            f.close()
            continue
        def rl2(a, b):
            return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))
        diff = {"grid_velocity_rel_l2": rl2(g["u"], ref["u"]), "pressure_rel_l2": rl2(g["p"], ref["p"]),
                "pcg_iters_step0": {"precision_64": ref["iters"], "precision_32": g["iters"]}}
        for _ in range(max(0, warmup - 1)):
            dt, st = f.step(dt)
        stats = []
        for _ in range(steps):
            dt, st = f.step(dt)
            stats.append(st.as_dict())
        f.close()
    dev_s = sum(s["ms_total"] for s in stats) * 1e-3
    return {"workload": f"dam-break FLIP full-tank fill, {n}^3, 8 particles/cell, params.precision = 32 (pcg_tol = FLT_EPSILON; state and CG arithmetic fp64)",
            "particles": total, "steps": steps, "warmup": warmup, "value": total * steps / dev_s, "unit": UNIT, "ms_per_step": dev_s / steps * 1e3,
            "passes": [s["passes"] for s in stats], "pcg_iters": [s["pcg_iters_total"] for s in stats],
            "pcg_rel_resid": [s["pcg_rel_resid"] for s in stats], "vs_precision_64_after_one_step": diff}


def bench_pcg(job, n):
    """BASELINE configs[4]: pressure-projection-only weak scaling.  The system is the one the first relaxation pass of the
    full-tank scene solves (b = divergence after one gravity kick); a step = PCG_BLOCK iterations of the product's CUDA-graph
    iteration without convergence exit (flipb200_time_kernel, CUDA events on the library's stream, max over ranks); the
    run-to-tolerance iteration count of the same system is reported next to it."""
    args, fb, world, rank = job.args, job.fb, job.world, job.rank
    nzg = n * world
    lo, hi = full_tank_box(n, nzg)
    total = box_count(lo, hi) * 8
    f = job.make(n, nzg, int(total // world * 1.06) + 65536)
    f.seed_box(lo, hi, seed=1234)
    check = {"hydrostatic": hydrostatic_check(job, f, n, nzg, top=2 + (n - 5))}
    dt = 0.1
    n_fluid = f.classify()
    f.rhs_div(dt)
    iters_tol, rel = f.pcg_solve(dt)                       # run to 2.2e-16 (also the warm-up of the iteration graph)
    blocks = max(3, min(10, args.steps))
    for _ in range(3):
        f.time_kernel(fb.KERNEL_PCG_ITER, 20)
    job.barrier()
    ms_block = 0.0
    nodes = 0
    for _ in range(blocks):
        ms_iter, nodes = f.time_kernel(fb.KERNEL_PCG_ITER, PCG_BLOCK)
        ms_block += ms_iter * PCG_BLOCK
    job.barrier()
    info = f.debug_info()
    ms_block = job.reduce(ms_block)
    it_s = PCG_BLOCK * blocks / (ms_block * 1e-3)
    ms_spmv, _ = f.time_kernel(fb.KERNEL_SPMV, args.kernel_iters)
    ms_spmv = job.reduce(ms_spmv)
    mixed = int(getattr(f.params, "mixed", 0)) if (n >= 128 and args.precond == 1) else 0
    spmv_bytes = 13.0 if mixed & 1 else 17.0
    ach = spmv_bytes * nodes / (ms_spmv * 1e-3) / 1e9
    its = job.gather_ints([iters_tol, n_fluid])
    check["same_on_all_ranks"] = {"ok": all(v == its[0] for v in its)}
    check["ok"] = all(v["ok"] for v in check.values() if isinstance(v, dict))
    f.close()
    return {"metric": "pcg_iterations_per_s", "value": it_s, "unit": "iterations/s", "n_gpus": world, "scaling": "weak", "dtype": "f64",
            "steps": blocks, "ms_per_step": ms_block / blocks,
            "config": {"workload": f"pressure projection only: 7-point {'MG-' if args.precond == 1 else 'Jacobi-'}PCG (fp64) on the full-tank system, "
                                   f"{n}x{n}x{nzg} grid = {n}^3 per GPU, {n_fluid} unknowns, b = divergence after one gravity kick; step = {PCG_BLOCK} "
                                   f"iterations without convergence exit (BASELINE configs[4])", "grid": [n, n, nzg], "engine": info},
            "us_per_iteration": ms_block / (PCG_BLOCK * blocks) * 1e3, "iterations_to_tolerance": iters_tol, "rel_residual": rel,
            "unknowns": n_fluid, "swept_nodes_per_gpu": nodes, "node_updates_per_s": it_s * nodes * world,
            "stencil": {"kernel": "k_spmv", "us_per_launch": ms_spmv * 1e3, "gbs_per_gpu": ach, "frac": ach / job.peak, "bytes_per_node": spmv_bytes},
            "jacobi_equivalent_frac": 91.0 * nodes * it_s / 1e9 / job.peak, "peak": job.peak,
            "gpu_launches": int(info["kernels_per_iteration"] * PCG_BLOCK * blocks), "check": check}


def bench_cube_drop(job, n, steps, warmup):
    """BASELINE configs[3]: the reference's cube drop scaled to an n^3 box cut into `world` z-slabs -- STRONG scaling.
    For world >= 2 the first 3 steps are also run on ONE GPU (rank 0) and compared: pass counts and cell counts exact, iteration
    counts within 2, dt and max speed to 1e-9 -- real particle migration and an unbalanced particle load go through this check."""
    args, fb, world, rank = job.args, job.fb, job.world, job.rank
    lo, hi = cube_box(n)
    side = hi[0] - lo[0] + 1
    total = side ** 3 * 8
    # static load balancing along z: the cube only occupies the planes lo-1 .. hi+1, so the slab boundaries are chosen to
    # give every rank the same share of THOSE planes (flipb200_create_slab_split); --even-split keeps the even cut
    zs = job.fdist.slab_split(n, world)[0] if (args.even_split or world == 1) else job.fdist.balanced_split(n, world, lo[2] - 1, hi[2] + 2)
    own = max(0, min(hi[2] + 1, zs[rank + 1]) - max(lo[2], zs[rank]))
    cap = int((side * side * own * 8 + side * side * 8 * 4) * 1.5) + 65536
    NCHK = 3
    ref_stats = None
    if world > 1:
        if rank == 0:
            g = fb.Flip(nx=n, ny=n, nz=n, ox=n // 2, oy=n // 2, oz=n // 2, max_particles=total + 65536, precond=args.precond, device=job.local_rank)
            g.seed_box(lo, hi, seed=1234)
            d = 0.1
            ref_stats = []
            for _ in range(NCHK):
                d, st = g.step(d)
                ref_stats.append(st.as_dict())
            g.close()
        job.barrier()
    f = job.make(n, n, cap, zsplit=None if (args.even_split or world == 1) else zs)
    f.seed_box(lo, hi, seed=1234)
    first = []
    dt = 0.1
    for _ in range(min(NCHK, warmup)):
        dt, st = f.step(dt)
        first.append(st.as_dict())
    dev_s, wall, stats, clocks, launches, dt = run_steps(job, f, steps, max(0, warmup - NCHK), dt=dt, sample_clocks=False)
    check = {"steps": step_consistency(job, f, stats, total)}
    nf0 = (side + 2) ** 3                                   # nodes the splat of a free-floating cube reaches: voxels lo..hi -> nodes lo-1..hi+1
    check["first_step_cells"] = {"ok": bool(first[0]["n_fluid"] == nf0) if first else True, "n_fluid": first[0]["n_fluid"] if first else None, "closed_form": nf0}
    if world > 1 and rank == 0 and ref_stats and first:
        worst = 0.0
        ok = True
        for a, b in zip(first, ref_stats):
            ok &= a["passes"] == b["passes"] and a["n_fluid"] == b["n_fluid"] and abs(a["pcg_iters_total"] - b["pcg_iters_total"]) <= 2 * max(1, a["passes"])
            for k in ("dt_out", "max_speed", "last_error"):
                if np.isfinite(a[k]) and np.isfinite(b[k]) and b[k] != 0:
                    worst = max(worst, abs(a[k] - b[k]) / abs(b[k]))
        ok &= worst <= 1e-9
        check["vs_single_gpu"] = {"ok": bool(ok), "steps": NCHK, "max_rel_diff_dt_maxspeed_error": worst, "rtol": 1e-9,
                                  "slab": [(s["passes"], s["pcg_iters_total"], s["n_fluid"]) for s in first],
                                  "single": [(s["passes"], s["pcg_iters_total"], s["n_fluid"]) for s in ref_stats]}
    okf = all(v["ok"] for v in check.values() if isinstance(v, dict))
    okf = bool(job.reduce(1.0 if okf else 0.0, "min") > 0.5)
    check["ok"] = okf
    info = f.debug_info()
    moved = [s["ms_sort"] for s in stats]
    f.close()
    return {"metric": METRIC, "value": total * steps / dev_s, "unit": UNIT, "n_gpus": world, "scaling": "strong", "dtype": "f64", "steps": steps, "warmup": warmup,
            "ms_per_step": dev_s / steps * 1e3, "wall_ms_per_step": wall / steps * 1e3,
            "config": {"workload": f"cube-drop FLIP, {n}^3 grid, cube of side {side} voxels, 8 particles/cell ({total} particles), fp64, "
                                   f"{'MG-' if args.precond == 1 else 'Jacobi-'}PCG to 2.2e-16 (BASELINE configs[3])", "grid": [n, n, n], "particles_total": total,
                       "parallelism": "single GPU" if world == 1 else f"{world} z-slabs with boundaries {zs} ({'even' if args.even_split else 'balanced over the planes that hold the cube'}), "
                                                                                  "halo exchange + PCG all-reduces + particle migration",
                       "slab_boundaries": zs, "engine": info},
            "gpu_launches": int(launches), "check": check, "per_step": per_step(stats), "ms_sort_mean": float(np.mean(moved)) if moved else None}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="flipb200", choices=["flipb200", "reference"])
    ap.add_argument("--grid-n", dest="n", type=int, default=int(os.environ.get("FLIPB200_BENCH_N", "256")))
    ap.add_argument("--workload", default="flip", choices=["flip", "dam-break-column", "cube-drop", "pcg"],
                    help="flip: dam-break full tank, weak scaling (headline, carries the other workloads as extra keys); "
                         "dam-break-column: SURVEY 8d fill (ii); cube-drop: n^3 box, strong scaling; pcg: pressure solve only")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="headline workload only (no pcg / cube_drop_512 / column / 128 / default scene)")
    ap.add_argument("--even-split", action="store_true", help="cube-drop on N GPUs: even z-slabs instead of the load-balanced boundaries")
    ap.add_argument("--kernel-iters", type=int, default=50)
    ap.add_argument("--precond", type=int, default=int(os.environ.get("FLIPB200_PRECOND", "1")), help="0 Jacobi, 1 multigrid V-cycle")
    args = ap.parse_args()

    if args.impl == "reference":
        # the reference's CPU implementation of the path, host cores only; rank 0 alone runs it
        if int(os.environ.get("RANK", "0")) != 0:
            return 0
        return reference_arm(args)

    job = Job(args)
    n, world, rank = args.n, job.world, job.rank
    rc = 0
    if args.workload == "pcg":
        line = bench_pcg(job, n)
        line.update({"warmup": args.warmup, "higher_is_better": True, "vs_baseline": None, "data": "synthetic", "e2e": None, "cpu_baseline": None})
    elif args.workload == "cube-drop":
        line = bench_cube_drop(job, n, args.steps, args.warmup)
        line.update({"higher_is_better": True, "vs_baseline": None, "data": "synthetic", "e2e": None, "cpu_baseline": None})
    elif args.workload == "dam-break-column":
        line = bench_flip(job, n, "column", args.steps, args.warmup, headline=True, tag="column")
        line["cpu_baseline"] = None
    else:
        line = bench_flip(job, n, "full", args.steps, args.warmup, headline=True, tag="headline")
        extra_steps = max(3, min(args.steps, 20))
        if not args.no_extra:
            # BASELINE configs[4], [3] and SURVEY 8d fill (ii) in the same invocation, so that the driver's run carries them
            line["pcg"] = bench_pcg(job, n)
            line["dam_break_column"] = {k: v for k, v in bench_flip(job, n, "column", extra_steps, 3, headline=False, tag="column").items()
                                        if k not in ("roofline", "data", "higher_is_better", "vs_baseline")}
            if n >= 256:
                line["cube_drop_512"] = bench_cube_drop(job, 512, extra_steps, 4)
            if world == 1 and rank == 0:
                line["fp32_mode"] = bench_fp32_mode(job, n, extra_steps, 3)
            if world == 1 and rank == 0 and n != 128:
                lo, hi = full_tank_box(128)
                line["grid_128"] = bench_small_scene(job, 128, lo, hi, 64, extra_steps, 3, "dam-break FLIP full-tank fill, 128^3, 8 particles/cell (BASELINE configs[1])")
            if world == 1 and rank == 0:
                line["default_scene"] = bench_small_scene(job, 121, (40, 40, 40), (80, 80, 80), 60, extra_steps, 1,
                                                          "the reference's own scene: 121^3 box, cube [-20,20]^3 (fluid.cc:1159,1176), 8 particles/voxel stratified "
                                                          "(551,368 particles; the reference scatters 689,210 = 10 per voxel) (BASELINE configs[0]); step 0 takes 8 relaxation passes")
        cpu = None
        if rank == 0 and world == 1 and not args.no_cpu:
            threads = os.cpu_count() or 1
            r = run_cpu_sample(5, 2, threads)
            cpu = {"value": r["value"], "unit": UNIT, "cores": threads, "kind": r["kind"], "sample": r["sample"],
                   "eigen_from_reference": r["eigen_from_reference"], "grid_n": r["grid_n"], "ms_per_step": r["ms_per_step"],
                   "passes": r["passes"], "pcg_iters": r["pcg_iters"]}
            lo, hi = full_tank_box(CPU_SAMPLE_N)
            g = bench_small_scene(job, CPU_SAMPLE_N, lo, hi, CPU_SAMPLE_N // 2, 5, 2, "the CUDA path on the CPU sample's scene and step window")
            cpu["same_config_gpu"] = g
            cpu["same_config_ratio_device_time"] = g["value"] / r["value"]
            cpu["same_config_ratio_wall_time"] = g["value_wall"] / r["value"]
        line["cpu_baseline"] = cpu
    oks = [line.get("check", {}).get("ok", True)] + [line[k]["check"]["ok"] for k in ("pcg", "dam_break_column", "cube_drop_512")
                                                       if isinstance(line.get(k), dict) and "check" in line[k]]
    all_ok = bool(job.reduce(1.0 if all(oks) else 0.0, "min") > 0.5)
    if rank == 0:
        line["checks_ok"] = all_ok
        print(json.dumps(line))
    job.finish()
    if not all_ok:
        if rank == 0:
            sys.stderr.write("bench.py: CHECK FAILED -- see the `check` blocks of the JSON line\n")
        rc = 3
    return rc


if __name__ == "__main__":
    sys.exit(main())
