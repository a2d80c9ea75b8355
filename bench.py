#!/usr/bin/env python
"""bench.py -- FLIP particle-steps/s (BASELINE.json metric) on N GPUs of one node.

  python bench.py --gpus N --steps K --warmup W            # CUDA path (this repo)
  python bench.py --impl reference --gpus N --steps K ...  # CPU oracle arm (reference's Eigen solve)

A "step" is one pass of the hot path (fluid.cc:1378-1490: sort, P2G, classify, relaxation loop with
the pressure solves, G2P, advection) over the synthetic dam-break full-tank scene (SURVEY.md 8d,
fill (i)): n^3 grid, exactly 8 particles per fluid voxel, zero initial velocity.  Default n = 256
(BASELINE configs[2], the configuration the metric is quoted on; ~127 M particles, fits one B200).

Prints ONE JSON line (rank 0).  `value` = particles x K / device time with inputs resident in HBM;
`e2e` = the same through flipb200_step_host with pinned HOST buffers (H2D + step + D2H per step);
`roofline` = the PCG stencil kernel (17 algorithmic bytes per swept node) against the measured HBM
peak; `cpu_baseline` = the CPU oracle on a bounded sample.

Other workloads (BASELINE configs[3] and [4]; parity-test configurations, not the headline line):
  --workload cube-drop --grid-n 512   cube of side n/3 (8 ppc) dropped in an n^3 box cut into N z-slabs
                                      of n/N planes -- STRONG scaling ("scaling": "strong")
  --workload pcg                      pressure projection only: the 7-point MG-PCG on the full-tank system
                                      (b = divergence after one gravity kick), a "step" = 200 iterations
                                      without convergence exit; metric pcg_iterations_per_s, weak scaling

N > 1 (torchrun, one rank per GPU): ONE tank of n x n x (n*N) nodes cut into N z-slabs (SURVEY.md
8e) -- weak scaling, every GPU owns an n^3 block and ~the same particles; ghost planes, PCG dot
products and particle migration go over the library's own NCCL communicator.  `value` counts the
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
CPU_SAMPLE_N = 64


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


def scene_box(n, nz=None):
    nz = n if nz is None else nz
    wall = 2
    lo = (wall, wall, wall)
    hi = (n - 1 - wall, n - 1 - wall - 2, nz - 1 - wall)     # full tank with a 2-cell air gap (SURVEY 8d fill (i))
    return lo, hi


def run_cpu_sample(steps, warmup, threads):
    """The CPU arm: oracle restatement + the reference's Eigen IC-PCG on the same scene family at
    n = CPU_SAMPLE_N (bounded sample).  Returns (particle-steps/s, description)."""
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
    kind = "port"   # restated particle/grid loops; the pressure solve IS the reference's own Eigen code
    desc = (f"dam-break full-tank fill at {n}^3 ({sc.num_particles} particles), {steps} full steps after "
            f"{warmup} warm-up: oracle restatement of fluid.cc with the reference's vendored Eigen IC-PCG "
            f"(single-threaded, as run.sh builds it) and OpenMP particle loops on {threads} threads; "
            f"{passes} relaxation passes, {iters} PCG iterations; generous to the CPU: IC-PCG iterations grow "
            f"~linearly with n and the restatement drops the reference's ValueAccessor/mutex overhead")
    return sc.num_particles * steps / el, el / steps * 1e3, kind, desc, o.has_eigen


PCG_BLOCK = 200      # iterations per "step" of the pcg workload


def bench_pcg(args, f, fb, config, rank, world, local_rank, barrier, max_over_ranks, n, nzg):
    """BASELINE configs[4]: pressure-projection-only weak scaling.  The system is the one the first
    relaxation pass of the full-tank scene solves (b = divergence after one gravity kick, non-zero in
    the layer next to the floor); a step = PCG_BLOCK iterations of the product's CUDA-graph iteration
    without convergence exit (flipb200_time_kernel, CUDA events on the library's stream, max over
    ranks); the run-to-tolerance iteration count of the same system is reported next to it."""
    import torch.distributed as dist
    dt = 0.1
    f.p2g()
    n_fluid = f.classify()
    f.rhs_div(dt)
    iters_tol, rel = f.pcg_solve(dt)                       # run to 2.2e-16 (also the warm-up of the iteration graph)
    for _ in range(args.warmup):
        f.time_kernel(fb.KERNEL_PCG_ITER, 20)
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0 and not os.environ.get("FLIPB200_BENCH_NOSMI"):
        sampler.start()
    ms_block = 0.0
    launches = 0
    nodes = 0
    for _ in range(args.steps):
        ms_iter, nodes = f.time_kernel(fb.KERNEL_PCG_ITER, PCG_BLOCK)
        ms_block += ms_iter * PCG_BLOCK
    barrier()
    clocks = sampler.stop()
    info = f.debug_info()
    launches = info["kernels_per_iteration"] * PCG_BLOCK * args.steps
    ms_block = max_over_ranks(ms_block)
    it_s = PCG_BLOCK * args.steps / (ms_block * 1e-3)
    peak, peak_src = measured_peak_gbs()
    ms_spmv, _ = f.time_kernel(fb.KERNEL_SPMV, args.kernel_iters)
    ach = 17.0 * nodes / (ms_spmv * 1e-3) / 1e9
    config = dict(config)
    config["workload"] = (f"pressure projection only: 7-point {'MG-' if args.precond == 1 else 'Jacobi-'}PCG (fp64) on the full-tank system, "
                          f"{n}x{n}x{nzg} grid = {n}^3 per GPU, {n_fluid} unknowns, b = divergence after one gravity kick; "
                          f"step = {PCG_BLOCK} iterations without convergence exit (BASELINE configs[4])")
    config["engine"] = info
    config["l2"] = "vectors larger than L2 (8 B x n^3 each)" if n >= 256 else "working set may fit L2 at this n"
    if rank == 0:
        line = {"metric": "pcg_iterations_per_s", "value": it_s, "unit": "iterations/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_block / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config, "clocks": clocks,
                "e2e": None, "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "kernel": "k_spmv (7-pt matrix-free A.p + p.Ap)", "achieved": ach, "peak": peak, "unit": "GB/s",
                             "frac": ach / peak, "traffic": None, "algorithmic_bytes": 17 * nodes, "peak_source": peak_src,
                             "swept_nodes_per_gpu": nodes, "us_per_launch": ms_spmv * 1e3},
                "cpu_baseline": None,
                "pcg": {"us_per_iteration": ms_block / (PCG_BLOCK * args.steps) * 1e3, "iterations_to_tolerance": iters_tol,
                        "rel_residual": rel, "unknowns": n_fluid, "swept_nodes_per_gpu": nodes,
                        "node_updates_per_s": it_s * nodes * world,
                        "jacobi_equivalent_frac": 91.0 * nodes * it_s / 1e9 / peak}}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="flipb200", choices=["flipb200", "reference"])
    ap.add_argument("--grid-n", dest="n", type=int, default=int(os.environ.get("FLIPB200_BENCH_N", "256")))
    ap.add_argument("--workload", default="flip", choices=["flip", "cube-drop", "pcg"],
                    help="flip: dam-break full tank, weak scaling (headline); cube-drop: n^3 box, strong scaling; pcg: pressure solve only")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--kernel-iters", type=int, default=50)
    ap.add_argument("--precond", type=int, default=int(os.environ.get("FLIPB200_PRECOND", "1")), help="0 Jacobi, 1 multigrid V-cycle")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    n = args.n
    if args.workload == "cube-drop":
        # BASELINE configs[3]: the reference's cube drop (41/121 of the box, fluid.cc:1176-1216) at n^3, cut into z-slabs
        nzg = n
        side = n // 3
        lo = tuple((n - side) // 2 for _ in range(3))
        hi = tuple(l + side - 1 for l in lo)
        total_particles = side ** 3 * 8
        from importlib import import_module
        zs, _a = import_module("fluid_simulation_b200.dist").slab_split(nzg, world)
        own = max(0, min(hi[2] + 1, zs[rank + 1]) - max(lo[2], zs[rank]))
        particles = side * side * own * 8              # this rank's seeds
        cap = int((particles + side * side * 8 * 2) * 1.25) + 4096
        config = {"workload": f"cube-drop FLIP, {n}^3 grid, cube of side {side} voxels, 8 particles/cell ({total_particles} particles), "
                              f"fp64, {'MG-' if args.precond == 1 else 'Jacobi-'}PCG to 2.2e-16 (BASELINE configs[3])",
                  "grid": [n, n, nzg], "particles_total": total_particles,
                  "parallelism": "single GPU" if world == 1 else f"{world} z-slabs of ~{n // world} planes, NCCL halo exchange + PCG all-reduces + particle migration",
                  "l2": "grid vectors (8 B x n^3) larger than L2; no flush between steps"}
        scaling = "strong"
    else:
        nzg = n * world                                    # weak scaling: an n^3 block per GPU, stacked along z
        lo, hi = scene_box(n, nz=nzg)
        total_particles = (hi[0] - lo[0] + 1) * (hi[1] - lo[1] + 1) * (hi[2] - lo[2] + 1) * 8
        particles = total_particles // world
        cap = int(particles * 1.06) + 4096
        config = {"workload": f"dam-break FLIP full-tank fill, {n}x{n}x{nzg} grid, 8 particles/cell ({total_particles} particles, "
                              f"~{particles}/GPU), fp64, {'MG-' if args.precond == 1 else 'Jacobi-'}PCG to 2.2e-16 "
                              f"(BASELINE configs[{2 if n == 256 else 1 if n == 128 else '-'}]" + (" per GPU" if world > 1 else "") + ")",
                  "grid": [n, n, nzg], "particles_per_gpu": particles, "particles_total": total_particles,
                  "parallelism": "single GPU" if world == 1 else f"{world} z-slabs of {n} planes, NCCL halo exchange + PCG all-reduces + particle migration",
                  "l2": "inputs larger than L2 (particle arrays 48 B x particles, grid vectors 8 B x n^3)" if n >= 256
                        else "working set may fit L2 at this n; no flush between steps"}
        scaling = "weak"

    if args.impl == "reference":
        # Reference arm: the reference's CPU implementation of the path, host cores only.
        if rank != 0:
            return 0
        threads = os.cpu_count() or 1
        v, ms, kind, desc, has_eigen = run_cpu_sample(args.steps, max(args.warmup, 0), threads)
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": scaling,
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "sample": desc,
                                 "eigen_from_reference": bool(has_eigen)},
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return 0

    import torch
    import torch.distributed as dist
    import fluid_simulation_b200 as fb
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the FLIP path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from fluid_simulation_b200 import dist as fdist

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    f = fdist.create_slab_flip(fb, rank, world, device=local_rank, nx=n, ny=n, nz=nzg, ox=n // 2, oy=n // 2, oz=nzg // 2,
                               max_particles=cap, precond=args.precond)
    f.seed_box(lo, hi, seed=1234)
    if world == 1:
        assert f.num_particles == particles
    if args.workload == "pcg":
        return bench_pcg(args, f, fb, config, rank, world, local_rank, barrier, max_over_ranks, n, nzg)
    dt = 0.1
    stats = []
    for _ in range(args.warmup):
        dt, st = f.step(dt)
    # one sampler for the job (rank 0, its own GPU): concurrent nvidia-smi polls from every rank
    # serialise on the driver and perturb the step they are meant to observe
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0 and not os.environ.get("FLIPB200_BENCH_NOSMI"):
        sampler.start()
    t0 = time.perf_counter()
    dev_ms = 0.0
    launches = 0
    for _ in range(args.steps):
        dt, st = f.step(dt)
        dev_ms += st.ms_total; launches += st.launches
        stats.append(st.as_dict())
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    # device time: CUDA events on the library's stream (flipb200_stats.ms_total), max over ranks
    dev_s = max_over_ranks(dev_ms * 1e-3)
    wall = max_over_ranks(wall)
    value = total_particles * args.steps / dev_s

    config["engine"] = f.debug_info()
    # ---- kernel roofline (dominant kernel: the PCG stencil) ----
    peak, peak_src = measured_peak_gbs()
    f.p2g(); f.classify(); f.rhs_div(dt)
    ms_spmv, nodes = f.time_kernel(fb.KERNEL_SPMV, args.kernel_iters)
    ms_iter, _ = f.time_kernel(fb.KERNEL_PCG_ITER, args.kernel_iters)
    ms_p2g, npart = f.time_kernel(fb.KERNEL_P2G, 5)
    ms_sort, _ = f.time_kernel(fb.KERNEL_SORT, 5)
    ach = 17.0 * nodes / (ms_spmv * 1e-3) / 1e9
    traffic = None          # dram__bytes_read.sum + dram__bytes_write.sum of one launch, from the committed ncu capture
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["k_spmv"]
        if tj["grid"] == [n, n, n]:
            traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "k_spmv (7-pt matrix-free A.p + p.Ap)", "achieved": ach, "peak": peak,
                "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "traffic_source": "profiles/traffic.json (ncu, per launch)" if traffic else None,
                "algorithmic_bytes": 17 * nodes, "peak_source": peak_src,
                "algorithmic_bytes_per_node": 17, "swept_nodes": nodes, "us_per_launch": ms_spmv * 1e3}
    kernels = {
        "pcg_iteration": {"bytes_per_node": 91, "us": ms_iter * 1e3, "gbs": 91.0 * nodes / (ms_iter * 1e-3) / 1e9,
                          "frac": 91.0 * nodes / (ms_iter * 1e-3) / 1e9 / peak, "iters_per_s": 1e3 / ms_iter},
        "p2g": {"bytes_per_particle": 52, "ms": ms_p2g, "gbs": 52.0 * npart / (ms_p2g * 1e-3) / 1e9,
                "frac": 52.0 * npart / (ms_p2g * 1e-3) / 1e9 / peak},
        "sort": {"ms": ms_sort},
    }
    # G2P + advect and the whole particle phase (the phase-level fraction is reported next to
    # the kernel-level ones), from the device-timed phases of the K timed steps of this rank
    ms_g2p_step = float(np.mean([s["ms_g2p"] for s in stats]))
    ms_part_step = float(np.mean([s["ms_sort"] + s["ms_p2g"] + s["ms_g2p"] for s in stats]))
    if ms_g2p_step > 0.0 and ms_part_step > 0.0:
        kernels["g2p_advect"] = {"bytes_per_particle": 99.1, "ms": ms_g2p_step,
                                 "gbs": 99.1 * npart / (ms_g2p_step * 1e-3) / 1e9,
                                 "frac": 99.1 * npart / (ms_g2p_step * 1e-3) / 1e9 / peak,
                                 "source": "flipb200_stats.ms_g2p (k_cell_delta + k_g2p_cell + k_advect), mean of the timed steps"}
        kernels["particle_phase"] = {"bytes_per_particle": 151.1, "ms": ms_part_step,
                                     "gbs": 151.1 * npart / (ms_part_step * 1e-3) / 1e9,
                                     "frac": 151.1 * npart / (ms_part_step * 1e-3) / 1e9 / peak,
                                     "source": "(52 + 99.1) B x particles / (ms_sort + ms_p2g + ms_g2p): sort traffic is overhead, not algorithmic"}

    # ---- e2e: the same step through the host-buffer entry point ----
    e2e = None
    if not args.no_e2e:
        import ctypes as C
        nloc = f.num_particles
        cap = nloc + nloc // 16 + 4096 if world > 1 else nloc
        arrs, ptrs = [], []
        for q in range(6):
            a, p = fb.pinned_empty(cap)
            arrs.append(a); ptrs.append(p)
        m = C.c_int64(0)
        fb._check(f.lib.flipb200_get_particles(f.h, *[fb._ptr(a) for a in arrs], cap, C.byref(m)), "get")
        out, outp = fb.pinned_empty(f.N, np.float32)          # this rank's planes of outputGrid
        dte = dt

        def host_step(nl, d):
            if world == 1:
                d, st_ = f.step_host(arrs, d, output_grid=out)
                return nl, d
            d, st_, nl = f.step_host_slab(arrs, nl, d, output_grid=out)
            return nl, d

        nloc, dte = host_step(nloc, dte)                       # warm-up
        barrier()
        t0 = time.perf_counter()
        moved = 0
        for _ in range(args.steps):
            moved += nloc
            nloc, dte = host_step(nloc, dte)
        barrier()
        el = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": total_particles * args.steps / el, "unit": UNIT,
               "h2d_bytes_per_step": 48 * total_particles, "d2h_bytes_per_step": 48 * total_particles + 4 * n * n * nzg,
               "ms_per_step": el / args.steps * 1e3,
               "api": ("flipb200_step_host" if world == 1 else "flipb200_step_host_slab per rank") +
                      " (pinned host SoA in/out + outputGrid)"}
        for p in ptrs + [outp]:
            fb.pinned_free(p)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        v, ms, kind, desc, has_eigen = run_cpu_sample(1, 1, threads)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "sample": desc,
               "eigen_from_reference": bool(has_eigen)}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_s / args.steps * 1e3, "higher_is_better": True,
                "scaling": scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
                "cpu_baseline": cpu, "kernels": kernels, "wall_ms_per_step": wall / args.steps * 1e3,
                "per_step": [{k: s[k] for k in ("passes", "pcg_iters_total", "ms_sort", "ms_p2g", "ms_grid", "ms_pcg",
                                                 "ms_g2p", "ms_total")} for s in stats]}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
