"""bench.py's JSON contract, checked on the build machine (no GPU): the reference arm runs for real
on a tiny budget; the CUDA arm's control flow runs against a fake binding (tests/bench_dryrun.py) so
that a typo in bench.py cannot first show up in the round-end run on the GPU box."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "cpu_baseline"}


def test_cuda_arm_control_flow_and_keys():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "bench_dryrun.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [json.loads(l) for l in r.stdout.strip().splitlines()]
    assert len(lines) == 3
    flip, cube, pcg = lines
    for d in lines:
        assert BASE_KEYS | {"check", "checks_ok"} <= set(d), sorted((BASE_KEYS | {"check", "checks_ok"}) - set(d))
        assert d["dtype"] == "f64" and d["data"] == "synthetic" and d["vs_baseline"] is None
        assert "workload" in d["config"] and "model" not in d["config"]
        assert d["gpu_launches"] > 0 and d["checks_ok"] is True and d["check"]["ok"] is True
    rf = flip["roofline"]
    assert rf["bound"] == "hbm" and rf["unit"] == "GB/s" and abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-12
    assert {"p2g", "g2p_advect", "sort", "particle_phase", "pcg_iteration", "pcg/k_spmv", "pcg/L0 halo+smooth", "pcg/k_update_p"} <= set(rf["by_kernel"])
    assert rf["by_kernel"]["pcg/L0 halo+smooth"]["bytes_per_node"] == 13.0 and 0 < rf["by_kernel"]["p2g"]["frac"] < 1
    assert rf["dominant"]["kernel"] and 0 < rf["dominant"]["share_of_step"] <= 1
    assert "clocks" in flip
    assert flip["metric"] == cube["metric"] == "flip_particle_steps_per_s"
    assert flip["scaling"] == "weak" and cube["scaling"] == "strong" and pcg["scaling"] == "weak"
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(flip["e2e"])
    assert flip["e2e"]["h2d_bytes_per_step"] == 48 * flip["config"]["particles_total"]
    assert flip["check"]["hydrostatic"]["ok"] and flip["check"]["steps"]["ok"]
    assert flip["check"]["hydrostatic"]["n_fluid"] == flip["check"]["hydrostatic"]["n_fluid_closed_form"]
    # the other BASELINE configs ride on the headline line
    assert flip["pcg"]["metric"] == "pcg_iterations_per_s" and "configs[4]" in flip["pcg"]["config"]["workload"]
    assert "fill (ii)" in flip["dam_break_column"]["config"]["workload"] and flip["dam_break_column"]["check"]["ok"]
    assert "configs[1]" in flip["grid_128"]["workload"] and "configs[0]" in flip["default_scene"]["workload"]
    assert "precision = 32" in flip["fp32_mode"]["workload"] and flip["fp32_mode"]["vs_precision_64_after_one_step"]["grid_velocity_rel_l2"] >= 0
    assert "configs[3]" in cube["config"]["workload"] and cube["check"]["first_step_cells"]["ok"]
    assert pcg["metric"] == "pcg_iterations_per_s" and "configs[4]" in pcg["config"]["workload"] and pcg["check"]["hydrostatic"]["ok"]


def test_reference_arm_line(oracle_mod):
    env = dict(os.environ, OMP_NUM_THREADS="4")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--no-extra"],
                       capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert BASE_KEYS <= set(d)
    assert d["impl"] == "reference" and d["metric"] == "flip_particle_steps_per_s" and d["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["gpu_launches"] == 0
    # the arm says what it ran: a bounded sample, not the 256^3 headline configuration
    assert d["config"]["grid"] == [48, 48, 48] and "BOUNDED SAMPLE" in d["config"]["workload"] and cb["grid_n"] == 48
