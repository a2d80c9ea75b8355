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
        assert BASE_KEYS | {"clocks", "roofline"} <= set(d), sorted((BASE_KEYS | {"clocks", "roofline"}) - set(d))
        assert d["dtype"] == "f64" and d["data"] == "synthetic" and d["vs_baseline"] is None
        assert "workload" in d["config"] and "model" not in d["config"]
        rf = d["roofline"]
        assert rf["bound"] == "hbm" and rf["unit"] == "GB/s" and abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-12
        assert d["gpu_launches"] > 0
    assert flip["metric"] == cube["metric"] == "flip_particle_steps_per_s"
    assert flip["scaling"] == "weak" and cube["scaling"] == "strong" and pcg["scaling"] == "weak"
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(flip["e2e"])
    assert flip["e2e"]["h2d_bytes_per_step"] == 48 * flip["config"]["particles_total"]
    assert {"pcg_iteration", "p2g", "sort", "g2p_advect", "particle_phase"} <= set(flip["kernels"])
    assert "configs[2]" in flip["config"]["workload"] and "configs[3]" in cube["config"]["workload"]
    assert pcg["metric"] == "pcg_iterations_per_s" and "configs[4]" in pcg["config"]["workload"]


def test_reference_arm_line(oracle_mod):
    env = dict(os.environ, OMP_NUM_THREADS="4")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert BASE_KEYS <= set(d)
    assert d["impl"] == "reference" and d["metric"] == "flip_particle_steps_per_s" and d["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["gpu_launches"] == 0
