"""Dry-run of bench.py's GPU arm with a FAKE fluid_simulation_b200 binding and a patched torch.cuda
(no GPU on the build machine): exercises the Python control flow and the JSON contract only.  Run as a
script in its own process by tests/test_bench_contract.py; prints one JSON line per workload."""
import contextlib
import io
import json
import os
import runpy
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["FLIPB200_BENCH_NOSMI"] = "1"
import torch
torch.cuda.is_available = lambda: True
torch.cuda.set_device = lambda d: None
torch.cuda.synchronize = lambda: None


class St:
    def __init__(s, n_fluid):
        s.ms_total = 40.0; s.launches = 1500; s.passes = 2; s.pcg_iters_total = 42; s.n_fluid = n_fluid

    def as_dict(s):
        return dict(passes=2, pcg_iters_total=42, pcg_iters_last=21, n_fluid=s.n_fluid, ms_sort=3.0, ms_p2g=2.5, ms_grid=2.0, ms_pcg=28.0,
                    ms_g2p=4.5, ms_total=40.0, last_error=0.05, dt_out=0.1, max_speed=1.0, pcg_rel_resid=1e-16)


class Lib:
    def flipb200_get_particles(self, h, *a):
        return 0


class Flip:
    def __init__(s, rank=0, world=1, **kw):
        s.kw = kw; s.n = 0; s.lib = Lib(); s.h = None
        s.nx, s.ny, s.nz = kw['nx'], kw['ny'], kw['nz']
        s.z0, s.z1 = 0, s.nz
        s.N = s.nx * s.ny * s.nz // world
        s.nf = 0
        s.params = types.SimpleNamespace(mixed=3)

    def seed_box(s, lo, hi, seed=0):
        s.n = (hi[0] - lo[0] + 1) * (hi[1] - lo[1] + 1) * (hi[2] - lo[2] + 1) * 8
        if tuple(lo) == (2, 2, 2) and hi[0] == s.nx - 3:
            s.nf = (s.nx - 4) * (s.ny - 5) * (s.nz - 4)
        else:
            s.nf = (hi[0] - lo[0] + 3) * (hi[1] - lo[1] + 3) * (hi[2] - lo[2] + 3)

    @property
    def num_particles(s): return s.n
    def step(s, dt): return dt, St(s.nf)
    def step_host(s, arrs, d, output_grid=None): return d, St(s.nf)
    def debug_info(s): return {"kernels_per_iteration": 30}
    def p2g(s): pass
    def classify(s): return s.nf
    def rhs_div(s, dt): return 1.0
    def pcg_solve(s, dt): return 21, 1e-16
    def close(s): pass

    def get_grid(s, which):
        top = s.ny - 3
        p = np.zeros((s.nz, s.ny, s.nx))
        p[:, 2:top, :] = (10.0 * (top - np.arange(2, top)))[None, :, None]
        return p

    def profile_iteration(s, iters=10):
        return [("k_spmv", 60.0), ("k_update_xr", 70.0), ("L0 halo+smooth", 43.0), ("L0 halo+resid", 50.0), ("L0 halo+restrict", 28.0),
                ("L1 halo+smooth", 9.0), ("L2+ halo+smooth", 6.0), ("L2+ halo+smooth", 6.0), ("tail (small levels + coarsest)", 36.0),
                ("L0 halo+prolong", 38.0), ("L0 halo+smooth (post 1)", 43.0), ("L0 halo+smooth (post 2)", 60.0), ("k_update_p", 55.0)]

    def time_kernel(s, which, iters): return (0.06, 16_000_000) if which != 2 else (2.4, s.n)


fb = types.ModuleType('fluid_simulation_b200')
fb.__path__ = [os.path.join(ROOT, 'fluid-simulation_b200')]
fb.Flip = Flip
fb.KERNEL_SPMV, fb.KERNEL_PCG_ITER, fb.KERNEL_P2G, fb.KERNEL_G2P, fb.KERNEL_SORT = 0, 1, 2, 3, 4
fb.GRID_P = 6
fb.GRID_U = 0
fb.comm_unique_id = lambda: b'0' * 128
fb.pinned_empty = lambda n, dtype=np.float64: (np.empty(8, dtype=dtype), 0)
fb.pinned_free = lambda p: None
fb._check = lambda rc, where: None
fb._ptr = lambda a: None
sys.modules['fluid_simulation_b200'] = fb

for argv in (["bench.py", "--steps", "2", "--warmup", "1", "--no-cpu", "--grid-n", "64"],
             ["bench.py", "--steps", "2", "--warmup", "4", "--no-cpu", "--workload", "cube-drop", "--grid-n", "512"],
             ["bench.py", "--steps", "2", "--warmup", "1", "--workload", "pcg", "--grid-n", "64"]):
    sys.argv = argv
    buf = io.StringIO()
    try:
        with contextlib.redirect_stdout(buf):
            runpy.run_path(os.path.join(ROOT, 'bench.py'), run_name='__main__')
    except SystemExit as e:
        assert e.code == 0, e.code
    d = json.loads(buf.getvalue().strip().splitlines()[-1])
    print(json.dumps(d))
