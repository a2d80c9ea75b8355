"""Golden vectors (tests/golden/*.npz, produced by tests/golden/make_golden.py with oracle/_ref, i.e.
the restatement of fluid.cc linked against the reference's own Eigen IC-PCG).

CPU: whichever oracle library is present here (the _ref build in the container, possibly only the
port elsewhere) must reproduce them -- pass counts and cell classes exactly, floating point to
round-off (the port replaces Eigen's IC-PCG by its own PCG, so it is held to the north_star bounds).
GPU: the CUDA path through the C ABI against the same vectors, north_star tolerances."""
import glob
import os

import numpy as np
import pytest

from helpers import rel_l2

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = sorted(glob.glob(os.path.join(HERE, "golden", "*.npz")))
TOL_GRID = 1e-5      # north_star: single-step grid velocity / pressure rel-L2
TOL_POS10 = 1e-4     # north_star: positions after 10 steps rel-L2


def test_golden_files_present():
    assert len(GOLDEN) >= 3


def _opts(g):
    """Optional parameters stored with a golden file (solid, extrapolate, flip_ratio)."""
    kw = {}
    for k in g.files:
        if k.startswith("opt_"):
            v = g[k]
            kw[k[4:]] = v if v.ndim else v.item()
    return kw


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_oracle_reproduces_golden(oracle_mod, path):
    g = np.load(path)
    nx, ny, nz = (int(v) for v in g["n"]); ox, oy, oz = (int(v) for v in g["origin"])
    o = oracle_mod.Oracle(nx=nx, ny=ny, nz=nz, ox=ox, oy=oy, oz=oz, wall=int(g["wall"]), **_opts(g))
    o.set_particles(g["pos0"], g["vel0"])
    exact = bool(o.has_eigen)
    dt = 0.1
    for s in range(len(g["passes"])):
        dt, st = o.step(dt)
        assert st.passes == g["passes"][s] and st.n_fluid == g["n_fluid"][s], s
        assert abs(dt - g["dt"][s]) <= 1e-12 * g["dt"][s]
        if s == 0:
            u = np.stack([o.dgrid(oracle_mod.D_U + a) for a in range(3)])
            assert np.array_equal(o.indices >= 0, g["fluid1"])
            assert rel_l2(u, g["u1"]) <= (1e-13 if exact else TOL_GRID)
            assert rel_l2(o.dgrid(oracle_mod.D_P), g["p1"]) <= (1e-12 if exact else TOL_GRID)
    p, v = o.get_particles()
    assert rel_l2(p, g["pos"]) <= (1e-13 if exact else TOL_POS10)


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_cuda_path_matches_golden(fb, path):
    g = np.load(path)
    nx, ny, nz = (int(v) for v in g["n"]); ox, oy, oz = (int(v) for v in g["origin"])
    f = fb.Flip(nx=nx, ny=ny, nz=nz, ox=ox, oy=oy, oz=oz, wall=int(g["wall"]), track_ids=1, **_opts(g))
    f.set_particles(g["pos0"], g["vel0"])
    dt = 0.1
    for s in range(len(g["passes"])):
        dt, st = f.step(dt)
        assert st.passes == g["passes"][s] and st.n_fluid == g["n_fluid"][s], (s, st.passes, g["passes"][s])
        assert abs(dt - g["dt"][s]) <= 1e-9 * g["dt"][s]
        if s == 0:
            u = np.stack([f.get_grid(fb.GRID_U + a) for a in range(3)])
            fl = f.get_grid(fb.GRID_FLAGS)
            assert np.array_equal((fl & fb.FB_FLUID) != 0, g["fluid1"])           # exact cell classification
            assert rel_l2(u, g["u1"]) <= TOL_GRID
            assert rel_l2(f.get_grid(fb.GRID_P), g["p1"]) <= TOL_GRID
            assert rel_l2(f.get_grid(fb.GRID_WEIGHT), g["w1"]) <= 5e-7
    p, v = f.get_particles()
    assert rel_l2(p, g["pos"]) <= TOL_POS10 and rel_l2(v, g["vel"]) <= 1e-3
    f.close()
