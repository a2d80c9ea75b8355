#!/usr/bin/env python
"""Generate tests/golden/*.npz with oracle/_ref -- the CPU restatement of fluid.cc linked against the
REFERENCE'S OWN Eigen 3.3.4 IC-PCG (compiled from /root/reference/Eigen; only available in the
build container, not on the GPU box).  Each file holds the seeded inputs and what the oracle
produced, so that (a) the oracle is regression-pinned on CPU (tests/test_golden.py) and (b) the CUDA
path is checked against vectors that were produced with the reference's solver even where
/root/reference does not exist.  Run from the repo root:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as O                       # noqa: E402
import fluid_simulation_b200  # noqa: E402,F401
from fluid_simulation_b200 import scenes    # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def run(name, sc, nsteps, **kw):
    """kw: optional oracle parameters stored with the vectors (solid, extrapolate, flip_ratio)."""
    nx, ny, nz = sc.n
    ox, oy, oz = sc.origin
    o = O.Oracle(nx=nx, ny=ny, nz=nz, ox=ox, oy=oy, oz=oz, wall=sc.wall, **kw)
    assert o.has_eigen, "goldens must come from oracle/_ref (reference Eigen), not the port"
    o.set_particles(sc.pos, sc.vel)
    out = dict(n=np.array(sc.n), origin=np.array(sc.origin), wall=np.array(sc.wall),
               pos0=sc.pos.astype(np.float64), vel0=sc.vel.astype(np.float64))
    for k, v in kw.items():
        out["opt_" + k] = np.asarray(v)
    dt = 0.1
    passes, nfluid, dts, errs = [], [], [], []
    for s in range(nsteps):
        dt, st = o.step(dt)
        passes.append(st.passes); nfluid.append(st.n_fluid); dts.append(dt); errs.append(st.last_error)
        if s == 0:   # single-step grid state (north_star: velocity / pressure rel-L2 <= 1e-5)
            out["u1"] = np.stack([o.dgrid(O.D_U + a) for a in range(3)]).astype(np.float64)
            out["p1"] = o.dgrid(O.D_P).astype(np.float64)
            out["w1"] = o.fgrid(O.F_W).astype(np.float32)
            out["fluid1"] = (o.indices >= 0)
    p, v = o.get_particles()
    out.update(passes=np.array(passes), n_fluid=np.array(nfluid), dt=np.array(dts), error=np.array(errs),
               pos=p, vel=v)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "particles", sc.num_particles, "passes", passes, "n_fluid", nfluid[0])


if __name__ == "__main__":
    O.build()
    run("cube_drop_25", scenes.small_cube_drop(n=25, half=4), 10)           # miniature of fluid.cc's default scene
    sc = scenes.dam_break(n=24, fill="column")
    rng = np.random.default_rng(11)
    sc.vel = rng.normal(0.0, 0.5, size=sc.pos.shape)
    run("dam_break_column_24", sc, 10)
    # the optional paths in one scene: interior obstacle (fluid.cc:1303-1345), extrapolate (:705-802),
    # PIC/FLIP blend 0.9 (:904-971, :125-207)
    sc = scenes.small_cube_drop(n=25, half=5, seed=3)
    sc.vel = np.random.default_rng(103).normal(0.0, 1.0, size=sc.pos.shape)
    solid = np.zeros((25, 25, 25), dtype=np.uint8)
    solid[10:15, 4:8, 10:15] = 1
    run("obstacle_extrapolate_pic_25", sc, 5, solid=solid, extrapolate=1, flip_ratio=0.9)
