"""z-slab decomposition on ONE GPU: the ranks run as host threads of this process
over the in-process transport (flipb200_comm_init_local), so ghost planes, the P2G ghost sum, the
distributed multigrid V-cycle, the all-reduced dot products and particle migration are all checked
against the single-context run of the same scene.  The NCCL transport swaps only the copies
(fluid-simulation_b200/csrc/comm.cu); it is exercised by tests/test_gpu_nccl.py on >= 2 GPUs.

Bars: integer results (fluid-cell count, relaxation passes, particle count, flags) identical;
floating point within the north_star tolerances against the single-GPU run, which itself is
parity-tested against the oracle in test_gpu_parity.py (the split changes only summation order)."""
import itertools
import os
import threading

import numpy as np
import pytest

from helpers import rel_l2

pytestmark = pytest.mark.gpu

_group = itertools.count(100)


@pytest.fixture(scope="module")
def scenes():
    from fluid_simulation_b200 import scenes as s
    return s


def run_ranks(fns):
    """Run one callable per rank concurrently; re-raise the first exception."""
    out = [None] * len(fns)
    err = []

    def wrap(i):
        try:
            out[i] = fns[i]()
        except BaseException as e:      # noqa: BLE001
            err.append(e)

    th = [threading.Thread(target=wrap, args=(i,)) for i in range(len(fns))]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=300)
    if err:
        raise err[0]
    assert all(not t.is_alive() for t in th), "a rank hung"
    return out


def make_slabs(fb, sc, world, **kw):
    nx, ny, nz = sc.n
    ox, oy, oz = sc.origin
    gid = next(_group)
    fs = [fb.Flip(nx=nx, ny=ny, nz=nz, ox=ox, oy=oy, oz=oz, wall=sc.wall, track_ids=1, rank=r, world=world, **kw)
          for r in range(world)]
    for f in fs:
        f.comm_init_local(gid)
    run_ranks([lambda f=f: f.set_particles(sc.pos, sc.vel) for f in fs])
    return fs


def gather_grid(fs, which):
    return np.concatenate([f.get_grid(which) for f in fs], axis=0)


def gather_particles(fs, n):
    pos = np.full((n, 3), np.nan); vel = np.full((n, 3), np.nan)
    seen = 0
    for f in fs:
        p, v = f.get_particles()
        ids = f.get_particle_ids()
        pos[ids] = p; vel[ids] = v
        seen += len(ids)
    assert seen == n and not np.isnan(pos).any(), "particles lost or duplicated by the migration"
    return pos, vel


def moving_scene(scenes, n=32, seed=5):
    """Column fill with a velocity field that pushes particles across the slab boundaries."""
    sc = scenes.dam_break(n=n, fill="column")
    rng = np.random.default_rng(seed)
    sc.vel = rng.normal(0.0, 1.0, size=sc.pos.shape)
    sc.vel[:, 2] += 3.0 * np.sin(sc.pos[:, 0] * 0.3)
    return sc


@pytest.mark.parametrize("world,precond,fused", [(2, 1, 1), (4, 1, 1), (2, 1, 0), (2, 0, 1)])
def test_slab_steps_match_single(fb, scenes, world, precond, fused, monkeypatch):
    """fused=1: ghost planes are pushed by the producing kernels (halo.cuh; here ordered by the host
    barrier of the in-process transport); fused=0: separate halo copies."""
    monkeypatch.setenv("FLIPB200_FUSED_HALO", str(fused))
    sc = moving_scene(scenes, n=32)
    nsteps = 6
    f1 = fb.Flip.from_scene(sc, track_ids=1, precond=precond)
    fs = make_slabs(fb, sc, world, precond=precond)
    assert sum(f.num_particles for f in fs) == sc.num_particles
    owner0 = np.empty(sc.num_particles, dtype=np.int64)
    for r, f in enumerate(fs):
        owner0[f.get_particle_ids()] = r
    assert [f.z0 for f in fs] == sorted(f.z0 for f in fs) and fs[0].z0 == 0 and fs[-1].z1 == 32
    dt1 = 0.1
    dts = [0.1] * world
    for step in range(nsteps):
        dt1, st1 = f1.step(dt1)
        res = run_ranks([lambda f=f, d=dts[r]: f.step(d) for r, f in enumerate(fs)])
        dts = [r[0] for r in res]
        sts = [r[1] for r in res]
        assert all(s.passes == st1.passes for s in sts), (step, st1.passes, [s.passes for s in sts])
        assert all(s.n_fluid == st1.n_fluid for s in sts)
        assert all(abs(d - dt1) <= 1e-12 * dt1 for d in dts)
        # grid state after the step
        fl1 = f1.get_grid(fb.GRID_FLAGS); fls = gather_grid(fs, fb.GRID_FLAGS)
        assert np.array_equal(fl1 & (fb.FB_FLUID | fb.FB_SOLID), fls & (fb.FB_FLUID | fb.FB_SOLID))
        for a in range(3):
            assert rel_l2(gather_grid(fs, fb.GRID_U + a), f1.get_grid(fb.GRID_U + a)) < 1e-9, (step, a)
        assert rel_l2(gather_grid(fs, fb.GRID_P), f1.get_grid(fb.GRID_P)) < 1e-8, step
    assert all(f.halo_mode == (1 if (fused and precond == 1) else 0) for f in fs)
    p1, v1 = f1.get_particles()
    ps, vs = gather_particles(fs, sc.num_particles)
    assert rel_l2(ps, p1) < 1e-10 and rel_l2(vs, v1) < 1e-9
    # the scene must actually have exercised migration
    owner1 = np.empty(sc.num_particles, dtype=np.int64)
    for r, f in enumerate(fs):
        owner1[f.get_particle_ids()] = r
    assert np.count_nonzero(owner0 != owner1) > 0
    for f in fs + [f1]:
        f.close()


@pytest.mark.parametrize("nd", [1, 2])
def test_slab_multigrid_levels(fb, scenes, nd):
    """Force 1 and 2 distributed multigrid levels (the rest replicated): same iteration count as the
    single-GPU V-cycle up to +-1 and the same pressure."""
    sc = scenes.dam_break(n=48, fill="full")
    os.environ["FLIPB200_MG_ND"] = str(nd)
    try:
        fs = make_slabs(fb, sc, 2)
    finally:
        os.environ.pop("FLIPB200_MG_ND", None)
    f1 = fb.Flip.from_scene(sc, track_ids=1)
    _, st1 = f1.step(0.1)
    res = run_ranks([lambda f=f: f.step(0.1) for f in fs])
    for _, st in res:
        assert st.passes == st1.passes and st.n_fluid == st1.n_fluid
        assert abs(st.pcg_iters_total - st1.pcg_iters_total) <= st1.passes, (st.pcg_iters_total, st1.pcg_iters_total)
    assert rel_l2(gather_grid(fs, fb.GRID_P), f1.get_grid(fb.GRID_P)) < 1e-8
    # nd = 1: level 1 (24^3) is the first replicated level and has its own kernels, so the restriction
    # into it pushes its planes to every rank (+20); nd = 2: level 2 sits in the one-block tail, the
    # all-gather stays a collective copy
    for f in fs:
        assert (f.debug_info()["halo_mode"] >= 20) == (nd == 1), f.debug_info()
    for f in fs + [f1]:
        f.close()


def test_slab_seed_box_matches_single(fb):
    """Device seeding on slabs draws the same particles as the single-GPU seeding."""
    n = 32
    lo, hi = (2, 2, 2), (n - 3, n - 5, n - 3)
    f1 = fb.Flip(nx=n, ny=n, nz=n, ox=n // 2, oy=n // 2, oz=n // 2)
    f1.seed_box(lo, hi, seed=99)
    gid = next(_group)
    fs = [fb.Flip(nx=n, ny=n, nz=n, ox=n // 2, oy=n // 2, oz=n // 2, rank=r, world=2) for r in range(2)]
    for f in fs:
        f.comm_init_local(gid)
    for f in fs:
        f.seed_box(lo, hi, seed=99)
    assert sum(f.num_particles for f in fs) == f1.num_particles
    p1, _ = f1.get_particles()
    ps = np.concatenate([f.get_particles()[0] for f in fs])
    k1 = np.lexsort(p1.T[::-1]); ks = np.lexsort(ps.T[::-1])
    assert np.array_equal(p1[k1], ps[ks])
    # one step: migration sorts the boundary particles out, totals are conserved
    d1, st1 = f1.step(0.1)
    res = run_ranks([lambda f=f: f.step(0.1) for f in fs])
    assert sum(f.num_particles for f in fs) == f1.num_particles
    assert all(r[1].passes == st1.passes and r[1].n_fluid == st1.n_fluid for r in res)
    for f in fs + [f1]:
        f.close()


@pytest.mark.parametrize("fused,zsplit", [(1, None), (0, None), (1, [0, 8, 16, 48]), (1, [0, 32, 40, 48])])
def test_slab_wide_rows_mixed_precision(fb, fused, zsplit, monkeypatch):
    """Rows of >= 128 nodes switch on the four-node sweeps and with them params.mixed (search direction stored as fp32,
    level-0 multigrid right-hand side / residual as fp16): on slabs the ghost planes of those arrays travel at their
    storage width (pushed by the producing kernels, fused = 1, or as halo copies, fused = 0).  Same bars as above, plus
    the warm-started passes of the relaxation loop."""
    monkeypatch.setenv("FLIPB200_FUSED_HALO", str(fused))
    nx, ny, nz = 136, 24, (32 if zsplit is None else zsplit[-1])
    world = 2 if zsplit is None else len(zsplit) - 1
    kw = dict(nx=nx, ny=ny, nz=nz, ox=nx // 2, oy=ny // 2, oz=nz // 2)
    lo, hi = (2, 2, 2), (nx - 3, ny - 7, nz - 3)
    f1 = fb.Flip(**kw)
    f1.seed_box(lo, hi, seed=5)
    gid = next(_group)
    fs = [fb.Flip(rank=r, world=world, zsplit=zsplit, **kw) for r in range(world)]
    for f in fs:
        f.comm_init_local(gid)
    for f in fs:
        f.seed_box(lo, hi, seed=5)
    assert f1.params.mixed == 3 and f1.params.warm_start == 1
    d1, ds = 0.1, [0.1] * world
    for step in range(4):
        d1, st1 = f1.step(d1)
        res = run_ranks([lambda f=f, d=ds[r]: f.step(d) for r, f in enumerate(fs)])
        ds = [r[0] for r in res]
        for _, st in res:
            assert st.passes == st1.passes and st.n_fluid == st1.n_fluid, step
            assert abs(st.pcg_iters_total - st1.pcg_iters_total) <= 2 * st1.passes, (step, st.pcg_iters_total, st1.pcg_iters_total)
            assert st.pcg_rel_resid < 2.220446049250313e-16
        assert all(abs(d - d1) <= 1e-12 * d1 for d in ds)
        # the tank is at rest: u_x, u_z are round-off of the solve, so the three components are compared as one field
        us = np.stack([gather_grid(fs, fb.GRID_U + a) for a in range(3)]); u1 = np.stack([f1.get_grid(fb.GRID_U + a) for a in range(3)])
        assert rel_l2(us, u1) < 1e-9, step
        assert rel_l2(gather_grid(fs, fb.GRID_P), f1.get_grid(fb.GRID_P)) < 1e-8, step
    assert sum(f.num_particles for f in fs) == f1.num_particles
    for f in fs + [f1]:
        f.close()


def test_slab_caller_chosen_split(fb, scenes):
    """flipb200_create_slab_split: uneven slab boundaries (static load balancing along z, dist.balanced_split) give the
    same step as the single context; a bad split is rejected."""
    from fluid_simulation_b200 import dist as fdist
    sc = moving_scene(scenes, n=48)
    nx, ny, nz = sc.n
    ox, oy, oz = sc.origin
    zs = [0, 8, 16, 48]
    assert fdist.balanced_split(512, 8, 170, 343) == [0, 192, 216, 232, 256, 280, 296, 320, 512]
    f1 = fb.Flip.from_scene(sc, track_ids=1)
    gid = next(_group)
    fs = [fb.Flip(nx=nx, ny=ny, nz=nz, ox=ox, oy=oy, oz=oz, wall=sc.wall, track_ids=1, rank=r, world=3, zsplit=zs) for r in range(3)]
    assert [(f.z0, f.z1) for f in fs] == [(0, 8), (8, 16), (16, 48)]
    for f in fs:
        f.comm_init_local(gid)
    run_ranks([lambda f=f: f.set_particles(sc.pos, sc.vel) for f in fs])
    d1, ds = 0.1, [0.1] * 3
    for step in range(4):
        d1, st1 = f1.step(d1)
        res = run_ranks([lambda f=f, d=ds[r]: f.step(d) for r, f in enumerate(fs)])
        ds = [r[0] for r in res]
        assert all(r[1].passes == st1.passes and r[1].n_fluid == st1.n_fluid for r in res), step
        assert rel_l2(gather_grid(fs, fb.GRID_P), f1.get_grid(fb.GRID_P)) < 1e-8, step
    p1, v1 = f1.get_particles()
    ps, vs = gather_particles(fs, sc.num_particles)
    assert rel_l2(ps, p1) < 1e-10 and rel_l2(vs, v1) < 1e-9
    with pytest.raises(fb.FlipError):
        fb.Flip(nx=nx, ny=ny, nz=nz, ox=ox, oy=oy, oz=oz, rank=0, world=3, zsplit=[0, 7, 16, 48])     # odd boundary
    for f in fs + [f1]:
        f.close()


def test_slab_fluid_free_ranks_and_arrival(fb, scenes):
    """A small cube in the middle slab thrown along +z: the outer slabs start WITHOUT fluid (their share of the active
    region, grid.cu fb_active_region, is empty: no classify / cell-delta launch, P2G clears ghost planes only) and the
    upper one receives the fluid after a few steps through the particle migration.  Every step as on a single GPU."""
    n = 48
    sc = scenes.small_cube_drop(n=n, half=4, ppv=8, seed=3)
    sc.vel = np.zeros_like(sc.pos); sc.vel[:, 2] = 6.0; sc.vel[:, 1] = 2.0
    nx, ny, nz = sc.n
    ox, oy, oz = sc.origin
    zs = [0, 16, 32, 48]
    f1 = fb.Flip.from_scene(sc, track_ids=1)
    gid = next(_group)
    fs = [fb.Flip(nx=nx, ny=ny, nz=nz, ox=ox, oy=oy, oz=oz, wall=sc.wall, track_ids=1, rank=r, world=3, zsplit=zs) for r in range(3)]
    for f in fs:
        f.comm_init_local(gid)
    run_ranks([lambda f=f: f.set_particles(sc.pos, sc.vel) for f in fs])
    assert [f.num_particles for f in fs] == [0, sc.num_particles, 0]
    d1, ds = 0.1, [0.1] * 3
    for step in range(14):
        d1, st1 = f1.step(d1)
        res = run_ranks([lambda f=f, d=ds[r]: f.step(d) for r, f in enumerate(fs)])
        ds = [r[0] for r in res]
        assert all(r[1].passes == st1.passes and r[1].n_fluid == st1.n_fluid for r in res), step
        assert all(abs(d - d1) <= 1e-9 * d1 for d in ds), step
    assert fs[2].num_particles > 0                                   # the fluid has crossed into the upper slab
    for which in (fb.GRID_U, fb.GRID_W, fb.GRID_WEIGHT):
        assert rel_l2(gather_grid(fs, which), f1.get_grid(which)) < 1e-7, which
    assert np.array_equal(gather_grid(fs, fb.GRID_FLAGS), f1.get_grid(fb.GRID_FLAGS))
    p1, v1 = f1.get_particles()
    ps, vs = gather_particles(fs, sc.num_particles)
    assert rel_l2(ps, p1) < 1e-8 and rel_l2(vs, v1) < 1e-7
    for f in fs + [f1]:
        f.close()
