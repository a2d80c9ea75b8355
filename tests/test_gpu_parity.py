"""GPU parity tests: the CUDA path, called through the C ABI (ctypes), against the CPU oracle on the
same seeded inputs.  Tolerances are the north_star's: single-step grid velocity and pressure
rel-L2 <= 1e-5, particle positions after 10 steps rel-L2 <= 1e-4 (fp64).  Integer/byte results
(cell classification, pass counts) must match exactly."""
import os

import numpy as np
import pytest

from helpers import rel_l2, oracle_for_scene, random_velocity_scene

pytestmark = pytest.mark.gpu

TOL_GRID = 1e-5      # north_star: single-step grid velocity / pressure rel-L2
TOL_POS10 = 1e-4     # north_star: positions after 10 steps rel-L2
# kernel variants kept after the round-2 measurements (FLIPB200_P2G / FLIPB200_G2P are read once, at create):
# cell = one thread per cell, particle = one thread per particle, tma = rows staged in shared memory by TMA bulk copies
PARTICLE_MODES = ["cell", "particle", "tma"]
G2P_MODES = ["tma", "cell", "particle"]      # tma: ONE pass G2P + advection betting on dt == dt_max, two kernels when the bet is lost


@pytest.fixture(scope="module")
def scenes():
    from fluid_simulation_b200 import scenes as s
    return s


def _pair(fb, oracle_mod, sc, **kw):
    f = fb.Flip.from_scene(sc, track_ids=1, **kw)
    o = oracle_for_scene(oracle_mod, sc)
    return f, o


def _grids(fb, f, base=0):
    return np.stack([f.get_grid(base + a) for a in range(3)])


def _ogrids(o, base=0):
    return np.stack([o.dgrid(base + a) for a in range(3)])


@pytest.mark.parametrize("mode", PARTICLE_MODES)
def test_p2g_classify(fb, oracle_mod, scenes, mode):
    """a5/a6/a7: splat + normalise + classification vs the oracle (both P2G kernels)."""
    sc = random_velocity_scene(scenes)
    os.environ["FLIPB200_P2G"] = mode
    try:
        f, o = _pair(fb, oracle_mod, sc)
        f.p2g(); nf = f.classify()
    finally:
        os.environ.pop("FLIPB200_P2G", None)
    o.clear_vel(); o.p2g(); no = o.classify()
    assert nf == no
    fl = f.get_grid(fb.GRID_FLAGS)
    assert np.array_equal((fl & fb.FB_FLUID) != 0, o.indices >= 0)          # exact cell classification
    assert np.array_equal((fl & fb.FB_SOLID) != 0, o.solid == 1)
    w = f.get_grid(fb.GRID_WEIGHT)
    assert rel_l2(w, o.fgrid(oracle_mod.F_W)) < 5e-7                         # fp32 storage, order-dependent sum
    assert rel_l2(w, o.fgrid(oracle_mod.F_CONTAINER)) < 5e-7                 # container := weights (a6 fused away)
    assert rel_l2(_grids(fb, f), _ogrids(o)) < 1e-6                          # u = sum(cw v)/float(w)
    assert rel_l2(_grids(fb, f, fb.GRID_UB), _ogrids(o)) < 1e-6              # velBeforeUpdate copy


def _inject_oracle_state(fb, oracle_mod, f, o):
    """Make the device hold exactly the oracle's post-P2G fields so later stages compare bit-tightly."""
    for a in range(3):
        f.set_grid(fb.GRID_U + a, o.dgrid(a))


def test_rhs_div_pcg_velupdate_relax(fb, oracle_mod, scenes):
    """a9-a16 stage by stage from identical inputs."""
    sc = random_velocity_scene(scenes, vscale=2.0)
    f, o = _pair(fb, oracle_mod, sc)
    dt = 0.1
    f.p2g(); f.classify()
    o.clear_vel(); o.p2g(); o.classify(); o.save_vel_before()
    _inject_oracle_state(fb, oracle_mod, f, o)
    # rhs + divergence: float grid, same statement order => bit-exact
    bn2 = f.rhs_div(dt)
    o.rhs_div(dt); o.assemble(dt)
    b = f.get_grid(fb.GRID_B)
    assert np.array_equal(b, o.fgrid(oracle_mod.F_DIVER))
    assert bn2 == pytest.approx(float(np.sum(o.b.astype(np.float64) ** 2)), rel=1e-12)
    # pressure solve: Jacobi-PCG (CUDA) vs the reference's IC-PCG (Eigen), both to eps*|b|
    it, rel = f.pcg_solve(dt)
    ito, relo = o.solve()
    p = f.get_grid(fb.GRID_P)
    assert rel < 2.220446049250313e-16 and relo < 2.220446049250313e-16     # converged to the reference's tolerance
    assert rel_l2(p, o.dgrid(oracle_mod.D_P)) < 1e-9
    assert 0 < it <= 2 * o.num_active
    # velocity update from identical pressure
    f.set_grid(fb.GRID_P, o.dgrid(oracle_mod.D_P))
    f.vel_update(dt / 10); o.vel_update(dt / 10)
    assert np.array_equal(_grids(fb, f), _ogrids(o))                        # same op order, no FMA => bit-exact
    # relaxation error
    e = f.relax_error(dt); eo = o.relax_error(dt)
    assert e == pytest.approx(eo, rel=1e-12)


@pytest.mark.parametrize("vscale", [6.0, 1.5])
@pytest.mark.parametrize("mode", G2P_MODES)
def test_g2p_advect(fb, oracle_mod, scenes, mode, vscale, monkeypatch):
    """a17/a18 from identical grids; includes particles that hit the solid shell (all G2P kernels).
    vscale 6: max speed > dx / dt_max, dt comes from the speed (the one-pass kernel's bet is lost, two kernels run);
    vscale 1.5: dt == dt_max, the one-pass kernel's result is the one compared."""
    monkeypatch.setenv("FLIPB200_G2P", mode)
    sc = random_velocity_scene(scenes, n=21, half=8, vscale=vscale)   # particles touching the wall (shell at |c| >= 9)
    f, o = _pair(fb, oracle_mod, sc)
    dt = 0.1
    f.p2g(); f.classify()
    o.clear_vel(); o.p2g(); o.classify(); o.save_vel_before()
    o.rhs_div(dt); o.assemble(dt); o.solve(); o.vel_update(dt / 10)
    for a in range(3):
        f.set_grid(fb.GRID_U + a, o.dgrid(a))
        f.set_grid(fb.GRID_UB + a, o.dgrid(3 + a))
    dt_g = f.g2p_advect()
    dt_o = o.flip_advect(dt)
    assert dt_g == pytest.approx(dt_o, rel=1e-13)
    assert (dt_o == 0.1) == (vscale < 2.0)
    pg, vg = f.get_particles(); po, vo = o.get_particles()
    assert rel_l2(vg, vo) < 1e-12
    assert rel_l2(pg, po) < 1e-13
    # the collision branch must have fired for some particles (velocity components zeroed, e = 0)
    assert np.count_nonzero(vo == 0.0) > 0
    assert np.array_equal(vg == 0.0, vo == 0.0)


@pytest.mark.parametrize("precond", [0, 1])
def test_single_step_small(fb, oracle_mod, scenes, precond):
    sc = random_velocity_scene(scenes, vscale=0.5)
    f, o = _pair(fb, oracle_mod, sc, precond=precond)
    dt_g, st = f.step(0.1)
    dt_o, so = o.step(0.1)
    assert (st.passes, st.n_fluid) == (so.passes, so.n_fluid)
    assert abs(st.last_error - so.last_error) < 1e-6
    assert rel_l2(_grids(fb, f), _ogrids(o)) < TOL_GRID
    assert rel_l2(f.get_grid(fb.GRID_P), o.dgrid(oracle_mod.D_P)) < TOL_GRID
    assert dt_g == pytest.approx(dt_o, rel=1e-9)
    pg, vg = f.get_particles(); po, vo = o.get_particles()
    assert rel_l2(pg, po) < 1e-7 and rel_l2(vg, vo) < 1e-5


def test_ten_steps_small_cube_drop(fb, oracle_mod, scenes):
    """Miniature of BASELINE config 1 from rest: pass counts (8 at step 0), then 10 steps."""
    sc = scenes.small_cube_drop(n=25, half=4)
    f, o = _pair(fb, oracle_mod, sc)
    dg = do = 0.1
    for s in range(10):
        dg, st = f.step(dg)
        do, so = o.step(do)
        assert (st.passes, st.n_fluid) == (so.passes, so.n_fluid), f"step {s}"
        if s == 0:
            assert st.passes == 8
            assert rel_l2(_grids(fb, f), _ogrids(o)) < TOL_GRID
            assert rel_l2(f.get_grid(fb.GRID_P), o.dgrid(oracle_mod.D_P)) < TOL_GRID
    pg, vg = f.get_particles(); po, vo = o.get_particles()
    assert rel_l2(pg, po) < TOL_POS10
    assert dg == pytest.approx(do, rel=1e-6)


def test_default_scene_single_step(fb, oracle_mod, scenes):
    """BASELINE config 1 at full size (121^3, 689,210 particles): one step vs the oracle."""
    sc = scenes.cube_drop_default()
    f, o = _pair(fb, oracle_mod, sc)
    dt_g, st = f.step(0.1)
    dt_o, so = o.step(0.1)
    assert (st.passes, st.n_fluid) == (so.passes, so.n_fluid) == (8, so.n_fluid)
    assert rel_l2(_grids(fb, f), _ogrids(o)) < TOL_GRID
    assert rel_l2(f.get_grid(fb.GRID_P), o.dgrid(oracle_mod.D_P)) < TOL_GRID
    w = f.get_grid(fb.GRID_WEIGHT)
    assert rel_l2(w, o.fgrid(oracle_mod.F_OUTPUT)) < 5e-7                   # outputGrid (fluid.cc:1434-1448)
    pg, vg = f.get_particles(); po, vo = o.get_particles()
    assert rel_l2(pg, po) < 1e-7


def test_hydrostatic_kat(fb):
    """Analytic KAT (analogue of openvdb/unittest/TestPoissonSolver.cc:215-290): open-top tank at
    rest, p = rho*|g|*dx*depth exactly when dt is exact in fp32."""
    n = 13
    xs = np.arange(2, n - 2) - 6
    ys = np.arange(2, 8) - 6
    X, Y, Z = np.meshgrid(xs, ys, xs, indexing="ij")
    pos = np.stack([X.ravel(), Y.ravel(), Z.ravel()], 1).astype(np.float64)
    f = fb.Flip(nx=n, ny=n, nz=n, ox=6, oy=6, oz=6)
    f.set_particles(pos, np.zeros_like(pos))
    dt = 0.125
    f.p2g(); assert f.classify() == 9 * 6 * 9
    f.rhs_div(dt); it, rel = f.pcg_solve(dt)
    p = f.get_grid(fb.GRID_P)
    for j in range(2, 8):
        assert np.allclose(p[2:n - 2, j, 2:n - 2], 10.0 * (8 - j), rtol=1e-12)
    assert it < 60
    f.vel_update(dt)
    for a in range(3):
        assert np.abs(f.get_grid(fb.GRID_U + a)[:, :8, :]).max() < 1e-12


def test_edge_cases(fb, oracle_mod, scenes):
    # empty particle set: a step is a no-op, error = 0/0 = NaN exits after one pass (fluid.cc:1483-1484)
    f = fb.Flip(nx=17, ny=17, nz=17, ox=8, oy=8, oz=8)
    f.set_particles(np.zeros((0, 3)), np.zeros((0, 3)))
    dt, st = f.step(0.1)
    assert st.n_fluid == 0 and st.passes == 1 and dt == 0.1
    # particles outside the box / inside the shell: no grid contribution, free flight (Appendix A.15)
    pos = np.array([[30.0, 0.2, 0.1], [-7.4, 0.3, 0.2], [0.25, 0.5, -0.75]])
    vel = np.array([[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.5, -0.5, 0.25]])
    f.set_particles(pos, vel)
    o = oracle_mod.Oracle(nx=17, ny=17, nz=17, ox=8, oy=8, oz=8)
    o.set_particles(pos, vel)
    f2 = fb.Flip(nx=17, ny=17, nz=17, ox=8, oy=8, oz=8, track_ids=1)
    f2.set_particles(pos, vel)
    dg, st = f2.step(0.1); do, so = o.step(0.1)
    assert (st.passes, st.n_fluid) == (so.passes, so.n_fluid)
    pg, vg = f2.get_particles(); po, vo = o.get_particles()
    assert rel_l2(pg, po) < 1e-9 and rel_l2(vg, vo) < 1e-7


def test_interior_obstacle(fb, oracle_mod, scenes):
    """Interior solids (the reference's commented-out block scenes, fluid.cc:1303-1345)."""
    sc = random_velocity_scene(scenes, n=25, half=5, vscale=0.5)
    solid = np.zeros((25, 25, 25), dtype=np.uint8)
    solid[10:15, 4:10, 10:15] = 1          # a block under the fluid cube (z,y,x order)
    f = fb.Flip(nx=25, ny=25, nz=25, ox=12, oy=12, oz=12, track_ids=1, solid=solid)
    f.set_particles(sc.pos, sc.vel)
    o = oracle_mod.Oracle(nx=25, ny=25, nz=25, ox=12, oy=12, oz=12, solid=solid)
    o.set_particles(sc.pos, sc.vel)
    dg = do = 0.1
    for s in range(3):
        dg, st = f.step(dg); do, so = o.step(do)
        assert (st.passes, st.n_fluid) == (so.passes, so.n_fluid), s
    assert rel_l2(_grids(fb, f), _ogrids(o)) < TOL_GRID
    pg, _ = f.get_particles(); po, _ = o.get_particles()
    assert rel_l2(pg, po) < 1e-6


def test_step_host_matches_step(fb, scenes):
    """The host-buffer entry point (what the reference's main() would call) == device-resident step."""
    sc = random_velocity_scene(scenes, vscale=0.5)
    f1 = fb.Flip.from_scene(sc, track_ids=1)
    f2 = fb.Flip(nx=25, ny=25, nz=25, ox=12, oy=12, oz=12, track_ids=1)
    d1, s1 = f1.step(0.1)
    arrs = [np.ascontiguousarray(sc.pos[:, i]) for i in range(3)] + [np.ascontiguousarray(sc.vel[:, i]) for i in range(3)]
    out = np.zeros((25, 25, 25), dtype=np.float32)
    d2, s2 = f2.step_host(arrs, 0.1, output_grid=out)
    assert (s1.passes, s1.n_fluid, d1) == (s2.passes, s2.n_fluid, d2)
    p1, v1 = f1.get_particles()
    assert rel_l2(np.stack(arrs[:3], 1), p1) < 1e-12 and rel_l2(np.stack(arrs[3:], 1), v1) < 1e-9
    assert np.array_equal(out, f2.get_grid(fb.GRID_WEIGHT))


def test_seed_box_and_sort_invariants(fb):
    """Device seeding: exactly 8 particles per voxel, inside their voxel, deterministic in the seed."""
    f = fb.Flip(nx=33, ny=33, nz=33, ox=16, oy=16, oz=16)
    f.seed_box((4, 4, 4), (19, 11, 27), seed=7)
    assert f.num_particles == 16 * 8 * 24 * 8
    p, v = f.get_particles()
    vox = np.floor(p + 0.5).astype(int) + 16
    assert vox.min(0).tolist() == [4, 4, 4] and vox.max(0).tolist() == [19, 11, 27]
    _, counts = np.unique(vox, axis=0, return_counts=True)
    assert np.all(counts == 8) and np.all(v == 0)
    g = fb.Flip(nx=33, ny=33, nz=33, ox=16, oy=16, oz=16)
    g.seed_box((4, 4, 4), (19, 11, 27), seed=7)
    assert np.array_equal(g.get_particles()[0], p)
    # P2G of a uniform velocity field reproduces it on every node that received weight (normalisation)
    f.set_particles(p, np.tile(np.array([[0.25, -1.5, 3.0]]), (p.shape[0], 1)))
    f.p2g(); f.classify()
    w = f.get_grid(fb.GRID_WEIGHT)
    for a, val in enumerate((0.25, -1.5, 3.0)):
        u = f.get_grid(fb.GRID_U + a)
        assert np.allclose(u[w > 0], val, rtol=1e-6)     # float(w) rounding only
        assert np.all(u[w == 0] == 0)


@pytest.mark.parametrize("n", [128, 256])
def test_full_size_properties(fb, n):
    """BASELINE configs[1] and [2] sizes (128^3: 15.0 M particles; 256^3: 127 M particles, the headline
    configuration; 8 ppc full tank) through size-independent properties: particle count conservation
    under the cell sort, exact cell classification, hydrostatic pressure at rest (p = rho*|g|*depth
    within the reference's own fp32-coefficient rounding; dt = 0.125 is exact in fp32)."""
    h = n // 2
    f = fb.Flip(nx=n, ny=n, nz=n, ox=h, oy=h, oz=h)
    f.seed_box((2, 2, 2), (n - 3, n - 5, n - 3), seed=1234)
    assert f.num_particles == (n - 4) * (n - 6) * (n - 4) * 8
    f.p2g(); nf = f.classify()
    assert f.num_particles == (n - 4) * (n - 6) * (n - 4) * 8
    assert nf == (n - 4) * (n - 5) * (n - 4)   # every interior node below the free surface (+1 layer reached by the splat)
    w = f.get_grid(fb.GRID_WEIGHT)
    assert w.sum(dtype=np.float64) > 0
    dt = 0.125
    f.rhs_div(dt); it, rel = f.pcg_solve(dt)
    assert rel < 2.220446049250313e-16
    p = f.get_grid(fb.GRID_P)
    top = 2 + (n - 5)                      # first air node index in y
    col = p[h, 2:top, h]
    expect = 10.0 * (top - np.arange(2, top))
    assert np.allclose(col, expect, rtol=1e-9)
    f.close()


# ---- multigrid-preconditioned CG (params.precond = 1) ---------------------------------------------
def test_mgpcg_solve_matches_reference(fb, oracle_mod, scenes):
    """Same system, three solvers: Eigen IC-PCG (reference), Jacobi-PCG and MG-PCG (CUDA)."""
    sc = random_velocity_scene(scenes, n=41, half=12, vscale=2.0)
    o = oracle_for_scene(oracle_mod, sc)
    o.clear_vel(); o.p2g(); o.classify(); o.rhs_div(0.1); o.assemble(0.1)
    ito, relo = o.solve()
    res = {}
    for precond in (0, 1):
        f = fb.Flip.from_scene(sc, precond=precond)
        f.p2g(); f.classify()
        for a in range(3):
            f.set_grid(fb.GRID_U + a, o.dgrid(a))
        f.rhs_div(0.1)
        it, rel = f.pcg_solve(0.1)
        assert rel < 2.220446049250313e-16
        assert rel_l2(f.get_grid(fb.GRID_P), o.dgrid(oracle_mod.D_P)) < 1e-9
        res[precond] = it
    assert res[1] < 40 and res[1] * 4 < res[0], res     # V-cycle must actually precondition


def test_mgpcg_steps(fb, oracle_mod, scenes):
    """Full steps with the multigrid preconditioner: same pass counts and parity bounds."""
    sc = scenes.small_cube_drop(n=33, half=6)
    f = fb.Flip.from_scene(sc, track_ids=1, precond=1)
    o = oracle_for_scene(oracle_mod, sc)
    dg = do = 0.1
    for s in range(10):
        dg, st = f.step(dg); do, so = o.step(do)
        assert (st.passes, st.n_fluid) == (so.passes, so.n_fluid), f"step {s}"
        if s == 0:
            assert st.passes == 8
            assert rel_l2(_grids(fb, f), _ogrids(o)) < TOL_GRID
            assert rel_l2(f.get_grid(fb.GRID_P), o.dgrid(oracle_mod.D_P)) < TOL_GRID
    pg, _ = f.get_particles(); po, _ = o.get_particles()
    assert rel_l2(pg, po) < TOL_POS10


def test_mgpcg_obstacle_and_tank_128(fb, oracle_mod, scenes):
    # interior obstacle (Neumann inside the domain) through the hierarchy
    sc = random_velocity_scene(scenes, n=25, half=5, vscale=0.5)
    solid = np.zeros((25, 25, 25), dtype=np.uint8)
    solid[10:15, 4:10, 10:15] = 1
    f = fb.Flip(nx=25, ny=25, nz=25, ox=12, oy=12, oz=12, track_ids=1, solid=solid, precond=1)
    f.set_particles(sc.pos, sc.vel)
    o = oracle_mod.Oracle(nx=25, ny=25, nz=25, ox=12, oy=12, oz=12, solid=solid)
    o.set_particles(sc.pos, sc.vel)
    dg, st = f.step(0.1); do, so = o.step(0.1)
    assert (st.passes, st.n_fluid) == (so.passes, so.n_fluid)
    assert rel_l2(_grids(fb, f), _ogrids(o)) < TOL_GRID
    # BASELINE config 2 size: hydrostatic tank at 128^3, analytic answer, iteration count bounded
    n = 128
    f = fb.Flip(nx=n, ny=n, nz=n, ox=64, oy=64, oz=64, precond=1)
    f.seed_box((2, 2, 2), (125, 123, 125), seed=1234)
    f.p2g(); f.classify()
    f.rhs_div(0.125); it, rel = f.pcg_solve(0.125)
    assert rel < 2.220446049250313e-16 and it < 60, (it, rel)
    p = f.get_grid(fb.GRID_P)
    assert np.allclose(p[64, 2:125, 64], 10.0 * (125 - np.arange(2, 125)), rtol=1e-9)


@pytest.mark.parametrize("ratio", [0.0, 0.5, 0.95])
def test_pic_flip_blend(fb, oracle_mod, scenes, ratio):
    """flip_ratio < 1 (extension; the reference runs 1.0): r*(FLIP) + (1-r)*(PIC of advect /
    clampedCatmullRom, fluid.cc:904-971, 125-207) against the oracle's restatement, 3 steps."""
    sc = random_velocity_scene(scenes)
    f = fb.Flip.from_scene(sc, track_ids=1, flip_ratio=ratio)
    o = oracle_for_scene(oracle_mod, sc, flip_ratio=ratio)
    dtg = dto = 0.1
    for _ in range(3):
        dtg, st = f.step(dtg)
        dto, so = o.step(dto)
        assert st.passes == so.passes and st.n_fluid == so.n_fluid
        assert abs(dtg - dto) <= 1e-9 * dto
    pg, vg = f.get_particles()
    po, vo = o.get_particles()
    # the PIC part reads u = sum(cw v)/float(w) itself (FLIP only reads differences of it), so the documented
    # fp32 summation-order deviation of w (<= 5e-7) shows: measured 9e-8 at r = 0
    assert rel_l2(vg, vo) < 2e-6 and rel_l2(pg, po) < 1e-6


def test_stencil4_matches_stencil2(fb, monkeypatch):
    """The four-nodes-per-lane sweep kernels (rows >= 128 nodes: k_spmv4, k_mg_smooth4, k_mg_resid4,
    k_mg_restrict4, k_mg_prolong4) against the two-nodes-per-lane ones on a wide, flat tank with an
    x extent that is not a multiple of 4 (150 nodes): same iteration count (+-1) and the same pressure."""
    res = {}
    for st4 in ("1", "0"):
        monkeypatch.setenv("FLIPB200_ST4", st4)
        f = fb.Flip(nx=150, ny=20, nz=24, ox=75, oy=10, oz=12, precond=1)
        f.seed_box((2, 2, 2), (147, 14, 21), seed=99)       # the whole interior up to y = 14 (inclusive box): walls on every side
        f.p2g(); f.classify()
        f.rhs_div(0.125)
        it, rel = f.pcg_solve(0.125)
        assert rel < 2.220446049250313e-16
        res[st4] = (it, f.get_grid(fb.GRID_P))
        f.close()
    assert abs(res["1"][0] - res["0"][0]) <= 1, (res["1"][0], res["0"][0])
    assert rel_l2(res["1"][1], res["0"][1]) < 1e-9
    # hydrostatic answer: particles in the voxels up to y = 14 mark the nodes up to y = 15 fluid (the splat reaches
    # one node up, as in test_mgpcg_obstacle_and_tank_128), p = rho*|g|*dx*depth
    p = res["1"][1]
    assert np.allclose(p[12, 2:14, 70], 10.0 * (16 - np.arange(2, 14)), rtol=1e-9)     # (z, y, x)


def test_emit_box_interp_from_grid(fb, oracle_mod, scenes):
    """Re-seeding hook (fluid.cc:1374-1382, 1495-1497): particles appended on the device get the PIC
    velocity of PointList::interpFromGrid (:883-895, clampedCatmullRom :125-207) from the grid
    velocities of the step that just ran; the old particles are untouched."""
    sc = random_velocity_scene(scenes, n=25, half=5, vscale=1.0)
    f, o = _pair(fb, oracle_mod, sc)
    f.step(0.1); o.step(0.1)
    n0 = sc.num_particles
    p_before, v_before = f.get_particles()
    for a in range(3):                                    # same grid velocities on both sides: isolates the gather
        f.set_grid(fb.GRID_U + a, o.dgrid(a))
    f.emit_box((9, 14, 9), (15, 20, 15), seed=5)          # index box straddling the top of the fluid cube, up to the shell
    pg, vg = f.get_particles()
    add = pg.shape[0] - n0
    assert add == 8 * 7 * 7 * 7
    assert np.array_equal(pg[:n0], p_before) and np.array_equal(vg[:n0], v_before)
    po, vo = o.get_particles()
    o.set_particles(np.concatenate([po, pg[n0:]]), np.concatenate([vo, np.zeros((add, 3))]))
    o.interp_from_grid(n0, add)
    _, vo2 = o.get_particles()
    assert np.count_nonzero(np.abs(vo2[n0:]).sum(axis=1) > 0) > add // 4      # the box does overlap moving fluid
    assert np.max(np.abs(vg[n0:] - vo2[n0:])) < 1e-13 * max(1.0, np.abs(vo2[n0:]).max())
    # and the enlarged particle set steps on in step with the oracle
    dg, st = f.step(0.1); do, so = o.step(0.1)
    assert (st.passes, st.n_fluid) == (so.passes, so.n_fluid)
    pg2, _ = f.get_particles(); po2, _ = o.get_particles()
    assert rel_l2(pg2, po2) < 1e-7


def test_reference_obstacle_scene(fb, oracle_mod, scenes):
    """One of the reference driver's commented-out scenes at full size (121^3): the "side" fluid slab
    (fluid.cc:1192-1203) against the "big wall" obstacle (:1334-1345), one step against the oracle."""
    sc = scenes.reference_scene("side", obstacle="big_wall")
    f, o = _pair(fb, oracle_mod, sc)
    dg, st = f.step(0.1); do, so = o.step(0.1)
    assert (st.passes, st.n_fluid) == (so.passes, so.n_fluid)
    assert dg == pytest.approx(do, rel=1e-9)
    assert rel_l2(_grids(fb, f), _ogrids(o)) < TOL_GRID
    assert rel_l2(f.get_grid(fb.GRID_P), o.dgrid(oracle_mod.D_P)) < TOL_GRID
    pg, _ = f.get_particles(); po, _ = o.get_particles()
    assert rel_l2(pg, po) < 1e-7


def test_extrapolate(fb, oracle_mod, scenes):
    """params.extrapolate = 1 runs the reference's extrapolate (fluid.cc:705-802, call commented out at
    :1147) on the P2G velocities: wave by wave every air node gets the average of the nodes defined one
    wave earlier.  Grid after P2G and one full step against the oracle, with an interior obstacle."""
    sc = random_velocity_scene(scenes, n=25, half=5, vscale=1.0)
    solid = np.zeros((25, 25, 25), dtype=np.uint8)
    solid[10:15, 4:8, 10:15] = 1
    f = fb.Flip(nx=25, ny=25, nz=25, ox=12, oy=12, oz=12, track_ids=1, solid=solid, extrapolate=1)
    f.set_particles(sc.pos, sc.vel)
    o = oracle_mod.Oracle(nx=25, ny=25, nz=25, ox=12, oy=12, oz=12, solid=solid, extrapolate=1)
    o.set_particles(sc.pos, sc.vel)
    f.p2g(); f.classify()
    o.clear_vel(); o.p2g(); o.classify(); o.save_vel_before()
    ug, uo = _grids(fb, f), _ogrids(o)
    air = (o.fgrid(oracle_mod.F_W) == 0) & (solid == 0)
    assert np.count_nonzero(uo[0][air]) > 5000                 # the air really was filled
    # same bound as test_p2g_classify (u = sum(cw v) / float(w) carries the fp32 weight rounding), now
    # over the whole box, and in the air alone
    assert rel_l2(ug, uo) < 1e-6 and rel_l2(ug[:, air], uo[:, air]) < 1e-6
    assert np.max(np.abs(ug - uo)) < 1e-6 * np.abs(uo).max()
    assert rel_l2(_grids(fb, f, fb.GRID_UB), _ogrids(o, 3)) < 1e-6              # velBeforeUpdate is the extrapolated field too
    f.set_particles(sc.pos, sc.vel); o.set_particles(sc.pos, sc.vel)
    dg, st = f.step(0.1); do, so = o.step(0.1)
    assert (st.passes, st.n_fluid) == (so.passes, so.n_fluid)
    assert rel_l2(_grids(fb, f), _ogrids(o)) < TOL_GRID
    pg, _ = f.get_particles(); po, _ = o.get_particles()
    assert rel_l2(pg, po) < 1e-7
    with pytest.raises(fb.FlipError):                          # z-slab contexts do not support it
        fb.Flip(nx=25, ny=25, nz=25, ox=12, oy=12, oz=12, extrapolate=1, rank=0, world=2)


def test_bench_scene_steps_vs_oracle(fb, oracle_mod, scenes):
    """The scene bench.py times (dam-break full-tank fill, 8 particles per voxel, at rest) at 64^3, stepped 5 times against
    the oracle: relaxation pass counts exact, grid velocity / pressure <= 1e-5 after the first step, positions <= 1e-4."""
    n = 64
    sc = scenes.dam_break(n=n, fill="full")
    f, o = _pair(fb, oracle_mod, sc)
    dg = do = 0.1
    for s in range(5):
        dg, st = f.step(dg); do, so = o.step(do)
        assert (st.passes, st.n_fluid) == (so.passes, so.n_fluid), s
        assert dg == pytest.approx(do, rel=1e-9)
        if s == 0:
            assert st.n_fluid == (n - 4) * (n - 5) * (n - 4)
            assert rel_l2(_grids(fb, f), _ogrids(o)) < TOL_GRID
            assert rel_l2(f.get_grid(fb.GRID_P), o.dgrid(oracle_mod.D_P)) < TOL_GRID
    pg, vg = f.get_particles(); po, vo = o.get_particles()
    assert rel_l2(pg, po) < TOL_POS10 and rel_l2(vg, vo) < 1e-5


def test_against_the_reference_main(fb):
    """The CUDA path against THE REFERENCE ITSELF: fluid.cc main() compiled unmodified (oracle/_ref/libfluid_ref.so, see
    tests/test_ref_pin.py) runs one step of its own loop on its own scene; the CUDA path steps the same initial particles."""
    import ref_py
    if not ref_py.available():
        pytest.skip("oracle/_ref/libfluid_ref.so not built")
    m = ref_py.MainRun(1)
    f = fb.Flip(nx=121, ny=121, nz=121, ox=60, oy=60, oz=60, track_ids=1)          # the literals of fluid.cc:1159,1264
    f.set_particles(m.pos0, np.zeros_like(m.pos0))
    dt, st = f.step(0.1)                                                            # dt0 = 0.1, fluid.cc:1367
    assert st.passes == m.passes[0] == 8                                            # do..while passes
    assert dt == pytest.approx(m.dts[-1], rel=1e-12)
    assert st.last_error == pytest.approx(m.errors[0], rel=1e-6)
    u = np.stack([f.get_grid(fb.GRID_U + a) for a in range(3)])
    assert rel_l2(u, m.vels) < TOL_GRID                                             # north_star: single-step grid velocity <= 1e-5
    assert rel_l2(f.get_grid(fb.GRID_WEIGHT), m.f["output"]) < 5e-7                 # outputGrid, fluid.cc:1434-1448
    fl = f.get_grid(fb.GRID_FLAGS)
    assert np.array_equal((fl & fb.FB_FLUID) != 0, m.indices >= 0)                  # the cells main() numbered, :1416-1433
    pg, vg = f.get_particles()
    assert rel_l2(pg, m.pos) < 1e-7 and rel_l2(vg, m.vel) < 1e-5


def test_fp32_mode(fb, oracle_mod, scenes):
    """params.precision = 32 ("fp32 mode" of BASELINE.json's north_star): the pressure solve stops at FLT_EPSILON relative
    residual instead of DBL_EPSILON.  The north_star states no fp32 tolerance ("reported separately"); the one stated HERE:
    against the fp64 ORACLE on the same scene classification and pass counts exact, single-step grid velocity <= 1e-5 (the
    fp64 bar; measured ~1e-7, the solve's residual), pressure <= 1e-4, fewer CG iterations than precision = 64, positions
    after 5 steps <= 1e-4."""
    TOL_FP32 = 1e-4
    sc = scenes.dam_break(n=48, fill="full")
    f32, o = _pair(fb, oracle_mod, sc, precision=32)
    f64 = fb.Flip.from_scene(sc, track_ids=1)
    d32 = d64 = do = 0.1
    it32 = it64 = 0
    for s in range(5):
        d32, s32 = f32.step(d32); d64, s64 = f64.step(d64); do, so = o.step(do)
        assert (s32.passes, s32.n_fluid) == (so.passes, so.n_fluid), s
        assert s32.pcg_rel_resid < 1.2e-7
        it32 += s32.pcg_iters_total; it64 += s64.pcg_iters_total
        if s == 0:
            assert rel_l2(_grids(fb, f32), _ogrids(o)) < TOL_GRID
            assert rel_l2(f32.get_grid(fb.GRID_P), o.dgrid(oracle_mod.D_P)) < TOL_FP32
    assert it32 < it64
    pg, vg = f32.get_particles(); po, vo = o.get_particles()
    assert rel_l2(pg, po) < TOL_POS10 and rel_l2(vg, vo) < TOL_FP32
    with pytest.raises(fb.FlipError):
        fb.Flip(nx=25, ny=25, nz=25, ox=12, oy=12, oz=12, precision=16)


def test_sparse_region_matches_full_sweeps(fb, scenes, monkeypatch):
    """The per-node passes of a step (P2G clears and cell sweep, k_classify, k_cell_delta) only visit the last step's wet
    bounding box + 3 nodes (grid.cu fb_active_region; FLIPB200_SPARSE=0: the whole grid every step).  A cube falling and
    spreading for 12 steps, then a second body of fluid emitted far away (the region must open up again): every grid and
    every particle as with whole-grid sweeps, and no stale weight / velocity / flag left where the fluid has been."""
    res = {}
    for sparse in ("1", "0"):
        monkeypatch.setenv("FLIPB200_SPARSE", sparse)
        sc = random_velocity_scene(scenes, n=41, half=5, seed=7, vscale=3.0)
        f = fb.Flip.from_scene(sc, track_ids=1)
        dt = 0.1
        for s in range(12):
            dt, st = f.step(dt)
        f.emit_box((30, 25, 30), (35, 30, 35), seed=5)
        for s in range(3):
            dt, st = f.step(dt)
        res[sparse] = (np.stack([f.get_grid(w).astype(np.float64) for w in (fb.GRID_U, fb.GRID_V, fb.GRID_W, fb.GRID_UB, fb.GRID_P, fb.GRID_WEIGHT)]),
                       f.get_grid(fb.GRID_FLAGS), f.get_particles(), (st.passes, st.n_fluid, dt))
        f.close()
    a, b = res["1"], res["0"]
    # (the two runs differ by the order of the P2G atomics, like two runs of one setting: ~1e-10 after 15 steps)
    assert a[3][:2] == b[3][:2] and a[3][2] == pytest.approx(b[3][2], rel=1e-7)
    assert np.array_equal(a[1], b[1])                                   # flags: identical bytes
    assert np.array_equal(a[0][5] > 0, b[0][5] > 0)                     # wet nodes: identical set
    for k in range(6):
        assert rel_l2(a[0][k], b[0][k]) < 1e-7, k
    assert rel_l2(a[2][0], b[2][0]) < 1e-8 and rel_l2(a[2][1], b[2][1]) < 1e-7
