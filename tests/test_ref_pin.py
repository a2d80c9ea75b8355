"""Pins the oracle (oracle/flip_oracle.cc, the restatement every CUDA parity test is compared with) to the
REFERENCE'S OWN CODE: /root/reference/fluid.cc compiled UNMODIFIED (main renamed) over the dense-array
OpenVDB/TBB stand-in of oracle/shim/ and the reference's vendored Eigen (oracle/ref_bridge.cc ->
oracle/_ref/libfluid_ref.so, `make -C oracle fluidref`).

Every comparison below is BIT FOR BIT (np.array_equal) unless a tolerance is written next to it: the
restatement and the reference are both compiled with the reference's flags (run.sh:3) and particle
loops run serially in index order on both sides (one legal tbb::parallel_for schedule).

CPU only (no GPU, no CUDA path): this is the oracle's own pinning, the `-m gpu` tests then compare the
CUDA path with the pinned oracle.
"""
import numpy as np
import pytest

import ref_py

pytestmark = pytest.mark.skipif(not ref_py.available(), reason="neither /root/reference/fluid.cc nor a prebuilt oracle/_ref/libfluid_ref.so")

N, O = ref_py.N, ref_py.O
G = (0.0, -10.0, 0.0)          # fluid.cc:1357
DX, RHO = 1.0, 1.0             # fluid.cc:1358, :1471


def shell():
    c = np.abs(np.arange(N) - O) > 58                    # fluid.cc:1264
    return (c[:, None, None] | c[None, :, None] | c[None, None, :])


def box_mask(lo, hi):
    m = np.zeros((N, N, N), bool)
    m[lo[2] + O:hi[2] + O + 1, lo[1] + O:hi[1] + O + 1, lo[0] + O:hi[0] + O + 1] = True
    return m


def scene(kind, seed=11):
    rng = np.random.default_rng(seed)
    solid = shell()
    if kind == "corner_splash":
        # a blob against three walls and an interior block, random velocities: wall and obstacle collisions,
        # solid-neighbour RHS terms on every axis, air/fluid/solid mixes in setA
        n = 40000
        pos = np.stack([rng.uniform(-57.9, -30.0, n), rng.uniform(-57.9, -38.0, n), rng.uniform(-57.9, -30.0, n)], 1)
        vel = rng.normal(0.0, 3.0, (n, 3))
        solid = solid | box_mask((-45, -58, -50), (-40, -46, -42))           # like the "blocks" obstacles, fluid.cc:1303-1327
        pos = pos[~solid[np.round(pos[:, 2]).astype(int) + O, np.round(pos[:, 1]).astype(int) + O, np.round(pos[:, 0]).astype(int) + O]]
        vel = vel[:len(pos)]
    elif kind == "ceiling_and_strays":
        # fluid under the +y / +x / +z walls, plus strays inside the shell and outside the box (window clamp :271-276,
        # out-of-box reads return the background, particles that sit in solid cells)
        n = 20000
        pos = np.stack([rng.uniform(35.0, 57.9, n), rng.uniform(40.0, 57.9, n), rng.uniform(35.0, 57.9, n)], 1)
        vel = rng.normal(0.0, 4.0, (n, 3)) + np.array([2.0, 3.0, 2.0])
        strays = np.stack([rng.uniform(57.0, 61.3, 300), rng.uniform(-61.3, 61.3, 300), rng.uniform(-61.3, -57.0, 300)], 1)
        pos = np.concatenate([pos, strays]); vel = np.concatenate([vel, rng.normal(0.0, 4.0, (300, 3))])
    elif kind == "sparse_sheet":
        # less than one particle per cell: isolated fluid cells, cells with all-air neighbours, ties at x.5
        n = 6000
        pos = np.stack([rng.uniform(-20.0, 20.0, n), rng.uniform(-3.0, 3.0, n), rng.uniform(-20.0, 20.0, n)], 1)
        pos[:500] = np.round(pos[:500]) + rng.choice([-0.5, 0.0, 0.5], (500, 3))   # exact ties of round()
        vel = rng.normal(0.0, 1.0, (n, 3))
    else:
        raise KeyError(kind)
    return solid, pos, vel


def make_pair(oracle_mod, solid, pos, vel, **okw):
    r = ref_py.RefWorld(solid=solid.astype(np.float32))
    r.set_particles(pos, vel)
    o = oracle_mod.Oracle(nx=N, ny=N, nz=N, ox=O, oy=O, oz=O, solid=(solid & ~shell()).astype(np.uint8), **okw)
    assert np.array_equal(o.solid != 0, solid)
    o.set_particles(pos, vel)
    return r, o


def ovel(o, first=0):
    return np.stack([o.dgrid(first + a) for a in range(3)])


def test_spline_is_the_references(oracle_mod):
    xs = np.concatenate([np.linspace(-1.5, 1.5, 3001), np.random.default_rng(0).uniform(-1.2, 1.2, 5000), [0.0, 0.5, -0.5, 1.0, -1.0]])
    for x in xs:
        assert oracle_mod.spline(x) == ref_py.spline(x)


@pytest.mark.parametrize("kind", ["corner_splash", "ceiling_and_strays", "sparse_sheet"])
def test_step_functions_bit_for_bit(oracle_mod, kind):
    """Every function of the step body, called in main()'s order on the same state, two relaxation passes."""
    F = oracle_mod
    solid, pos, vel = scene(kind)
    r, o = make_pair(oracle_mod, solid, pos, vel)
    dt = 0.0625 if kind == "sparse_sheet" else 0.1

    # vels.fill(0) :1378 + PointList::P2Gtransfer :1384 -> :1106-1148, p2gCatmullRom :265-299
    r.clear_vel(); r.P2Gtransfer()
    o.clear_vel(); o.p2g()
    assert np.array_equal(r.get_f(ref_py.F_WEIGHTS), o.fgrid(F.F_W))
    assert np.array_equal(r.get_v(ref_py.V_VELS), ovel(o))
    assert np.count_nonzero(o.fgrid(F.F_W)) > 1000

    # container = 0 :1396-1409, PointList::interpolate :1413 -> :843-882, numbering :1416-1433, output :1434-1448
    r.clear_container(); r.interpolate()
    n_o = o.classify()
    cont = r.get_f(ref_py.F_CONTAINER)
    assert np.array_equal(cont, o.fgrid(F.F_CONTAINER))
    idx, n_r = ref_py.number_cells(cont, r.get_f(ref_py.F_SOLID))
    assert n_r == n_o and np.array_equal(idx, o.indices)
    r.set_indices(idx, n_r)
    r.save_vel_before(); o.save_vel_before()                                  # :1455

    for _pass in range(2):
        # :1461-1472
        r.clear_pass_grids(True); r.setRHS(DX, G, dt); r.setDiver(DX); r.setA(DX, dt, RHO)
        outer, inner, vals = r.setA2()
        o.rhs_div(dt); o.assemble(dt)
        assert np.array_equal(r.get_f(ref_py.F_RHS), o.fgrid(F.F_RHS))
        assert np.array_equal(r.get_f(ref_py.F_DIVER), o.fgrid(F.F_DIVER))
        for a, b in ((ref_py.F_ADIAG, F.F_ADIAG), (ref_py.F_APLUSI, F.F_APLUSI), (ref_py.F_APLUSJ, F.F_APLUSJ), (ref_py.F_APLUSK, F.F_APLUSK)):
            assert np.array_equal(r.get_f(a), o.fgrid(b))
        assert np.array_equal(r.vec(0), o.b)
        # the matrix setA2 hands to Eigen (:540) == the one the oracle hands to the same Eigen
        rows, cols, tv = o.triplets()
        import scipy.sparse as sp
        Ao = sp.coo_matrix((tv, (rows, cols)), shape=(n_o, n_o)).tocsc(); Ao.sort_indices()
        Ar = sp.csc_matrix((vals, inner, outer), shape=(n_r, n_r)); Ar.sort_indices()
        assert np.array_equal(Ar.indptr, Ao.indptr) and np.array_equal(Ar.indices, Ao.indices) and np.array_equal(Ar.data, Ao.data)
        # cg.compute(A); p = cg.solve(b) :1473-1474 (the reference's Eigen on both sides)
        it_r, err_r = r.solve()
        it_o, err_o = o.solve()
        assert it_r == it_o and np.array_equal(r.vec(2), o.p)
        # velUpdate(dt/10) :1475 -> :612-703
        r.velUpdate(DX, dt / 10, RHO, G); o.vel_update(dt / 10)
        assert np.array_equal(r.get_v(ref_py.V_VELS), ovel(o))
        # :1477-1483
        r.clear_pass_grids(False); r.setRHS(DX, G, dt); r.setDiver(DX)
        e_r = r.setOnlyB_error(); e_o = o.relax_error(dt)
        assert np.array_equal(r.get_f(ref_py.F_DIVER), o.fgrid(F.F_DIVER))
        assert np.array_equal(r.vec(1), o.b2)
        # Eigen's norm() sums in packets, the oracle sequentially: not the same rounding
        assert e_r == pytest.approx(e_o, rel=1e-12)

    # PointList::FLIPadvect :1490 -> :972-1038 (CatmullRomFLIP :210-263, getVelocity :59-70)
    dt_r = r.FLIPadvect(0.1, DX, dt)
    dt_o = o.flip_advect(dt)
    assert dt_r == dt_o
    pr, vr = r.get_particles(); po, vo = o.get_particles()
    assert np.array_equal(pr, po) and np.array_equal(vr, vo)
    moved = np.linalg.norm(pr - pos, axis=1)
    assert moved.max() > 0.05
    if kind != "sparse_sheet":
        # the collision branch (:1011-1030) fired: some component became -0.0 / 0.0
        assert np.count_nonzero(vr == 0.0) > 0


def test_extrapolate_bit_for_bit(oracle_mod):
    """extrapolate (fluid.cc:705-802), called where the commented-out call sits (:1147, end of P2Gtransfer)."""
    solid, pos, vel = scene("corner_splash", seed=5)
    pos, vel = pos[:6000], vel[:6000]
    r, o = make_pair(oracle_mod, solid, pos, vel, extrapolate=1)
    r.clear_vel(); r.P2Gtransfer(); r.extrapolate()
    o.clear_vel(); o.p2g()
    assert np.array_equal(r.get_v(ref_py.V_VELS), ovel(o))
    assert np.count_nonzero(ovel(o)) > 3 * np.count_nonzero(o.fgrid(oracle_mod.F_W))   # velocities spread far beyond the splat


def test_pic_advect_and_interp_from_grid_bit_for_bit(oracle_mod):
    """PointList::advect (:904-971, PIC, e = 0.5) == oracle flip_ratio 0; PointList::interpFromGrid (:883-895)."""
    solid, pos, vel = scene("corner_splash", seed=9)
    r, o = make_pair(oracle_mod, solid, pos, vel, flip_ratio=0.0, restitution=0.5)
    r.clear_vel(); r.P2Gtransfer(); r.clear_container(); r.interpolate()
    o.clear_vel(); o.p2g(); o.classify()
    r.save_vel_before(); o.save_vel_before()
    # interpFromGrid on a range of "new" particles
    r.mark_uninitialized(100, 5000); r.interpFromGrid()
    o.interp_from_grid(100, 5000)
    pr, vr = r.get_particles(); po, vo = o.get_particles()
    assert np.array_equal(vr, vo) and not np.array_equal(vr, vel)
    dt_r = r.advect(0.1, DX, 0.1)
    dt_o = o.flip_advect(0.1)
    assert dt_r == dt_o
    pr, vr = r.get_particles(); po, vo = o.get_particles()
    assert np.array_equal(pr, po) and np.array_equal(vr, vo)


def test_reference_main_three_steps_bit_for_bit(oracle_mod):
    """fluid.cc main() :1151-1514 itself -- scene construction, its own 500-step loop left after step 3 through the
    io::File::write hook -- against oracle_step x 3 from the same initial particles.  Particles: the stand-in
    scatterer's (count rule of PointScatter.h:150: Index64(10.f)*41^3 = 689,210, fluid.cc:1176,1349)."""
    K = 3
    m = ref_py.MainRun(K)
    assert m.pos0.shape == (689210, 3) and m.pos.shape == (689210, 3)
    o = oracle_mod.Oracle(nx=N, ny=N, nz=N, ox=O, oy=O, oz=O)
    o.set_particles(m.pos0, np.zeros_like(m.pos0))              # PointList::add zeroes the velocity, :841
    dt = 0.1                                                     # :1367
    passes, errors, dts = [], [], []
    for _ in range(K):
        dts.append(dt)
        dt, st = o.step(dt)
        dts.append(dt)
        passes.append(st.passes); errors.append(st.last_error)
    assert passes == m.passes                                    # do..while passes (Adiag fills, :1461): 8 at step 0
    assert passes[0] == 8
    assert dts == m.dts                                          # "DT" lines :1386, :1491, all digits
    assert errors == pytest.approx(m.errors, rel=1e-12)          # "Error:" :1499 (Eigen packet sum vs sequential sum)
    po, vo = o.get_particles()
    assert np.array_equal(po, m.pos) and np.array_equal(vo, m.vel)
    assert np.array_equal(ovel(o), m.vels)
    assert np.array_equal(o.fgrid(oracle_mod.F_CONTAINER), m.f["container"])
    assert np.array_equal(o.fgrid(oracle_mod.F_OUTPUT), m.f["output"])
    assert np.array_equal(o.fgrid(oracle_mod.F_DIVER), m.f["diver"])
    assert np.array_equal(o.fgrid(oracle_mod.F_RHS), m.f["rhs"])
    assert np.array_equal(o.fgrid(oracle_mod.F_ADIAG), m.f["adiag"])
    assert np.array_equal(o.indices, m.indices)
