"""CPU tests of the oracle itself (no GPU): pins the restatement against the known-answer tests
the reference tree holds next to the path, and against the hand analysis in SURVEY.md 3.2."""
import numpy as np
import pytest

from helpers import oracle_for_scene


def test_spline_known_values(oracle_mod):
    # fluid.cc:22-37: spline(0)=1, spline(.5)=.25, spline(1)=0, even, zero outside radius 1
    s = oracle_mod.spline
    assert s(0.0) == pytest.approx(1.0, abs=1e-15)
    assert s(0.5) == pytest.approx(0.25, abs=1e-15)
    assert s(1.0) == 0.0 and s(1.5) == 0.0 and s(-2.0) == 0.0
    for x in np.linspace(0, 1, 37):
        assert s(x) == s(-x)
    # C1-continuous at 0.5 (both branches meet)
    assert s(0.5 - 1e-9) == pytest.approx(s(0.5 + 1e-9), abs=1e-8)


def test_eigen_kat_5x5(oracle_mod):
    """openvdb/unittest/TestConjGradient.cc:66-92: 5x5 SPD system, b = 1, expected x to 1e-5."""
    if not oracle_mod.load().oracle_has_eigen():
        pytest.skip("oracle built without the reference's Eigen")
    A = np.zeros((5, 5))
    for (i, j, v) in [(0, 0, 24.0), (0, 2, 6.0), (1, 1, 8.0), (1, 2, 2.0), (2, 0, 6.0), (2, 1, 2.0), (2, 2, 8.0),
                      (2, 3, -6.0), (2, 4, 2.0), (3, 2, -6.0), (3, 3, 24.0), (4, 2, 2.0), (4, 4, 8.0)]:
        A[i, j] = v
    x, it = oracle_mod.kat_solve_dense(A, np.ones(5))
    expected = np.array([0.0104167, 0.09375, 0.125, 0.0729167, 0.09375])
    assert np.allclose(x, expected, atol=1e-5)
    assert 0 <= it <= 20   # IC is an exact factor of this tiny matrix: Eigen breaks inside iteration 0


def test_single_particle_on_node(oracle_mod):
    """A particle sitting exactly on a node gives that node weight 1 and its own velocity."""
    o = oracle_mod.Oracle(nx=13, ny=13, nz=13, ox=6, oy=6, oz=6)
    o.set_particles(np.array([[1.0, -2.0, 0.0]]), np.array([[0.3, -0.7, 1.1]]))
    o.clear_vel(); o.p2g()
    w = o.fgrid(oracle_mod.F_W)
    assert w[6, 4, 7] == pytest.approx(1.0)
    assert np.count_nonzero(w) == 1
    for a, v in enumerate((0.3, -0.7, 1.1)):
        assert o.dgrid(a)[6, 4, 7] == pytest.approx(v, rel=1e-15)


def _hydrostatic(oracle_mod, solver):
    """Analogue of openvdb/unittest/TestPoissonSolver.cc:215-290 (open-top tank, P=0 at the free
    surface, Neumann walls): fluid at rest under gravity gives p = rho*|g|*dx*depth exactly, where
    depth counts cells up to the first air cell."""
    n = 13
    o = oracle_mod.Oracle(nx=n, ny=n, nz=n, ox=6, oy=6, oz=6, solver=solver)
    # fill interior x,z fully, y from bottom (index 2) up to index 7 with one particle per node
    xs = np.arange(2, n - 2) - 6
    ys = np.arange(2, 8) - 6
    X, Y, Z = np.meshgrid(xs, ys, xs, indexing="ij")
    pos = np.stack([X.ravel(), Y.ravel(), Z.ravel()], 1).astype(np.float64)
    o.set_particles(pos, np.zeros_like(pos))
    # dt = 1/8 is exact in fp32, so the FloatGrid-stored matrix coefficients (fluid.cc:333-335) carry
    # no rounding and the hydrostatic answer is exact; with dt = 0.1 the reference's own fp32
    # coefficient rounding perturbs p by ~4e-7 relative (checked below at 2e-6).
    dt = 0.125
    o.clear_vel(); o.p2g(); nf = o.classify()
    assert nf == 9 * 6 * 9
    o.rhs_div(dt); o.assemble(dt)
    it, err = o.solve()
    p = o.dgrid(oracle_mod.D_P)
    for j in range(2, 8):
        depth = 8 - j
        assert np.allclose(p[2:n - 2, j, 2:n - 2], 10.0 * depth, rtol=1e-12), (j, p[6, j, 6])
    # full-dt update leaves the fluid at rest (gravity balanced by the pressure gradient)
    o.vel_update(dt)
    for a in range(3):
        assert np.abs(o.dgrid(a)[:, :8, :]).max() < 1e-12
    # ... except the free-surface face: the air cell above gets +s*p_top but no gravity kick
    # (velUpdate fluid.cc:651-656 vs :638-642) -- the G0 term of SURVEY.md 3.2
    assert np.allclose(o.dgrid(1)[2:n - 2, 8, 2:n - 2], 0.125 * 10.0)
    o.clear_vel(); o.p2g(); o.classify(); o.rhs_div(0.1); o.assemble(0.1); o.solve()
    p = o.dgrid(oracle_mod.D_P)
    for j in range(2, 8):
        assert np.allclose(p[2:n - 2, j, 2:n - 2], 10.0 * (8 - j), rtol=2e-6)
    return it


def test_hydrostatic_reference_eigen(oracle_mod):
    if not oracle_mod.load().oracle_has_eigen():
        pytest.skip("no Eigen")
    assert _hydrostatic(oracle_mod, 0) < 60


def test_hydrostatic_port_jacobi(oracle_mod):
    _hydrostatic(oracle_mod, 1)


def test_step0_hand_analysis(oracle_mod):
    """from rest, pass 1 has b = 0 (x = 0, error = inf), then b_{k+1} = 0.9 b_k + G0:
    errors 0.9, 0.43, 0.27, 0.19, 0.14, 0.11, 0.092 => exit after 8 passes."""
    from fluid_simulation_b200 import scenes
    sc = scenes.small_cube_drop(n=25, half=4)
    o = oracle_for_scene(oracle_mod, sc)
    dt, st = o.step(0.1)
    err, its = o.pass_history()
    assert st.passes == 8
    assert np.isinf(err[0]) and its[0] == 0
    expect = [0.9, 0.4263, 0.2690, 0.1908, 0.1442, 0.1134, 0.0917]
    assert np.allclose(err[1:], expect, atol=2e-3)
    assert dt == pytest.approx(0.1)   # gravity kick 8*0.01*10 = 0.8 < dx/0.1


def test_jacobi_port_matches_eigen(oracle_mod):
    """The oracle's own Jacobi-PCG (used only where IC is too slow) agrees with the reference solver."""
    if not oracle_mod.load().oracle_has_eigen():
        pytest.skip("no Eigen")
    from fluid_simulation_b200 import scenes
    from helpers import random_velocity_scene
    sc = random_velocity_scene(scenes)
    res = []
    for solver in (0, 1):
        o = oracle_for_scene(oracle_mod, sc, solver=solver)
        o.clear_vel(); o.p2g(); o.classify(); o.rhs_div(0.1); o.assemble(0.1)
        o.solve()
        res.append(o.p.copy())
    assert np.linalg.norm(res[0] - res[1]) / np.linalg.norm(res[0]) < 1e-10


def test_reference_alternative_scenes():
    """The reference driver's commented-out scenes (fluid.cc:1176-1216 fluid shapes, :1303-1345
    obstacles): voxel boxes and scatter counts as the driver would produce them."""
    from fluid_simulation_b200 import scenes
    sc = scenes.reference_scene("normal")
    assert sc.num_particles == 689210                      # 10 * 41^3, all inside |p| < 58 (fluid.cc:841)
    assert sc.solid is None and sc.n == (121, 121, 121)
    pea = scenes.reference_scene("pea")
    assert pea.num_particles == 10 * 27 + 10 * 36          # cube [-1,1]^3 and the drop above it (j = 6..9)
    assert pea.pos[:, 1].max() < 9.5 and pea.pos[:, 1].min() >= -1.5
    side = scenes.reference_scene("side", obstacle="three_blocks")
    # boxes reach coordinate 57: voxel 57 spans [56.5, 57.5), PointList::add keeps |p| < 58
    assert np.all(np.abs(side.pos) < 58)
    assert int(side.solid.sum()) == 3 * 5 * 51 * 7         # x 5 wide, y -58..-8, z -3..3
    assert side.solid[60, 10, 60] == 1 and side.solid[60, 10, 64] == 0     # (z, y, x): centre block spans x in [-2, 2]
    c = np.rint(side.pos).astype(int) + 60
    assert not np.any(side.solid[c[:, 2], c[:, 1], c[:, 0]])
    wall = scenes.reference_scene("stable", obstacle="big_wall")
    assert int(wall.solid.sum()) == 117 * 9 * 6


def test_interp_from_grid_uniform_field(oracle_mod):
    """interpFromGrid / clampedCatmullRom (fluid.cc:883-895, 125-207) reproduces a uniform velocity
    field exactly wherever the weight sum is non-zero, and returns 0 where no node is in bounds."""
    n = 13
    o = oracle_mod.Oracle(nx=n, ny=n, nz=n, ox=6, oy=6, oz=6)
    for a, val in enumerate((1.5, -2.0, 0.25)):
        o.dgrid(a)[...] = val
    pos = np.array([[0.2, -0.3, 0.4], [3.9, 3.9, -3.9], [0.0, 0.0, 0.0]])
    o.set_particles(pos, np.full_like(pos, 7.0))
    o.interp_from_grid(0, 2)
    _, v = o.get_particles()
    assert np.allclose(v[:2], [[1.5, -2.0, 0.25]] * 2, rtol=1e-14)
    assert np.all(v[2] == 7.0)                             # outside [first, first + count): untouched


def test_extrapolate_uniform_velocity(oracle_mod):
    """extrapolate (fluid.cc:705-802, run at the end of P2Gtransfer when params.extrapolate = 1): a fluid
    blob moving with one velocity fills every non-solid in-bounds air node with that velocity; solid
    and out-of-bounds nodes stay 0; without the flag the air stays 0 (what main() runs)."""
    from fluid_simulation_b200 import scenes
    sc = scenes.small_cube_drop(n=17, half=3)
    sc.vel = np.tile(np.array([0.5, -1.25, 2.0]), (sc.num_particles, 1))
    solid = np.zeros((17, 17, 17), dtype=np.uint8)
    solid[3:6, 3:6, 3:6] = 1
    for ext in (0, 1):
        o = oracle_for_scene(oracle_mod, sc, solid=solid, extrapolate=ext)
        o.clear_vel(); o.p2g()
        u = np.stack([o.dgrid(a) for a in range(3)], -1)
        inside = np.zeros((17, 17, 17), dtype=bool); inside[2:15, 2:15, 2:15] = True
        air = inside & (solid == 0) & (o.fgrid(oracle_mod.F_W) == 0)
        assert air.sum() > 1000
        if ext:
            assert np.allclose(u[air], [0.5, -1.25, 2.0], rtol=1e-6)      # node velocities carry the fp32 weight-sum rounding (1e-8)
        else:
            assert np.all(u[air] == 0.0)
        assert np.all(u[~inside] == 0.0) and np.all(u[solid == 1] == 0.0)
