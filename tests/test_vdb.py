"""mygrids.vdb output: the self-contained VDB-224 writer round-trips through an
independent reader written from the reference's read path (tests/vdb_reader.py)."""
import ctypes as C
import importlib.util
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def vdblib():
    spec = importlib.util.spec_from_file_location("flipb200_build", os.path.join(ROOT, "fluid-simulation_b200", "build.py"))
    b = importlib.util.module_from_spec(spec); spec.loader.exec_module(b)
    lib = C.CDLL(b.build_host())
    lib.flipvdb_write_file.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                       C.POINTER(C.c_void_p), C.c_int]
    return lib


def _write(lib, path, grids, lo, zip_):
    dim = grids[0].shape[::-1]
    ptrs = (C.c_void_p * len(grids))(*[g.ctypes.data for g in grids])
    rc = lib.flipvdb_write_file(str(path).encode(), len(grids), (C.c_int * 3)(*lo), (C.c_int * 3)(*dim), ptrs, zip_)
    assert rc == 0


@pytest.mark.parametrize("zip_", [0, 1])
def test_roundtrip_reference_box(vdblib, tmp_path, zip_):
    """The reference's box [-60,60]^3 (8 root children, partially filled leaves), 2 grids per file."""
    from vdb_reader import read_vdb
    rng = np.random.default_rng(5)
    n = 121
    grids = [np.ascontiguousarray(rng.random((n, n, n), dtype=np.float32)) for _ in range(2)]
    grids[1][30:60] = 0.0                               # runs of zeros (compressible, still ACTIVE voxels)
    p = tmp_path / "mygrids.vdb"
    _write(vdblib, p, grids, (-60, -60, -60), zip_)
    out = read_vdb(str(p))
    assert len(out) == 2
    for k, g in enumerate(out):
        assert g["name"] == "\x1e%d" % k                # unnamed grids get the suffix (GridDescriptor.cc:178-182)
        assert g["flags"] == (2 | zip_)
        assert tuple(g["meta"]["file_bbox_min"]) == (-60, -60, -60) and tuple(g["meta"]["file_bbox_max"]) == (60, 60, 60)
        assert g["meta"]["file_voxel_count"] == n ** 3
        assert np.array_equal(g["dense"], grids[k])     # bit-exact floats, every voxel active
    if zip_:
        assert os.path.getsize(p) < 2 * 4 * n ** 3      # zip actually shrank the zero slab


def test_small_offcentre_box(vdblib, tmp_path):
    from vdb_reader import read_vdb
    g = np.arange(5 * 9 * 17, dtype=np.float32).reshape(5, 9, 17)       # (z,y,x)
    p = tmp_path / "a.vdb"
    _write(vdblib, p, [g], (3, -4, 120), 1)                             # crosses a 128-block boundary in z
    out = read_vdb(str(p))
    assert np.array_equal(out[0]["dense"], g) and tuple(out[0]["lo"]) == (3, -4, 120)


@pytest.mark.gpu
def test_run_sh_fluid_drop_in(tmp_path):
    """`./run.sh fluid` (the reference's entry point, run.sh:1-7): compiles the C++ driver, runs the
    default scene for a few steps, prints the reference's stdout lines and writes a readable
    mygrids.vdb whose grids carry the splatted particle weights."""
    import subprocess
    from vdb_reader import read_vdb
    env = dict(os.environ, FLUID_STEPS="3")
    r = subprocess.run([os.path.join(ROOT, "run.sh"), "fluid"], cwd=tmp_path, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    out = r.stdout
    for tok in ("Compiled", "DT 0.1", "Before", "After", "Error:\t", "Iteration:\t3", "Time delta:\t", "Time Taken"):
        assert tok in out, tok
    grids = read_vdb(str(tmp_path / "mygrids.vdb"))
    assert len(grids) == 3
    for g in grids:
        d = g["dense"]
        assert d.shape == (121, 121, 121) and np.isfinite(d).all() and d.min() >= 0.0
        assert d[0].max() == 0.0 and d[:, 0].max() == 0.0          # solid shell stays 0 (fluid.cc:1442-1445)
        assert 0.4 * 689210 < d.sum(dtype=np.float64) < 689210      # sum of spline weights per particle in (0.125, 1]


@pytest.mark.gpu
def test_run_sh_alternative_scene_with_reseeding(tmp_path):
    """The driver's commented-out scenes (FLUID_SCENE / FLUID_OBSTACLE, fluid.cc:1178-1216, 1303-1345)
    and the re-seeding hook (FLUID_RESEED, :1374-1382 / :1495-1497): the particle mass in the output
    grids grows when new particles are emitted."""
    import subprocess
    from vdb_reader import read_vdb
    env = dict(os.environ, FLUID_STEPS="4", FLUID_SCENE="pea", FLUID_OBSTACLE="three_blocks", FLUID_RESEED="2")
    r = subprocess.run([os.path.join(ROOT, "run.sh"), "fluid"], cwd=tmp_path, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "Iteration:\t4" in r.stdout
    grids = read_vdb(str(tmp_path / "mygrids.vdb"))
    assert len(grids) == 4
    mass = [float(g["dense"].sum(dtype=np.float64)) for g in grids]
    assert 0.4 * 630 < mass[0] < 630 and abs(mass[1] - mass[0]) < 0.2 * mass[0]      # 10 * (27 + 36) particles
    assert mass[2] > mass[1] + 0.4 * 8 * 63                                         # step 2 emitted 8 per voxel into both boxes
