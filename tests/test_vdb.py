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


def test_header_and_descriptor_bytes(vdblib, tmp_path):
    """Byte-level pin of the writer against the REFERENCE'S writer code (not against the sibling reader): every field of
    the file header and of the first grid descriptor, with the reference line that emits it."""
    import struct
    vdblib.flipvdb_write_file_named.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                                C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.c_char_p)]
    g = [np.ones((2, 2, 2), np.float32) for _ in range(4)]
    ptrs = (C.c_void_p * 4)(*[a.ctypes.data for a in g])
    names = (C.c_char_p * 4)(b"", b"density", b"", b"density")
    p = tmp_path / "h.vdb"
    assert vdblib.flipvdb_write_file_named(str(p).encode(), 4, (C.c_int * 3)(0, 0, 0), (C.c_int * 3)(2, 2, 2), ptrs, 0, names) == 0
    raw = open(p, "rb").read()
    o = 0
    assert raw[o:o + 8] == struct.pack("<q", 0x56444220); o += 8        # OPENVDB_MAGIC (version.h:83), int64: Archive.cc:945-946
    assert raw[o:o + 4] == struct.pack("<I", 224); o += 4               # OPENVDB_FILE_VERSION (version.h:96): Archive.cc:949-950
    assert raw[o:o + 8] == struct.pack("<II", 4, 0); o += 8             # library major, minor (version.h:46-47): Archive.cc:953-956
    assert raw[o:o + 1] == b"\x01"; o += 1                              # hasGridOffsets (seekable file): Archive.cc:959-960
    uuid = raw[o:o + 36].decode("ascii"); o += 36                       # 36-char uuid string, no length prefix: Archive.cc:965-971
    assert [i for i, ch in enumerate(uuid) if ch == "-"] == [8, 13, 18, 23] and all(ch in "0123456789abcdef-" for ch in uuid)
    assert raw[o:o + 4] == struct.pack("<I", 0); o += 4                 # file-level MetaMap: count 0 (MetaMap.cc:117; File::write passes an empty map)
    assert raw[o:o + 4] == struct.pack("<i", 4); o += 4                 # grid count, int32: Archive.cc:1166-1170

    def pstr(off):                                                      # Name I/O: uint32 length + bytes (util/Name.h:58-63)
        n, = struct.unpack_from("<I", raw, off)
        return raw[off + 4:off + 4 + n], off + 4 + n
    nm, o = pstr(o)
    assert nm == b"\x1e0"                                               # empty name -> addSuffix(name, 0): Archive.cc:1203-1206, GridDescriptor.cc:178-182
    ty, o = pstr(o)
    assert ty == b"Tree_float_5_4_3"                                    # Tree::treeType(): tree/Tree.h:2129; GridDescriptor.cc:87
    par, o = pstr(o)
    assert par == b""                                                   # instance parent name: GridDescriptor.cc:90
    gpos, bpos, epos = struct.unpack_from("<qqq", raw, o); o += 24      # grid / block / end stream offsets: GridDescriptor.cc:93-99
    assert gpos == o and gpos < bpos < epos <= len(raw)
    assert raw[o:o + 4] == struct.pack("<I", 2)                         # per-grid compression flags, COMPRESS_ACTIVE_MASK (io/Compression.h:78-81): Archive.cc:1269
    # unique-name rule for the other three grids (Archive.cc:1199-1210)
    from vdb_reader import read_vdb
    assert [x["name"] for x in read_vdb(str(p))] == ["\x1e0", "density", "\x1e1", "density\x1e1"]
