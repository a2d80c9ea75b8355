"""ctypes binding of oracle/_ref/libfluid_ref.so -- the REFERENCE'S OWN fluid.cc, compiled unmodified
over the dense-array OpenVDB/TBB stand-in of oracle/shim/ (see oracle/ref_bridge.cc).

TEST INFRASTRUCTURE: only tests/ may import this module.  It exists to pin oracle/flip_oracle.cc (the
restatement the CUDA path is compared with) to the reference's own functions, bit for bit.
Grids are (121, 121, 121) numpy arrays indexed [z+60, y+60, x+60] -- the oracle's layout.
"""
import ctypes as C
import os
import re
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "_ref", "libfluid_ref.so")
N, O = 121, 60
SHAPE = (N, N, N)

F_WEIGHTS, F_CONTAINER, F_RHS, F_DIVER, F_ADIAG, F_APLUSI, F_APLUSJ, F_APLUSK, F_OUTPUT, F_SOLID = range(10)
V_VELS, V_VEL_BEFORE, V_NORMALS = range(3)

_lib = None


def available():
    return os.path.exists(LIB) or os.path.isfile("/root/reference/fluid.cc")


def build():
    if os.path.isfile("/root/reference/fluid.cc"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "fluidref"], env={**os.environ, "CXX": "/usr/bin/g++"})


def load():
    global _lib
    if _lib is not None:
        return _lib
    build()
    lib = C.CDLL(LIB)
    vp, d, i64 = C.c_void_p, C.c_double, C.c_int64
    lib.ref_spline.restype = d; lib.ref_spline.argtypes = [d]
    lib.ref_set_threads.argtypes = [C.c_int]
    lib.ref_world_create.restype = vp
    lib.ref_world_destroy.argtypes = [vp]
    for nm in ("ref_get_fgrid", "ref_set_fgrid"):
        getattr(lib, nm).argtypes = [vp, C.c_int, vp]
    for nm in ("ref_get_vgrid", "ref_set_vgrid"):
        getattr(lib, nm).argtypes = [vp, C.c_int, vp, vp, vp]
    lib.ref_get_indices.argtypes = [vp, vp]
    lib.ref_set_indices.argtypes = [vp, vp, C.c_int]
    lib.ref_get_defined.argtypes = [vp, C.c_int, vp]
    lib.ref_set_defined.argtypes = [vp, C.c_int, vp]
    lib.ref_set_particles.argtypes = [vp] + [vp] * 6 + [i64]
    lib.ref_num_particles.restype = i64; lib.ref_num_particles.argtypes = [vp]
    lib.ref_get_particles.argtypes = [vp] + [vp] * 6
    lib.ref_mark_uninitialized.argtypes = [vp, i64, i64]
    for nm in ("ref_clear_vel", "ref_P2Gtransfer", "ref_extrapolate", "ref_clear_container", "ref_interpolate",
               "ref_save_vel_before", "ref_interpFromGrid"):
        getattr(lib, nm).argtypes = [vp]; getattr(lib, nm).restype = None
    lib.ref_clear_pass_grids.argtypes = [vp, C.c_int]
    lib.ref_setRHS.argtypes = [vp, d, d, d, d, d]
    lib.ref_setDiver.argtypes = [vp, d]
    lib.ref_setA.argtypes = [vp, d, d, d]
    lib.ref_setA2.argtypes = [vp]; lib.ref_setA2.restype = i64
    lib.ref_get_A.argtypes = [vp, vp, vp, vp]
    lib.ref_get_b.argtypes = [vp, C.c_int, vp]
    lib.ref_solve.argtypes = [vp, C.POINTER(d)]; lib.ref_solve.restype = C.c_int
    lib.ref_set_p.argtypes = [vp, vp]
    lib.ref_velUpdate.argtypes = [vp, d, d, d, d, d, d]
    lib.ref_setOnlyB_error.argtypes = [vp]; lib.ref_setOnlyB_error.restype = d
    lib.ref_FLIPadvect.argtypes = [vp, d, d, d]; lib.ref_FLIPadvect.restype = d
    lib.ref_advect.argtypes = [vp, d, d, d]; lib.ref_advect.restype = d
    lib.ref_main_run.argtypes = [C.c_int]; lib.ref_main_run.restype = C.c_int
    lib.ref_main_num_particles.restype = i64
    lib.ref_main_num_initial.restype = i64
    lib.ref_main_passes.argtypes = [vp]
    lib.ref_main_particles.argtypes = [vp, vp, vp]
    lib.ref_main_vels.argtypes = [vp, vp, vp]
    lib.ref_main_fgrid.argtypes = [C.c_int, vp]
    lib.ref_main_indices.argtypes = [vp]
    lib.ref_main_log.argtypes = [vp, i64]; lib.ref_main_log.restype = i64
    _lib = lib
    return lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def set_threads(t):
    """threads of the tbb::parallel_for stand-in: 1 = serial in index order (pinning), > 1 = timing runs"""
    load().ref_set_threads(int(t))


def spline(x):
    return load().ref_spline(float(x))


class RefWorld:
    """main()'s grids (fluid.cc:1157-1246) + a PointList, with the reference's functions as methods."""

    def __init__(self, solid=None):
        self.lib = load()
        self.h = self.lib.ref_world_create()
        self.num_active = 0
        if solid is not None:
            self.set_f(F_SOLID, np.asarray(solid, dtype=np.float32))

    def close(self):
        if self.h:
            self.lib.ref_world_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- data ----
    def get_f(self, which):
        a = np.empty(SHAPE, np.float32); self.lib.ref_get_fgrid(self.h, which, _p(a)); return a

    def set_f(self, which, a):
        a = np.ascontiguousarray(a, dtype=np.float32); assert a.shape == SHAPE
        self.lib.ref_set_fgrid(self.h, which, _p(a))

    def get_v(self, which=V_VELS):
        u = [np.empty(SHAPE, np.float64) for _ in range(3)]
        self.lib.ref_get_vgrid(self.h, which, *[_p(x) for x in u]); return np.stack(u)

    def set_v(self, which, uvw):
        u = [np.ascontiguousarray(uvw[a], dtype=np.float64) for a in range(3)]
        self.lib.ref_set_vgrid(self.h, which, *[_p(x) for x in u])

    def get_indices(self):
        a = np.empty(SHAPE, np.int32); self.lib.ref_get_indices(self.h, _p(a)); return a

    def set_indices(self, idx, num_active):
        a = np.ascontiguousarray(idx, dtype=np.int32); self.num_active = int(num_active)
        self.lib.ref_set_indices(self.h, _p(a), self.num_active)

    def get_defined(self, which=0):
        a = np.empty(SHAPE, np.uint8); self.lib.ref_get_defined(self.h, which, _p(a)); return a

    def set_particles(self, pos, vel):
        a = [np.ascontiguousarray(pos[:, i], dtype=np.float64) for i in range(3)]
        b = [np.ascontiguousarray(vel[:, i], dtype=np.float64) for i in range(3)]
        self.lib.ref_set_particles(self.h, *[_p(x) for x in a + b], a[0].size)

    def get_particles(self):
        n = self.lib.ref_num_particles(self.h)
        o = [np.empty(n, np.float64) for _ in range(6)]
        self.lib.ref_get_particles(self.h, *[_p(x) for x in o])
        return np.stack(o[:3], 1), np.stack(o[3:], 1)

    def mark_uninitialized(self, first, count):
        self.lib.ref_mark_uninitialized(self.h, int(first), int(count))

    # ---- the reference's functions ----
    def clear_vel(self): self.lib.ref_clear_vel(self.h)
    def P2Gtransfer(self): self.lib.ref_P2Gtransfer(self.h)
    def extrapolate(self): self.lib.ref_extrapolate(self.h)
    def clear_container(self): self.lib.ref_clear_container(self.h)
    def interpolate(self): self.lib.ref_interpolate(self.h)
    def save_vel_before(self): self.lib.ref_save_vel_before(self.h)
    def clear_pass_grids(self, all_grids=True): self.lib.ref_clear_pass_grids(self.h, 1 if all_grids else 0)
    def setRHS(self, dx, g, dt): self.lib.ref_setRHS(self.h, dx, g[0], g[1], g[2], dt)
    def setDiver(self, dx): self.lib.ref_setDiver(self.h, dx)
    def setA(self, dx, dt, rho): self.lib.ref_setA(self.h, dx, dt, rho)

    def setA2(self):
        nnz = self.lib.ref_setA2(self.h)
        n = self.num_active
        outer = np.empty(n + 1, np.int32); inner = np.empty(nnz, np.int32); vals = np.empty(nnz, np.float64)
        self.lib.ref_get_A(self.h, _p(outer), _p(inner), _p(vals))
        return outer, inner, vals

    def vec(self, which):
        """0 = b, 1 = b2, 2 = p"""
        a = np.empty(self.num_active, np.float64); self.lib.ref_get_b(self.h, which, _p(a)); return a

    def solve(self):
        e = C.c_double(0); it = self.lib.ref_solve(self.h, C.byref(e)); return it, e.value

    def set_p(self, p):
        a = np.ascontiguousarray(p, dtype=np.float64); assert a.size == self.num_active
        self.lib.ref_set_p(self.h, _p(a))

    def velUpdate(self, dx, dt_sub, rho, g): self.lib.ref_velUpdate(self.h, dx, dt_sub, rho, g[0], g[1], g[2])
    def setOnlyB_error(self): return self.lib.ref_setOnlyB_error(self.h)
    def FLIPadvect(self, max_dt, dx, dt): return self.lib.ref_FLIPadvect(self.h, max_dt, dx, dt)
    def advect(self, max_dt, dx, dt): return self.lib.ref_advect(self.h, max_dt, dx, dt)
    def interpFromGrid(self): self.lib.ref_interpFromGrid(self.h)


def number_cells(container, solid):
    """main()'s inline numbering loop (fluid.cc:1416-1433) in numpy: scan x outer, y, z inner;
    `!isSolid && |c| <= 58 && container > 0` gets numActive++ (others keep the -1 of :1388)."""
    c = np.arange(N) - O
    inside = (np.abs(c) <= 58)
    m = (solid != 1) & (container > 0) & inside[:, None, None] & inside[None, :, None] & inside[None, None, :]
    idx = np.full(SHAPE, -1, np.int32)
    mt = np.ascontiguousarray(m.transpose(2, 1, 0))            # [x, y, z]: C order = x outer, z inner
    it = np.full(mt.shape, -1, np.int32)
    it[mt] = np.arange(int(mt.sum()), dtype=np.int32)
    idx[...] = it.transpose(2, 1, 0)
    return idx, int(m.sum())


class MainRun:
    """fluid.cc main() run for `steps` steps of its own loop on its own scene."""

    def __init__(self, steps):
        lib = load()
        rc = lib.ref_main_run(int(steps))
        if rc != 0:
            raise RuntimeError("ref_main returned without reaching the requested step")
        n = lib.ref_main_num_particles(); n0 = lib.ref_main_num_initial()
        self.pos0 = np.empty((n0, 3)); self.pos = np.empty((n, 3)); self.vel = np.empty((n, 3))
        lib.ref_main_particles(_p(self.pos0), _p(self.pos), _p(self.vel))
        u = [np.empty(SHAPE) for _ in range(3)]
        lib.ref_main_vels(*[_p(x) for x in u]); self.vels = np.stack(u)
        self.f = {}
        for k, nm in enumerate(("container", "output", "diver", "rhs", "adiag")):
            a = np.empty(SHAPE, np.float32); lib.ref_main_fgrid(k, _p(a)); self.f[nm] = a
        self.indices = np.empty(SHAPE, np.int32); lib.ref_main_indices(_p(self.indices))
        ps = (C.c_long * steps)(); lib.ref_main_passes(ps); self.passes = list(ps)
        ln = lib.ref_main_log(None, 0); buf = C.create_string_buffer(ln + 1); lib.ref_main_log(buf, ln + 1)
        self.log = buf.value.decode()
        self.dts = [float(x) for x in re.findall(r"^DT (\S+)$", self.log, re.M)]          # :1386 and :1491, two per step
        self.errors = [float(x) for x in re.findall(r"^Error:\t(\S+)$", self.log, re.M)]  # :1499
