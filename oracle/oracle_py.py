"""ctypes binding of the CPU ORACLE (test infrastructure only -- see oracle/flip_oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(_HERE, "_ref", "libflip_oracle.so")      # with the reference's Eigen
PORT_LIB = os.path.join(_HERE, "libflip_oracle_port.so")        # restatement only


class OracleParams(C.Structure):
    _fields_ = [
        ("nx", C.c_int32), ("ny", C.c_int32), ("nz", C.c_int32),
        ("ox", C.c_int32), ("oy", C.c_int32), ("oz", C.c_int32),
        ("wall", C.c_int32), ("relax_div", C.c_int32),
        ("dx", C.c_double), ("rho", C.c_double),
        ("gx", C.c_double), ("gy", C.c_double), ("gz", C.c_double),
        ("dt_max", C.c_double), ("relax_tol", C.c_double), ("restitution", C.c_double),
        ("solver", C.c_int32), ("max_passes", C.c_int32), ("threads", C.c_int32),
        ("pcg_max_iter", C.c_int32),
        ("flip_ratio", C.c_double),
        ("extrapolate", C.c_int32), ("reserved0", C.c_int32),
    ]


class OracleStats(C.Structure):
    _fields_ = [
        ("passes", C.c_int32), ("n_fluid", C.c_int32), ("pcg_iters_total", C.c_int32),
        ("pcg_iters_last", C.c_int32),
        ("last_error", C.c_double), ("dt_out", C.c_double), ("max_speed", C.c_double),
        ("t_p2g", C.c_double), ("t_grid", C.c_double), ("t_assemble", C.c_double),
        ("t_solve", C.c_double), ("t_g2p", C.c_double),
    ]


F_W, F_CONTAINER, F_RHS, F_DIVER, F_ADIAG, F_APLUSI, F_APLUSJ, F_APLUSK, F_OUTPUT = range(9)
D_U, D_V, D_W, D_UB, D_VB, D_WB, D_P = range(7)

_lib = None


def build(force=False):
    """Run oracle/Makefile (make decides what is stale): the port always; oracle/_ref (restatement + the
    reference's Eigen, and the reference's own fluid.cc over oracle/shim) only when /root/reference is
    present -- on the GPU box the prebuilt files are used as they travelled."""
    have_ref = os.path.isdir("/root/reference/Eigen")
    if force or have_ref or not os.path.exists(PORT_LIB):
        subprocess.check_call(["make", "-s", "-C", _HERE, "all"], env={**os.environ, "CXX": "/usr/bin/g++"})


def load(prefer_ref=True):
    global _lib
    if _lib is not None:
        return _lib
    path = REF_LIB if (prefer_ref and os.path.exists(REF_LIB)) else PORT_LIB
    if not os.path.exists(path):
        build()
        path = REF_LIB if (prefer_ref and os.path.exists(REF_LIB)) else PORT_LIB
    lib = C.CDLL(path)
    dp = C.POINTER(C.c_double)
    lib.oracle_default_params.argtypes = [C.POINTER(OracleParams)]
    lib.oracle_has_eigen.restype = C.c_int
    lib.oracle_create.restype = C.c_void_p
    lib.oracle_create.argtypes = [C.POINTER(OracleParams), C.c_void_p]
    lib.oracle_destroy.argtypes = [C.c_void_p]
    lib.oracle_set_particles.argtypes = [C.c_void_p] + [C.c_void_p] * 6 + [C.c_int64]
    lib.oracle_num_particles.restype = C.c_int64
    lib.oracle_num_particles.argtypes = [C.c_void_p]
    lib.oracle_get_particles.argtypes = [C.c_void_p] + [C.c_void_p] * 6
    lib.oracle_spline.restype = C.c_double
    lib.oracle_spline.argtypes = [C.c_double]
    for name in ("oracle_clear_vel", "oracle_p2g", "oracle_save_vel_before"):
        getattr(lib, name).argtypes = [C.c_void_p]
        getattr(lib, name).restype = None
    lib.oracle_classify.argtypes = [C.c_void_p]
    lib.oracle_classify.restype = C.c_int
    lib.oracle_rhs_div.argtypes = [C.c_void_p, C.c_double]
    lib.oracle_assemble.argtypes = [C.c_void_p, C.c_double]
    lib.oracle_solve.argtypes = [C.c_void_p, dp]
    lib.oracle_solve.restype = C.c_int
    lib.oracle_vel_update.argtypes = [C.c_void_p, C.c_double]
    lib.oracle_relax_error.argtypes = [C.c_void_p, C.c_double]
    lib.oracle_relax_error.restype = C.c_double
    lib.oracle_flip_advect.argtypes = [C.c_void_p, dp]
    lib.oracle_interp_from_grid.argtypes = [C.c_void_p, C.c_int64, C.c_int64]
    lib.oracle_interp_from_grid.restype = None
    lib.oracle_extrapolate.argtypes = [C.c_void_p]
    lib.oracle_extrapolate.restype = None
    lib.oracle_step.argtypes = [C.c_void_p, dp, C.POINTER(OracleStats)]
    lib.oracle_step.restype = C.c_int
    lib.oracle_pass_history.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
    lib.oracle_pass_history.restype = C.c_int
    lib.oracle_fgrid.argtypes = [C.c_void_p, C.c_int]
    lib.oracle_fgrid.restype = C.POINTER(C.c_float)
    lib.oracle_dgrid.argtypes = [C.c_void_p, C.c_int]
    lib.oracle_dgrid.restype = dp
    lib.oracle_indices.argtypes = [C.c_void_p]
    lib.oracle_indices.restype = C.POINTER(C.c_int32)
    lib.oracle_solid.argtypes = [C.c_void_p]
    lib.oracle_solid.restype = C.POINTER(C.c_uint8)
    lib.oracle_num_active.argtypes = [C.c_void_p]
    lib.oracle_num_active.restype = C.c_int
    lib.oracle_b.argtypes = [C.c_void_p]
    lib.oracle_b.restype = dp
    lib.oracle_p.argtypes = [C.c_void_p]
    lib.oracle_p.restype = dp
    lib.oracle_b2.argtypes = [C.c_void_p]
    lib.oracle_b2.restype = dp
    lib.oracle_triplets.argtypes = [C.c_void_p, C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.POINTER(C.c_int32)), C.POINTER(dp)]
    lib.oracle_triplets.restype = C.c_int64
    lib.oracle_kat_solve_dense.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.oracle_kat_solve_dense.restype = C.c_int
    _lib = lib
    return lib


def default_params(**kw):
    lib = load()
    p = OracleParams()
    lib.oracle_default_params(C.byref(p))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    """Dense-array oracle context.  Grid arrays are (nz, ny, nx) numpy views (x fastest)."""

    def __init__(self, params=None, solid=None, **kw):
        self.lib = load()
        self.params = params if params is not None else default_params(**kw)
        if self.params.solver == 0 and not self.lib.oracle_has_eigen():
            self.params.solver = 1
        self.shape = (self.params.nz, self.params.ny, self.params.nx)
        self.N = int(np.prod(self.shape))
        sp = None
        if solid is not None:
            self._solid_in = np.ascontiguousarray(solid, dtype=np.uint8).reshape(-1)
            assert self._solid_in.size == self.N
            sp = _ptr(self._solid_in)
        self.h = self.lib.oracle_create(C.byref(self.params), sp)

    def close(self):
        if self.h:
            self.lib.oracle_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def has_eigen(self):
        return bool(self.lib.oracle_has_eigen())

    def set_particles(self, pos, vel):
        """pos, vel: (n,3) or tuple of 3 arrays, reference coordinates."""
        a = [np.ascontiguousarray(pos[:, i] if isinstance(pos, np.ndarray) else pos[i], dtype=np.float64) for i in range(3)]
        b = [np.ascontiguousarray(vel[:, i] if isinstance(vel, np.ndarray) else vel[i], dtype=np.float64) for i in range(3)]
        n = a[0].size
        self.lib.oracle_set_particles(self.h, *[_ptr(x) for x in a + b], n)

    def get_particles(self):
        n = self.lib.oracle_num_particles(self.h)
        out = [np.empty(n, dtype=np.float64) for _ in range(6)]
        self.lib.oracle_get_particles(self.h, *[_ptr(x) for x in out])
        return np.stack(out[:3], 1), np.stack(out[3:], 1)

    def fgrid(self, which):
        return np.ctypeslib.as_array(self.lib.oracle_fgrid(self.h, which), shape=self.shape)

    def dgrid(self, which):
        return np.ctypeslib.as_array(self.lib.oracle_dgrid(self.h, which), shape=self.shape)

    @property
    def indices(self):
        return np.ctypeslib.as_array(self.lib.oracle_indices(self.h), shape=self.shape)

    @property
    def solid(self):
        return np.ctypeslib.as_array(self.lib.oracle_solid(self.h), shape=self.shape)

    @property
    def num_active(self):
        return self.lib.oracle_num_active(self.h)

    @property
    def b(self):
        n = self.num_active
        return np.ctypeslib.as_array(self.lib.oracle_b(self.h), shape=(n,)) if n else np.zeros(0)

    @property
    def p(self):
        n = self.num_active
        return np.ctypeslib.as_array(self.lib.oracle_p(self.h), shape=(n,)) if n else np.zeros(0)

    @property
    def b2(self):
        n = self.num_active
        return np.ctypeslib.as_array(self.lib.oracle_b2(self.h), shape=(n,)) if n else np.zeros(0)

    def triplets(self):
        """(rows, cols, vals) of the list setA2 builds (fluid.cc:498-540), push order."""
        r = C.POINTER(C.c_int32)(); c_ = C.POINTER(C.c_int32)(); v = C.POINTER(C.c_double)()
        n = self.lib.oracle_triplets(self.h, C.byref(r), C.byref(c_), C.byref(v))
        if n == 0:
            return np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0)
        return (np.ctypeslib.as_array(r, shape=(n,)).copy(), np.ctypeslib.as_array(c_, shape=(n,)).copy(),
                np.ctypeslib.as_array(v, shape=(n,)).copy())

    # stages
    def clear_vel(self): self.lib.oracle_clear_vel(self.h)
    def p2g(self): self.lib.oracle_p2g(self.h)
    def classify(self): return self.lib.oracle_classify(self.h)
    def save_vel_before(self): self.lib.oracle_save_vel_before(self.h)
    def rhs_div(self, dt): self.lib.oracle_rhs_div(self.h, dt)
    def assemble(self, dt): self.lib.oracle_assemble(self.h, dt)

    def solve(self):
        err = C.c_double(0)
        it = self.lib.oracle_solve(self.h, C.byref(err))
        return it, err.value

    def vel_update(self, dt_sub): self.lib.oracle_vel_update(self.h, dt_sub)
    def relax_error(self, dt): return self.lib.oracle_relax_error(self.h, dt)

    def interp_from_grid(self, first, count):
        """PointList::interpFromGrid (fluid.cc:883-895) for particles [first, first + count)."""
        self.lib.oracle_interp_from_grid(self.h, int(first), int(count))

    def flip_advect(self, dt):
        d = C.c_double(dt)
        self.lib.oracle_flip_advect(self.h, C.byref(d))
        return d.value

    def step(self, dt):
        d = C.c_double(dt)
        st = OracleStats()
        self.lib.oracle_step(self.h, C.byref(d), C.byref(st))
        return d.value, st

    def pass_history(self, cap=256):
        e = np.zeros(cap, dtype=np.float64)
        it = np.zeros(cap, dtype=np.int32)
        n = self.lib.oracle_pass_history(self.h, _ptr(e), _ptr(it), cap)
        return e[:n].copy(), it[:n].copy()


def spline(x):
    return load().oracle_spline(float(x))


def kat_solve_dense(A, b):
    A = np.ascontiguousarray(A, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    x = np.zeros_like(b)
    it = load().oracle_kat_solve_dense(b.size, _ptr(A), _ptr(b), _ptr(x))
    return x, it
