"""flip-b200: ctypes host binding of libflipb200.so (the C ABI in include/flipb200.h).

This is the Python face of the drop-in boundary: class `Flip` mirrors the reference's step-loop
seam (PointList::P2Gtransfer / setRHS / setDiver / cg.solve / velUpdate / PointList::FLIPadvect,
fluid.cc:1378-1490) one method per C entry point.  There is no CPU fallback here: if the CUDA
library is missing or no device is present every compute call raises.
Nothing in this package imports oracle/.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FLIPB200_LIB_PATH") or os.path.join(_HERE, "libflipb200.so")      # override: A/B builds (tools/)

OK = 0
ERR_ARG, ERR_CUDA, ERR_BREAKDOWN, ERR_NOCONV, ERR_STATE, ERR_COMM, ERR_LIMIT = -1, -2, -3, -4, -5, -6, -7
NOALLOC_MAX_NODES = 2097152      # FLIPB200_NOALLOC_MAX_NODES: static scratch of flipb200_p2g_noalloc

GRID_U, GRID_V, GRID_W, GRID_UB, GRID_VB, GRID_WB, GRID_P, GRID_WEIGHT, GRID_B, GRID_FLAGS = range(10)
_GRID_DTYPE = {GRID_WEIGHT: np.float32, GRID_B: np.float32, GRID_FLAGS: np.uint8}
KERNEL_SPMV, KERNEL_PCG_ITER, KERNEL_P2G, KERNEL_G2P, KERNEL_SORT = range(5)

FB_FLUID, FB_SOLID = 64, 128


class Params(C.Structure):
    _fields_ = [
        ("nx", C.c_int32), ("ny", C.c_int32), ("nz", C.c_int32),
        ("ox", C.c_int32), ("oy", C.c_int32), ("oz", C.c_int32),
        ("wall", C.c_int32), ("relax_div", C.c_int32),
        ("dx", C.c_double), ("rho", C.c_double),
        ("gx", C.c_double), ("gy", C.c_double), ("gz", C.c_double),
        ("dt_max", C.c_double), ("relax_tol", C.c_double), ("restitution", C.c_double),
        ("flip_ratio", C.c_double), ("pcg_tol", C.c_double),
        ("pcg_max_iter", C.c_int32), ("max_passes", C.c_int32),
        ("precond", C.c_int32), ("device", C.c_int32),
        ("track_ids", C.c_int32), ("extrapolate", C.c_int32),
        ("max_particles", C.c_int64),
        ("warm_start", C.c_int32), ("mixed", C.c_int32), ("precision", C.c_int32), ("reserved", C.c_int32 * 5),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("passes", C.c_int32), ("n_fluid", C.c_int32), ("pcg_iters_total", C.c_int32),
        ("pcg_iters_last", C.c_int32),
        ("last_error", C.c_double), ("dt_out", C.c_double), ("max_speed", C.c_double),
        ("pcg_rel_resid", C.c_double),
        ("ms_sort", C.c_float), ("ms_p2g", C.c_float), ("ms_grid", C.c_float), ("ms_pcg", C.c_float),
        ("ms_g2p", C.c_float), ("ms_total", C.c_float),
        ("launches", C.c_int64),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


EXPORTS = [
    "flipb200_default_params", "flipb200_strerror", "flipb200_version", "flipb200_create",
    "flipb200_destroy", "flipb200_set_particles", "flipb200_get_particles", "flipb200_num_particles",
    "flipb200_seed_box", "flipb200_emit_box", "flipb200_step", "flipb200_step_host", "flipb200_p2g", "flipb200_classify",
    "flipb200_rhs_div", "flipb200_pcg_solve", "flipb200_vel_update", "flipb200_relax_error",
    "flipb200_g2p_advect", "flipb200_get_grid", "flipb200_set_grid", "flipb200_time_kernel",
    "flipb200_host_alloc", "flipb200_host_free",
    "flipb200_create_slab", "flipb200_slab_range", "flipb200_comm_unique_id", "flipb200_comm_init_nccl",
    "flipb200_comm_init_local", "flipb200_get_particle_ids", "flipb200_step_host_slab", "flipb200_halo_mode", "flipb200_debug_info",
    "flipb200_profile_iteration", "flipb200_create_slab_split", "flipb200_p2g_noalloc",
]

_lib = None


class FlipError(RuntimeError):
    def __init__(self, status, where):
        self.status = status
        msg = load().flipb200_strerror(status).decode() if _lib is not None else str(status)
        super().__init__(f"{where}: {msg} ({status})")


def load():
    """dlopen libflipb200.so; fails loudly if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} not built: run `python __graft_entry__.py build` "
                          "(the FLIP hot path has no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, dp, ip = C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int32)
    lib.flipb200_default_params.argtypes = [C.POINTER(Params)]
    lib.flipb200_default_params.restype = None
    lib.flipb200_strerror.argtypes = [C.c_int]
    lib.flipb200_strerror.restype = C.c_char_p
    lib.flipb200_version.restype = C.c_int
    lib.flipb200_create.argtypes = [C.POINTER(Params), vp, C.POINTER(vp)]
    lib.flipb200_destroy.argtypes = [vp]
    lib.flipb200_destroy.restype = None
    lib.flipb200_set_particles.argtypes = [vp] + [vp] * 6 + [C.c_int64]
    lib.flipb200_get_particles.argtypes = [vp] + [vp] * 6 + [C.c_int64, C.POINTER(C.c_int64)]
    lib.flipb200_num_particles.argtypes = [vp, C.POINTER(C.c_int64)]
    lib.flipb200_seed_box.argtypes = [vp, ip, ip, C.c_uint64]
    lib.flipb200_emit_box.argtypes = [vp, ip, ip, C.c_uint64]
    lib.flipb200_step.argtypes = [vp, dp, C.POINTER(Stats)]
    lib.flipb200_step_host.argtypes = [vp] + [vp] * 6 + [C.c_int64, dp, vp, C.POINTER(Stats)]
    lib.flipb200_p2g.argtypes = [vp]
    lib.flipb200_classify.argtypes = [vp, ip]
    lib.flipb200_rhs_div.argtypes = [vp, C.c_double, dp]
    lib.flipb200_pcg_solve.argtypes = [vp, C.c_double, ip, dp]
    lib.flipb200_vel_update.argtypes = [vp, C.c_double]
    lib.flipb200_relax_error.argtypes = [vp, C.c_double, dp]
    lib.flipb200_g2p_advect.argtypes = [vp, dp]
    lib.flipb200_get_grid.argtypes = [vp, C.c_int, vp, C.c_size_t]
    lib.flipb200_set_grid.argtypes = [vp, C.c_int, vp, C.c_size_t]
    lib.flipb200_time_kernel.argtypes = [vp, C.c_int, C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_int64)]
    lib.flipb200_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
    lib.flipb200_host_free.argtypes = [vp]
    lib.flipb200_create_slab.argtypes = [C.POINTER(Params), vp, C.c_int32, C.c_int32, C.POINTER(vp)]
    lib.flipb200_create_slab_split.argtypes = [C.POINTER(Params), vp, C.c_int32, C.c_int32, ip, C.POINTER(vp)]
    lib.flipb200_slab_range.argtypes = [vp, ip, ip]
    lib.flipb200_comm_unique_id.argtypes = [vp]
    lib.flipb200_comm_init_nccl.argtypes = [vp, vp]
    lib.flipb200_comm_init_local.argtypes = [vp, C.c_int32]
    lib.flipb200_get_particle_ids.argtypes = [vp, vp, C.c_int64]
    lib.flipb200_halo_mode.argtypes = [vp]
    lib.flipb200_debug_info.argtypes = [vp, ip]
    lib.flipb200_profile_iteration.argtypes = [vp, C.c_int32, C.c_int32, vp, vp, ip]
    lib.flipb200_step_host_slab.argtypes = [vp] + [vp] * 6 + [C.c_int64, C.c_int64, C.POINTER(C.c_int64), dp, vp,
                                                              C.POINTER(Stats)]
    lib.flipb200_p2g_noalloc.argtypes = [C.c_int32] * 7 + [vp] + [vp] * 6 + [C.c_int64] + [vp] * 4 + [vp]
    _lib = lib
    return lib


def strerror(status):
    return load().flipb200_strerror(int(status)).decode()


def comm_unique_id():
    """128-byte NCCL unique id (call on rank 0, broadcast to the other ranks)."""
    buf = (C.c_char * 128)()
    _check(load().flipb200_comm_unique_id(buf), "comm_unique_id")
    return bytes(buf)


def default_params(**kw):
    p = Params()
    load().flipb200_default_params(C.byref(p))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _check(rc, where):
    if rc != OK:
        raise FlipError(rc, where)


def pinned_empty(n, dtype=np.float64):
    """numpy array backed by cudaHostAlloc memory (for the e2e host-buffer path)."""
    lib = load()
    nbytes = int(n) * np.dtype(dtype).itemsize
    p = C.c_void_p()
    _check(lib.flipb200_host_alloc(C.byref(p), max(nbytes, 8)), "host_alloc")
    buf = (C.c_char * max(nbytes, 8)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(n))
    return arr, p


def pinned_free(p):
    load().flipb200_host_free(p)


class Flip:
    """One simulation context on one GPU.  Grid arrays are (nz, ny, nx) numpy arrays (x fastest)."""

    def __init__(self, params=None, solid=None, rank=0, world=1, zsplit=None, **kw):
        """world > 1: this context is z-slab `rank` of `world` (connect with comm_init_* before use);
        zsplit: world + 1 slab boundaries chosen by the caller (dist.balanced_split), default an even split."""
        self.lib = load()
        self.params = params if params is not None else default_params(**kw)
        self.rank, self.world = rank, world
        gshape = (self.params.nz, self.params.ny, self.params.nx)
        sp = None
        if solid is not None:
            self._solid = np.ascontiguousarray(solid, dtype=np.uint8).reshape(-1)
            if self._solid.size != int(np.prod(gshape)):
                raise ValueError("solid must have nx*ny*nz entries")
            sp = _ptr(self._solid)
        h = C.c_void_p()
        if world == 1:
            _check(self.lib.flipb200_create(C.byref(self.params), sp, C.byref(h)), "flipb200_create")
        elif zsplit is not None:
            zs = (C.c_int32 * (world + 1))(*[int(z) for z in zsplit])
            _check(self.lib.flipb200_create_slab_split(C.byref(self.params), sp, rank, world, zs, C.byref(h)), "flipb200_create_slab_split")
        else:
            _check(self.lib.flipb200_create_slab(C.byref(self.params), sp, rank, world, C.byref(h)), "flipb200_create_slab")
        self.h = h
        z0, z1 = C.c_int32(0), C.c_int32(0)
        _check(self.lib.flipb200_slab_range(self.h, C.byref(z0), C.byref(z1)), "slab_range")
        self.z0, self.z1 = z0.value, z1.value
        self.shape = (self.z1 - self.z0, self.params.ny, self.params.nx)     # owned planes
        self.N = int(np.prod(self.shape))

    # ---- z-slab plumbing ----
    def comm_init_local(self, group_id):
        _check(self.lib.flipb200_comm_init_local(self.h, group_id), "comm_init_local")

    def comm_init_nccl(self, unique_id):
        buf = (C.c_char * 128).from_buffer_copy(bytes(unique_id))
        _check(self.lib.flipb200_comm_init_nccl(self.h, buf), "comm_init_nccl")

    def debug_info(self):
        out = (C.c_int32 * 8)()
        _check(self.lib.flipb200_debug_info(self.h, out), "debug_info")
        keys = ["halo_mode", "iteration_graph", "mg_levels", "mg_distributed_levels", "kernels_per_iteration",
                "pushing_kernels_per_iteration", "rank", "world"]
        return dict(zip(keys, list(out)))

    @property
    def halo_mode(self):
        return self.lib.flipb200_halo_mode(self.h)

    def get_particle_ids(self):
        n = self.num_particles
        ids = np.empty(n, dtype=np.uint32)
        _check(self.lib.flipb200_get_particle_ids(self.h, _ptr(ids), n), "get_particle_ids")
        return ids

    @classmethod
    def from_scene(cls, scene, **kw):
        nx, ny, nz = scene.n
        ox, oy, oz = scene.origin
        if getattr(scene, "solid", None) is not None and "solid" not in kw:
            kw["solid"] = scene.solid
        f = cls(nx=nx, ny=ny, nz=nz, ox=ox, oy=oy, oz=oz, wall=scene.wall, **kw)
        f.set_particles(scene.pos, scene.vel)
        return f

    def close(self):
        if getattr(self, "h", None):
            self.lib.flipb200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- particles (PointList state, fluid.cc:804-813) ----
    def set_particles(self, pos, vel):
        a = [np.ascontiguousarray(pos[:, i], dtype=np.float64) for i in range(3)]
        b = [np.ascontiguousarray(vel[:, i], dtype=np.float64) for i in range(3)]
        _check(self.lib.flipb200_set_particles(self.h, *[_ptr(x) for x in a + b], a[0].size), "set_particles")

    @property
    def num_particles(self):
        n = C.c_int64(0)
        _check(self.lib.flipb200_num_particles(self.h, C.byref(n)), "num_particles")
        return n.value

    def get_particles(self):
        n = self.num_particles
        out = [np.empty(n, dtype=np.float64) for _ in range(6)]
        m = C.c_int64(0)
        _check(self.lib.flipb200_get_particles(self.h, *[_ptr(x) for x in out], n, C.byref(m)), "get_particles")
        return np.stack(out[:3], 1), np.stack(out[3:], 1)

    def seed_box(self, lo, hi, seed=1234):
        lo = (C.c_int32 * 3)(*lo)
        hi = (C.c_int32 * 3)(*hi)
        _check(self.lib.flipb200_seed_box(self.h, lo, hi, seed), "seed_box")

    def emit_box(self, lo, hi, seed=1234):
        """Re-seeding hook (fluid.cc:1374-1382, 1495-1497): seed_box + interpFromGrid (:883-895)."""
        lo = (C.c_int32 * 3)(*lo)
        hi = (C.c_int32 * 3)(*hi)
        _check(self.lib.flipb200_emit_box(self.h, lo, hi, seed), "emit_box")

    # ---- the step (fluid.cc:1378-1490) ----
    def step(self, dt):
        d = C.c_double(dt)
        st = Stats()
        _check(self.lib.flipb200_step(self.h, C.byref(d), C.byref(st)), "step")
        return d.value, st

    def step_host(self, arrays, dt, output_grid=None):
        """arrays: six 1-D float64 host arrays px,py,pz,vx,vy,vz updated in place (cell-sorted order)."""
        d = C.c_double(dt)
        st = Stats()
        n = arrays[0].size
        _check(self.lib.flipb200_step_host(self.h, *[_ptr(x) for x in arrays], n, C.byref(d),
                                           _ptr(output_grid), C.byref(st)), "step_host")
        return d.value, st

    def step_host_slab(self, arrays, n, dt, output_grid=None):
        """Slab version of step_host: arrays hold capacity >= n entries; returns (dt, stats, n_out)."""
        d = C.c_double(dt)
        st = Stats()
        m = C.c_int64(0)
        _check(self.lib.flipb200_step_host_slab(self.h, *[_ptr(x) for x in arrays], n, arrays[0].size, C.byref(m),
                                                C.byref(d), _ptr(output_grid), C.byref(st)), "step_host_slab")
        return d.value, st, m.value

    # ---- stages ----
    def p2g(self):
        _check(self.lib.flipb200_p2g(self.h), "p2g")

    def classify(self):
        n = C.c_int32(0)
        _check(self.lib.flipb200_classify(self.h, C.byref(n)), "classify")
        return n.value

    def rhs_div(self, dt):
        b2 = C.c_double(0)
        _check(self.lib.flipb200_rhs_div(self.h, dt, C.byref(b2)), "rhs_div")
        return b2.value

    def pcg_solve(self, dt):
        it = C.c_int32(0)
        rel = C.c_double(0)
        _check(self.lib.flipb200_pcg_solve(self.h, dt, C.byref(it), C.byref(rel)), "pcg_solve")
        return it.value, rel.value

    def vel_update(self, dt_sub):
        _check(self.lib.flipb200_vel_update(self.h, dt_sub), "vel_update")

    def relax_error(self, dt):
        e = C.c_double(0)
        _check(self.lib.flipb200_relax_error(self.h, dt, C.byref(e)), "relax_error")
        return e.value

    def g2p_advect(self):
        d = C.c_double(0)
        _check(self.lib.flipb200_g2p_advect(self.h, C.byref(d)), "g2p_advect")
        return d.value

    # ---- grids ----
    def get_grid(self, which):
        a = np.empty(self.shape, dtype=_GRID_DTYPE.get(which, np.float64))
        _check(self.lib.flipb200_get_grid(self.h, which, _ptr(a), a.nbytes), "get_grid")
        return a

    def set_grid(self, which, arr):
        a = np.ascontiguousarray(arr, dtype=_GRID_DTYPE.get(which, np.float64))
        if a.shape != self.shape:
            raise ValueError("grid shape")
        _check(self.lib.flipb200_set_grid(self.h, which, _ptr(a), a.nbytes), "set_grid")

    def profile_iteration(self, iters=10):
        """[(stage name, microseconds)] of one PCG iteration, in launch order (plain launches, CUDA events)."""
        cap, W = 96, 40
        names = C.create_string_buffer(cap * W)
        us = (C.c_float * cap)()
        n = C.c_int32(0)
        _check(self.lib.flipb200_profile_iteration(self.h, iters, cap, names, us, C.byref(n)), "profile_iteration")
        raw = names.raw
        return [(raw[m * W:(m + 1) * W].split(b"\0", 1)[0].decode(), float(us[m])) for m in range(n.value)]

    def time_kernel(self, which, iters):
        ms = C.c_float(0)
        units = C.c_int64(0)
        _check(self.lib.flipb200_time_kernel(self.h, which, iters, C.byref(ms), C.byref(units)), "time_kernel")
        return ms.value, units.value
