"""flipb200_p2g_noalloc (include/flipb200.h): the P2G transfer as one call that allocates nothing -- its scratch is a static
__device__ array -- against the existing entry points flipb200_p2g + flipb200_classify on the same particles.

Collected last on purpose.  Called through the C ABI with ctypes; torch only owns the
caller-side device buffers.  Two calls in a row with DIFFERENT particles on the same grid: the second must not see the
first one's sums in the shared scratch.  Tolerance: the north_star's grid bar of the other GPU tests (rel-L2 <= 1e-5);
the two paths add the same fp64 terms in a different order, so the measured distance is round-off (~1e-16).
"""
import ctypes as C

import numpy as np
import pytest

from helpers import rel_l2, random_velocity_scene

pytestmark = pytest.mark.gpu

TOL_GRID = 1e-5
N, O, WALL = 40, 20, 2


def _existing(fb, pos, vel):
    f = fb.Flip(nx=N, ny=N, nz=N, ox=O, oy=O, oz=O, wall=WALL)
    f.set_particles(pos, vel)
    f.p2g()
    f.classify()
    u = np.stack([f.get_grid(fb.GRID_U + a) for a in range(3)])
    w = f.get_grid(fb.GRID_WEIGHT)
    f.close()
    return u, w


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def test_p2g_noalloc_matches_p2g_twice(fb):
    import torch
    from fluid_simulation_b200 import scenes
    lib = fb.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    nodes = N * N * N
    # caller-owned buffers, allocated ONCE before the calls
    u = torch.full((3, nodes), 7.0, dtype=torch.float64, device=dev)       # garbage on purpose: the call must clear its outputs
    w = torch.full((nodes,), 7.0, dtype=torch.float32, device=dev)
    sc_a = random_velocity_scene(scenes, n=N, half=9, seed=5, vscale=2.0)
    sc_b = random_velocity_scene(scenes, n=N, half=6, seed=11, vscale=0.5)
    sc_b.pos = sc_b.pos + np.array([3.25, -4.5, 1.75])                    # other cells than the first call's
    for sc in (sc_a, sc_b):
        pos = np.ascontiguousarray(sc.pos, dtype=np.float64)
        vel = np.ascontiguousarray(sc.vel, dtype=np.float64)
        n = pos.shape[0]
        p = torch.from_numpy(np.ascontiguousarray(pos.T)).to(dev)
        v = torch.from_numpy(np.ascontiguousarray(vel.T)).to(dev)
        torch.cuda.synchronize()
        rc = lib.flipb200_p2g_noalloc(N, N, N, O, O, O, WALL, None, _ptr(p[0]), _ptr(p[1]), _ptr(p[2]), _ptr(v[0]), _ptr(v[1]), _ptr(v[2]),
                                      C.c_int64(n), _ptr(u[0]), _ptr(u[1]), _ptr(u[2]), _ptr(w), None)
        assert rc == 0, fb.strerror(rc)
        torch.cuda.synchronize()
        u_new = u.cpu().numpy().reshape(3, N, N, N)
        w_new = w.cpu().numpy().reshape(N, N, N)
        u_old, w_old = _existing(fb, pos, vel)
        assert np.count_nonzero(w_old) > 1000
        assert np.array_equal(w_new > 0, w_old > 0)
        assert rel_l2(w_new, w_old) < TOL_GRID
        assert rel_l2(u_new, u_old) < TOL_GRID


def test_p2g_noalloc_rejects_sizes_beyond_the_static_scratch(fb):
    lib = fb.load()
    assert 128 ** 3 == fb.NOALLOC_MAX_NODES
    rc = lib.flipb200_p2g_noalloc(129, 128, 128, 64, 64, 64, 2, None, None, None, None, None, None, None, C.c_int64(0), None, None, None, None, None)
    assert rc == fb.ERR_LIMIT
    assert "limit" in fb.strerror(rc)
