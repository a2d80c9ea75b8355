"""Shared helpers for the parity tests (test infrastructure)."""
import numpy as np


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    den = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / den) if den > 0 else float(np.linalg.norm(a - b))


def sort_particles(pos, vel):
    """Canonical order for comparing particle sets (the CUDA path returns cell-sorted order)."""
    idx = np.lexsort((pos[:, 2], pos[:, 1], pos[:, 0]))
    return pos[idx], vel[idx]


def match_particles(pos0, pos_a, vel_a, pos0_b):
    """Reorder set A (initial positions pos0) to the order of set B's initial positions."""
    ia = np.lexsort((pos0[:, 2], pos0[:, 1], pos0[:, 0]))
    ib = np.lexsort((pos0_b[:, 2], pos0_b[:, 1], pos0_b[:, 0]))
    inv = np.empty_like(ib); inv[ib] = np.arange(ib.size)
    order = ia[inv]
    return pos_a[order], vel_a[order]


def oracle_for_scene(oracle_mod, sc, **kw):
    nx, ny, nz = sc.n
    ox, oy, oz = sc.origin
    if getattr(sc, "solid", None) is not None and "solid" not in kw:
        kw["solid"] = sc.solid
    o = oracle_mod.Oracle(nx=nx, ny=ny, nz=nz, ox=ox, oy=oy, oz=oz, wall=sc.wall, **kw)
    o.set_particles(sc.pos, sc.vel)
    return o


def random_velocity_scene(scenes, n=25, half=5, seed=3, vscale=1.0):
    sc = scenes.small_cube_drop(n=n, half=half, seed=seed)
    rng = np.random.default_rng(seed + 100)
    sc.vel = rng.normal(0.0, vscale, size=sc.pos.shape)
    return sc
