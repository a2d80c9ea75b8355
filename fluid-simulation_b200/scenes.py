"""Synthetic scenes for the FLIP hot path (host side, numpy only).

Scene constants follow the reference driver: index box [-60,60]^3 with a 2-cell solid shell
(fluid.cc:1159,1264-1266), fluid cube [-20,20]^3 (fluid.cc:1176) seeded by
openvdb::tools::UniformPointScatter with 10 points per voxel volume and mt19937(0)
(fluid.cc:1348-1350; openvdb/tools/PointScatter.h:143-185: sorted uniform voxel picks, then one
Rand01 offset per axis in [c-0.5, c+0.5)), zero initial velocity, PointList::add's |p| < bound-2
filter (fluid.cc:841).  The scatter is distribution-equivalent, not bit-identical: the
reference's draws go through libstdc++'s std::uniform_*_distribution and the VDB tree's
iteration order, and both the oracle and the CUDA path consume the SAME arrays.

Benchmark scenes: "dam-break" full-tank and column fills and the scaled cube
drop, exactly `ppc` particles per fluid voxel with stratified jitter in 2x2x2 sub-cells.
All positions are in reference coordinates (index i = c + origin).
"""
from dataclasses import dataclass
from typing import Optional

import numpy as np


@dataclass
class Scene:
    name: str
    n: tuple          # (nx, ny, nz)
    origin: tuple     # (ox, oy, oz)
    wall: int
    pos: np.ndarray   # (P, 3) float64, reference coordinates
    vel: np.ndarray   # (P, 3) float64
    solid: Optional[np.ndarray] = None   # (nz, ny, nx) uint8 interior obstacles (the shell is implied by `wall`)

    @property
    def num_particles(self):
        return self.pos.shape[0]


_EPS = 1e-9


def _jitter(rng, shape):
    # open interval so that a particle never sits exactly on a voxel face (x.5 ties)
    return rng.random(shape) * (1.0 - 2 * _EPS) + _EPS


def uniform_point_scatter(lo, hi, points_per_voxel, seed):
    """UniformPointScatter restatement over the voxel box [lo, hi]^3 (inclusive, reference coords)."""
    lo = np.asarray(lo, dtype=np.int64)
    hi = np.asarray(hi, dtype=np.int64)
    dims = hi - lo + 1
    nvox = int(np.prod(dims))
    target = int(points_per_voxel) * nvox                 # Index64(ppv*dx*dy*dz)*voxelCount, PointScatter.h:150
    rng = np.random.Generator(np.random.MT19937(seed))
    ids = np.sort(rng.integers(0, nvox, size=target, dtype=np.int64))   # :158-161
    # voxel order: x outer, y, z inner (a dense box's natural VDB order is tree-dependent; equivalent in distribution)
    cz = ids % dims[2]
    cy = (ids // dims[2]) % dims[1]
    cx = ids // (dims[2] * dims[1])
    c = np.stack([cx + lo[0], cy + lo[1], cz + lo[2]], axis=1).astype(np.float64)
    pos = c - 0.5 + _jitter(rng, (target, 3))             # addPoint: coord - 0.5 + Rand01, :420-428
    return pos


def cube_drop_default(seed=0, half=20, n=121, origin=60, wall=2, ppv=10):
    """BASELINE config 1: the reference's built-in scene (fluid.cc:1159-1176,1348-1350)."""
    pos = uniform_point_scatter([-half] * 3, [half] * 3, ppv, seed)
    bound = origin
    keep = np.all(np.abs(pos) < bound - 2, axis=1)        # PointList::add, fluid.cc:841
    pos = np.ascontiguousarray(pos[keep])
    return Scene("cube_drop_default", (n, n, n), (origin,) * 3, wall, pos, np.zeros_like(pos))


def _fill_box(ilo, ihi, origin, ppc, seed):
    """Exactly ppc (=8) particles per voxel in index box [ilo, ihi] (inclusive), stratified 2x2x2."""
    assert ppc == 8, "stratified fill is defined for 8 particles per cell"
    ilo = np.asarray(ilo); ihi = np.asarray(ihi)
    xs = np.arange(ilo[0], ihi[0] + 1); ys = np.arange(ilo[1], ihi[1] + 1); zs = np.arange(ilo[2], ihi[2] + 1)
    rng = np.random.Generator(np.random.PCG64(seed))
    out = []
    # chunk over z to bound host memory
    sub = np.array([[a, b, c] for c in (0, 1) for b in (0, 1) for a in (0, 1)], dtype=np.float64)
    for z in zs:
        X, Y = np.meshgrid(xs, ys, indexing="xy")          # x fastest
        vox = np.stack([X.ravel(), Y.ravel(), np.full(X.size, z)], axis=1).astype(np.float64)
        p = vox[:, None, :] - 0.5 + (sub[None, :, :] + _jitter(rng, (vox.shape[0], 8, 3))) * 0.5
        out.append(p.reshape(-1, 3))
    pos = np.concatenate(out, axis=0) - np.asarray(origin, dtype=np.float64)[None, :]
    return np.ascontiguousarray(pos)


def dam_break(n=128, fill="full", ppc=8, seed=1234, wall=2, nz=None):
    """BASELINE configs 2/3.  fill="full": every interior voxel with y below a 2-cell air gap
    (count-matching headline fill); fill="column": x < 0.4n, y < 0.6n, full z (true dam break)."""
    nx = ny = n
    nz = n if nz is None else nz
    origin = (nx // 2, ny // 2, nz // 2)
    lo = [wall, wall, wall]
    if fill == "full":
        hi = [nx - 1 - wall, ny - 1 - wall - 2, nz - 1 - wall]
    elif fill == "column":
        hi = [int(0.4 * nx) - 1, int(0.6 * ny) - 1, nz - 1 - wall]
    else:
        raise ValueError(fill)
    pos = _fill_box(lo, hi, origin, ppc, seed)
    return Scene(f"dam_break_{fill}_{nx}x{ny}x{nz}", (nx, ny, nz), origin, wall, pos, np.zeros_like(pos))


def cube_drop(n=128, ppc=8, seed=1234, wall=2):
    """Scaled cube drop (BASELINE config 4 geometry): centred cube of side n/3."""
    origin = (n // 2,) * 3
    side = n // 3
    lo = [n // 2 - side // 2] * 3
    hi = [lo[0] + side - 1] * 3
    pos = _fill_box(lo, hi, origin, ppc, seed)
    return Scene(f"cube_drop_{n}", (n, n, n), origin, wall, pos, np.zeros_like(pos))


def small_cube_drop(n=25, half=4, ppv=10, seed=0):
    """A miniature of the default scene for fast parity tests (odd n, symmetric box)."""
    origin = n // 2
    pos = uniform_point_scatter([-half] * 3, [half] * 3, ppv, seed)
    return Scene(f"small_cube_drop_{n}", (n, n, n), (origin,) * 3, 2, pos, np.zeros_like(pos))


# ---- the reference driver's alternative scenes (commented out in fluid.cc) -------------------------
# fluid shapes: fluid.cc:1176 ("normal"), :1178-1190 ("pea"), :1192-1203 ("side"), :1205-1216 ("stable");
# obstacles inside the solid-shell loop: :1303-1315 ("two_blocks"), :1317-1329 ("three_blocks"),
# and after it :1334-1345 ("big_wall").  All in reference coordinates on the 121^3 box.
_REF_FLUIDS = {
    "normal": [((-20, -20, -20), (20, 20, 20))],
    "pea": [((-1, -1, -1), (1, 1, 1)), ((-1, 6, -1), (1, 9, 1))],
    "side": [((-57, -57, -57), (57, -40, -40))],
    "stable": [((-57, -57, -57), (57, -55, 57))],
}
_REF_OBSTACLES = {
    "two_blocks": [((-11, -58, -3), (-6, -8, 3)), ((6, -58, -3), (11, -8, 3))],
    "three_blocks": [((-11, -58, -3), (-7, -8, 3)), ((-2, -58, -3), (2, -8, 3)), ((7, -58, -3), (11, -8, 3))],
    "big_wall": [((-58, -58, -30), (58, -50, -25))],
}


def reference_scene(fluid="normal", obstacle=None, seed=0, ppv=10, n=121, origin=60, wall=2):
    """One of the reference driver's scenes: `fluid` in normal | pea | side | stable, `obstacle` in
    None | two_blocks | three_blocks | big_wall.  Particles: UniformPointScatter (10 per voxel volume)
    over every fluid box with mt19937(seed), as fluid.cc:1348-1350 does over fluidGrid."""
    boxes = _REF_FLUIDS[fluid]
    parts = [uniform_point_scatter(lo, hi, ppv, seed + 7919 * b) for b, (lo, hi) in enumerate(boxes)]
    pos = np.concatenate(parts, axis=0)
    pos = np.ascontiguousarray(pos[np.all(np.abs(pos) < origin - 2, axis=1)])       # PointList::add, fluid.cc:841
    solid = None
    if obstacle is not None:
        solid = np.zeros((n, n, n), dtype=np.uint8)                                  # (z, y, x)
        for lo, hi in _REF_OBSTACLES[obstacle]:
            solid[lo[2] + origin:hi[2] + origin + 1, lo[1] + origin:hi[1] + origin + 1, lo[0] + origin:hi[0] + origin + 1] = 1
        # particles are not seeded inside an obstacle
        c = np.rint(pos).astype(np.int64) + origin
        pos = np.ascontiguousarray(pos[solid[c[:, 2], c[:, 1], c[:, 0]] == 0])
    name = f"reference_{fluid}" + (f"_{obstacle}" if obstacle else "")
    return Scene(name, (n, n, n), (origin,) * 3, wall, pos, np.zeros_like(pos), solid)
