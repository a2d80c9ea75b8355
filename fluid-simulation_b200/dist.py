"""Host side of the z-slab decomposition: one process per GPU.

torch.distributed is only the bootstrap: rank 0 asks the library for an NCCL unique id
(flipb200_comm_unique_id) and broadcasts the 128 bytes; every rank then creates its slab context
(flipb200_create_slab) and joins the library's own NCCL communicator (flipb200_comm_init_nccl).
All data-path exchanges (ghost planes, P2G ghost sums, PCG dot products, particle migration) happen
inside libflipb200.so on the context's CUDA stream -- no tensor ever crosses torch.distributed.

The reference has no distributed mode (single process, tbb::parallel_for over particles,
fluid.cc:845-1126); this module has no counterpart there.
"""
import numpy as np


def slab_split(nz, world):
    """First owned global plane of every rank (+ end) and log2 of the boundary alignment.
    Mirror of slab_split() in csrc/ctx.cu: boundaries are multiples of 2^a so that `a` multigrid
    levels coarsen inside a slab, every slab keeps >= 2 planes on level a."""
    if world == 1:
        return [0, nz], 0
    base = nz // world
    if base < 4:
        raise ValueError(f"nz={nz} is too thin for {world} slabs (need >= 4 planes per rank)")
    a = 0
    while a < 6 and (4 << a) <= base:
        a += 1
    A = 1 << a
    zs = [((r * nz) // world) & ~(A - 1) for r in range(world)] + [nz]
    if any(zs[r + 1] - zs[r] < 2 * A for r in range(world)):
        raise ValueError("slab thinner than its alignment")
    return zs, a


def balanced_split(nz, world, z_lo, z_hi, max_imbalance=1.12):
    """Slab boundaries that give every rank about the same number of the planes [z_lo, z_hi) that hold fluid (the
    pressure solve and the particle kernels only work there).  Boundaries are multiples of 2^a, slabs >= 2 * 2^a planes;
    the largest a (more distributed multigrid levels) whose busiest rank stays within max_imbalance of the ideal share
    wins.  Returns the world + 1 boundaries for flipb200_create_slab_split."""
    if world == 1:
        return [0, nz]
    z_lo, z_hi = max(0, int(z_lo)), min(nz, int(z_hi))
    ideal = (z_hi - z_lo) / world
    best = None
    for a in range(5, 0, -1):
        A = 1 << a
        zs = [0]
        for r in range(1, world):
            b = int(round((z_lo + r * ideal) / A)) * A
            b = max(b, zs[-1] + 2 * A)
            zs.append(b)
        zs.append(nz)
        if any(zs[r + 1] - zs[r] < max(2 * A, 4) for r in range(world)):
            continue
        load = max(max(0, min(zs[r + 1], z_hi) - max(zs[r], z_lo)) for r in range(world))
        if best is None or load < best[0]:
            best = (load, zs)
        if load <= max_imbalance * ideal:
            return zs
    if best is None:
        raise ValueError("no aligned split")
    return best[1]


def owner_of(pz, oz, zs):
    """Rank that owns particles with reference z coordinate pz (array): the rank whose planes hold
    the base plane of the particle's 2x2x2 support (round(p) on the side of the fractional part,
    fluid.cc:267-276 restricted to the non-zero spline weights).  Particles whose base plane lies
    outside the box belong to rank 0 (they touch no node)."""
    pz = np.asarray(pz, dtype=np.float64)
    f = np.where(pz >= 0, np.floor(pz + 0.5), np.ceil(pz - 0.5))     # C round(): half away from zero
    base = (f + oz).astype(np.int64) - ((pz - f) < 0)
    nz = zs[-1]
    owner = np.searchsorted(np.asarray(zs[1:-1]), base, side="right")
    owner[(base < 0) | (base > nz - 2)] = 0
    return owner


def broadcast_unique_id(make_id, rank, world, group=None):
    """Rank 0 calls make_id() -> 128 bytes; everyone returns the same bytes (torch.distributed,
    any backend)."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return make_id()
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    if rank == 0:
        raw = make_id()
        if len(raw) != 128:
            raise ValueError("unique id must be 128 bytes")
        t = torch.tensor(list(raw), dtype=torch.uint8, device=dev)
    else:
        t = torch.zeros(128, dtype=torch.uint8, device=dev)
    dist.broadcast(t, src=0, group=group)
    return bytes(t.cpu().tolist())


def create_slab_flip(fb, rank, world, device=None, zsplit=None, **params):
    """Slab context of this rank, connected over NCCL.  `fb` is the fluid_simulation_b200 module;
    torch.distributed must be initialised when world > 1.  zsplit: see balanced_split."""
    kw = dict(params)
    if device is not None:
        kw["device"] = device
    f = fb.Flip(rank=rank, world=world, zsplit=zsplit if world > 1 else None, **kw)
    if world > 1:
        uid = broadcast_unique_id(fb.comm_unique_id, rank, world)
        f.comm_init_nccl(uid)
    return f
