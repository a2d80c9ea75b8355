"""CPU tests of the multi-GPU host logic: slab split, particle ownership, and the
torch.distributed bootstrap of the library's NCCL unique id over `gloo` with world_size 2.  No
compute call is made (there is no GPU here and the library has no CPU fallback)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _dist():
    import fluid_simulation_b200  # noqa: F401  (registers the package)
    from fluid_simulation_b200 import dist
    return dist


@pytest.mark.parametrize("nz,world", [(121, 2), (128, 2), (256, 4), (512, 8), (2048, 8), (64, 8), (48, 2), (32, 4)])
def test_slab_split_properties(nz, world):
    d = _dist()
    zs, a = d.slab_split(nz, world)
    assert zs[0] == 0 and zs[-1] == nz and len(zs) == world + 1
    A = 1 << a
    for r in range(world):
        assert zs[r] % A == 0                       # boundaries aligned: `a` levels coarsen inside a slab
        assert zs[r + 1] - zs[r] >= 2 * A           # >= 2 planes on the coarsest distributed level
    assert a >= 1
    sizes = [zs[r + 1] - zs[r] for r in range(world)]
    assert max(sizes) - min(sizes) <= 2 * A         # balanced up to the alignment


@pytest.mark.parametrize("nz,world,z_lo,z_hi", [(512, 8, 170, 343), (512, 4, 170, 343), (512, 2, 170, 343), (256, 4, 0, 256),
                                                 (256, 8, 100, 130), (128, 2, 2, 20), (2048, 8, 2, 2045)])
def test_balanced_split_properties(nz, world, z_lo, z_hi):
    """dist.balanced_split (boundaries for flipb200_create_slab_split): covers [0, nz], boundaries aligned to 2^a with
    slabs of at least two coarse planes (what fb_create's slab_split_given accepts), and the busiest rank holds no more
    fluid planes than an even split gives its busiest rank."""
    d = _dist()
    zs = d.balanced_split(nz, world, z_lo, z_hi)
    assert zs[0] == 0 and zs[-1] == nz and len(zs) == world + 1 and all(zs[r + 1] > zs[r] for r in range(world))
    a = min((b & -b).bit_length() - 1 for b in zs[1:-1])          # common alignment of the inner boundaries
    assert a >= 1
    assert all(zs[r + 1] - zs[r] >= max(2 << 1, 4) for r in range(world))
    load = lambda z: max(max(0, min(z[r + 1], z_hi) - max(z[r], z_lo)) for r in range(world))
    even, _ = d.slab_split(nz, world)
    assert load(zs) <= load(even)
    assert d.balanced_split(nz, 1, z_lo, z_hi) == [0, nz]


def test_slab_split_rejects_thin():
    d = _dist()
    with pytest.raises(ValueError):
        d.slab_split(12, 4)
    assert d.slab_split(77, 1) == ([0, 77], 0)


def test_owner_partition():
    """Every particle has exactly one owner; particles on a slab boundary follow the base plane."""
    d = _dist()
    nz, world, oz = 64, 4, 32
    zs, _ = d.slab_split(nz, world)
    rng = np.random.default_rng(0)
    pz = rng.uniform(-oz - 3, nz - oz + 3, size=20000)
    own = d.owner_of(pz, oz, zs)
    assert own.min() >= 0 and own.max() < world
    # hand cases around the boundary plane 16 (reference coordinate -16): base plane = 15 below it
    assert d.owner_of(np.array([-16.2]), oz, zs)[0] == 0      # round -> -16, frac < 0 -> base 15
    assert d.owner_of(np.array([-15.8]), oz, zs)[0] == 1      # round -> -16, frac > 0 -> base 16
    assert d.owner_of(np.array([-16.5 + 1e-9]), oz, zs)[0] == 0
    assert d.owner_of(np.array([1e6]), oz, zs)[0] == 0        # outside the box: rank 0 keeps strays


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import fluid_simulation_b200  # noqa: F401
    from fluid_simulation_b200 import dist as fd
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        calls = []

        def make_id():
            calls.append(rank)
            return bytes((7 * i + 3) % 256 for i in range(128))

        uid = fd.broadcast_unique_id(make_id, rank, world)
        zs, a = fd.slab_split(256, world)
        q.put((rank, uid, calls, zs[rank], zs[rank + 1]))
    finally:
        dist.destroy_process_group()


def test_unique_id_broadcast_gloo_world2():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(2))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    expect = bytes((7 * i + 3) % 256 for i in range(128))
    assert got[0][1] == expect and got[1][1] == expect
    assert got[0][2] == [0] and got[1][2] == []       # only rank 0 generates the id
    assert (got[0][3], got[0][4], got[1][3], got[1][4]) == (0, 128, 128, 256)
