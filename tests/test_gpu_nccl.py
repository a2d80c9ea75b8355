"""z-slab decomposition over the NCCL transport: one process per GPU (needs >= 2 GPUs; skipped on a
single-GPU box, where tests/test_gpu_slab.py covers the same slab logic over the in-process
transport).  Each rank steps its slab; rank 0 also steps the whole box on one GPU and compares.
Also run with the CUDA graph of the PCG iteration disabled (plain launches of the same NCCL calls)."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, nsteps, no_graph, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    if no_graph & 1:
        os.environ["FLIPB200_NO_GRAPH"] = "1"
    if no_graph & 2:
        os.environ["FLIPB200_FUSED_HALO"] = "0"
    for p in (ROOT, os.path.join(ROOT, "tests")):
        sys.path.insert(0, p)
    import torch
    import torch.distributed as dist
    import fluid_simulation_b200 as fb
    from fluid_simulation_b200 import dist as fd, scenes
    from test_gpu_slab import moving_scene
    from helpers import rel_l2
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc = moving_scene(scenes, n=48)
        nx, ny, nz = sc.n
        ox, oy, oz = sc.origin
        f = fd.create_slab_flip(fb, rank, world, device=rank, nx=nx, ny=ny, nz=nz, ox=ox, oy=oy, oz=oz, wall=sc.wall,
                                track_ids=1)
        f.set_particles(sc.pos, sc.vel)
        f1 = fb.Flip.from_scene(sc, track_ids=1, device=0) if rank == 0 else None
        dt, dt1 = 0.1, 0.1
        log = []
        for _ in range(nsteps):
            dt, st = f.step(dt)
            rec = (st.passes, st.n_fluid, st.pcg_iters_total)
            if f1 is not None:
                dt1, s1 = f1.step(dt1)
                assert (s1.passes, s1.n_fluid) == rec[:2], (rec, s1.passes, s1.n_fluid)
                assert abs(dt - dt1) <= 1e-12 * dt1
            log.append(rec)
        p, v = f.get_particles()
        ids = f.get_particle_ids()
        assert f.halo_mode == (0 if os.environ.get("FLIPB200_FUSED_HALO") == "0" else 2), f.halo_mode
        parts = [None] * world
        dist.gather_object((ids, p, v, f.get_grid(fb.GRID_P), log), parts if rank == 0 else None, dst=0)
        if rank == 0:
            n = sc.num_particles
            pos = np.full((n, 3), np.nan); vel = np.full((n, 3), np.nan)
            for ids_r, p_r, v_r, _, log_r in parts:
                pos[ids_r] = p_r; vel[ids_r] = v_r
                assert log_r == log                       # every rank saw the same pass / iteration counts
            assert not np.isnan(pos).any() and sum(len(x[0]) for x in parts) == n
            p1, v1 = f1.get_particles()
            assert rel_l2(pos, p1) < 1e-10 and rel_l2(vel, v1) < 1e-9
            pg = np.concatenate([x[3] for x in parts], axis=0)
            assert rel_l2(pg, f1.get_grid(fb.GRID_P)) < 1e-8
        q.put((rank, "ok"))
    except BaseException as e:       # noqa: BLE001
        q.put((rank, repr(e)))
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("no_graph", [0, 1, 2])   # bit 0: no CUDA graph, bit 1: NCCL halo copies instead of fused pushes
def test_nccl_slabs_match_single(no_graph):
    import torch
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 4 if ngpu >= 4 else 2
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, world, port, 5, no_graph, q)) for r in range(world)]
    for p in ps:
        p.start()
    try:
        res = [q.get(timeout=120) for _ in range(world)]
        for p in ps:
            p.join(timeout=60)
    finally:
        for p in ps:
            if p.is_alive():
                p.kill()
    assert all(r[1] == "ok" for r in res), res
    assert all(p.exitcode == 0 for p in ps)
