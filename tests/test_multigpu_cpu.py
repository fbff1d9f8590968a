"""Host-side logic of the z-slab decomposition, on CPU: the partition, the
local problems, and the node-local communicator (allreduce-max + barrier) run by
two processes that rendezvous over torch.distributed / gloo."""
import os
import sys

import numpy as np
import pytest

import spatial_model_editor_b200 as smeb
from spatial_model_editor_b200 import models as M
from spatial_model_editor_b200 import multigpu as MG

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("nz,world", [(512, 8), (7, 2), (10, 4), (5, 5), (64, 3)])
def test_slab_partition_covers_the_box(nz, world):
    ranges = [MG.slab_range(nz, r, world) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == nz
    for (a0, a1), (b0, b1) in zip(ranges, ranges[1:]):
        assert a1 == b0 and a1 > a0
    sizes = [b - a for a, b in ranges]
    assert max(sizes) - min(sizes) <= 1
    for r in range(world):
        g0, g1, lo, hi = MG.ghost_range(nz, r, world)
        assert (lo, hi) == (int(r > 0), int(r < world - 1))
        assert (g0, g1) == (ranges[r][0] - lo, ranges[r][1] + hi)


def test_local_problems_tile_the_global_state():
    p = M.box_problem(M.GRAY_SCOTT, 6, 5, 11)
    glob = p.compartments[0].conc0
    plane = 6 * 5
    owned = []
    for r in range(3):
        lp = MG.local_problem(p, r, 3, "x")
        c = lp.compartments[0]
        g0, g1, lo, hi = MG.ghost_range(11, r, 3)
        assert c.box == (6, 5, g1 - g0)
        assert np.array_equal(c.conc0, glob[g0 * plane:g1 * plane])
        assert (lp.rank, lp.n_ranks, lp.comm_name) == (r, 3, "x")
        owned.append(c.conc0[lo * plane:c.conc0.shape[0] - hi * plane])
        # make_slab supplies the same planes without the global array
        lp2 = MG.local_problem(p, r, 3, "x", make_slab=lambda a, b: glob[a * plane:b * plane])
        assert np.array_equal(lp2.compartments[0].conc0, c.conc0)
    assert np.array_equal(np.concatenate(owned), glob)
    with pytest.raises(ValueError):
        MG.local_problem(M.two_compartment_brusselator(8, 8), 0, 2, "x")


def test_slab_module_compiles_with_ghost_planes():
    """NVRTC compile of a rank's module (no device): ghost planes shrink the owned
    range but the kernels are the same templates"""
    p = M.box_problem(M.GRAY_SCOTT, 16, 16, 12)
    lp = MG.local_problem(p, 1, 3, "x")
    sim = smeb.PixelSim(lp, compile_only=True)
    assert sim.errorMessage() == ""
    assert "smeb_stage_tile" in sim.kernelSource()


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, ROOT)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import torch
        names = [f"smeb200_test_{os.getpid()}" if rank == 0 else None]
        dist.broadcast_object_list(names, src=0)
        comm = MG.Comm(names[0], rank, world)
        ok = True
        rng = np.random.default_rng(100 + rank)
        for it in range(200):
            v = rng.uniform(0, 1, size=2)
            if it % 17 == 0:
                v[1] = np.inf
            mine = comm.allreduce_max(v)
            t = torch.from_numpy(v.copy())
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ok = ok and np.array_equal(mine, t.numpy())
            if it % 50 == 0:
                comm.barrier()
        comm.barrier()
        comm.close()
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_communicator_allreduce_max_matches_gloo(world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    results = [q.get(timeout=180) for _ in range(world)]
    for pr in procs:
        pr.join(timeout=60)
    assert sorted(results) == [(r, True) for r in range(world)]
