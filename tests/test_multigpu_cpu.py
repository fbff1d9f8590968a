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


def _cell_problem():
    """outer box / inner box in x, y and z with a membrane between them: pairs in
    all six face directions, some straddling the slab boundaries"""
    from spatial_model_editor_b200.pde import ReactionDef
    lab = np.ones((9, 8, 7), dtype=np.int32)
    lab[2:7, 2:6, 2:5] = 2
    mem = dict(a=0, b=1, species_a=["X_a", "Y_a"], species_b=["X_b", "Y_b"],
               constants={"km": 0.3},
               reactions=[ReactionDef("flux", "km * (X_a - X_b)", {"X_a": -1.0, "X_b": 1.0})])
    p = M.labelled_problem(lab, [dict(label=1, id="outer", kin=M.BRUSSELATOR),
                                 dict(label=2, id="inner", kin=M.BRUSSELATOR)], [mem],
                           integrator=smeb.RK101)
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    p.max_timestep = 0.01
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_listed_local_problems_partition_voxels_and_pairs(world):
    p = _cell_problem()
    locs = [MG.local_problem(p, r, world, "x") for r in range(world)]
    for ci, c in enumerate(p.compartments):
        owned = []
        for r, lp in enumerate(locs):
            lc = lp.compartments[ci]
            lo, hi, first, last = lc.slab
            assert (lo == 0) == (r == 0 or lo == 0) and (r < world - 1 or hi == 0)
            owned.append(lc.conc0[lo:lc.n_voxels - hi])
            if r > 0:  # my low ghosts are the previous rank's last owned plane
                assert lo == locs[r - 1].compartments[ci].slab[3]
                assert locs[r - 1].compartments[ci].slab[1] == first
            # neighbours of owned voxels stay inside the local list
            assert lc.nn.max(initial=0) < max(lc.n_voxels, 1)
        assert np.array_equal(np.concatenate(owned), c.conc0)
    # every membrane pair is listed by the rank(s) owning its A voxel and its B voxel
    m = p.membranes[0]
    for d in range(6):
        n_a = n_b = 0
        for lp in locs:
            a, b = lp.compartments[0], lp.compartments[1]
            pr = lp.membranes[0].pairs[d].astype(np.int64).reshape(-1, 2)
            n_a += int(np.count_nonzero((pr[:, 0] >= a.slab[0])
                                        & (pr[:, 0] < a.n_voxels - a.slab[1])))
            n_b += int(np.count_nonzero((pr[:, 1] >= b.slab[0])
                                        & (pr[:, 1] < b.n_voxels - b.slab[1])))
        assert n_a == n_b == len(m.pairs[d])
    assert sum(len(x) for x in m.pairs[4:]) > 0  # there are z faces


@pytest.mark.parametrize("world", [2, 3])
def test_listed_slabs_emulated_on_cpu(world):
    """the decomposition itself, without GPUs: every rank's local problem stepped by
    the CPU oracle (forward Euler, one stage per step), ghosts refreshed from the
    neighbours after every step like the halo push does — bit-identical to the
    global run (ghost voxels are also stepped by the oracle, then overwritten)"""
    from oracle.oracle_sim import OraclePixelSim
    p = _cell_problem()
    dt, steps = p.max_timestep, 12
    ref = OraclePixelSim(p)
    ref.nextTimestep = dt
    locs = [MG.local_problem(p, r, world, "x") for r in range(world)]
    for lp in locs:
        lp.rank, lp.n_ranks, lp.comm_name = 0, 1, None
    sims = [OraclePixelSim(lp) for lp in locs]
    for _ in range(steps):
        assert ref.run(dt) == 1
        state = []
        for sim, lp in zip(sims, locs):
            assert sim.run(dt) == 1
            state.append([np.array(sim.getConcentrations(ci))
                          for ci in range(len(lp.compartments))])
        for r, lp in enumerate(locs):
            for ci, lc in enumerate(lp.compartments):
                lo, hi, first, last = lc.slab
                cur = state[r][ci]
                if lo:
                    prev, pc = state[r - 1][ci], locs[r - 1].compartments[ci]
                    end = pc.n_voxels - pc.slab[1]
                    cur[:lo] = prev[end - pc.slab[3]:end]
                if hi:
                    nxt, nc = state[r + 1][ci], locs[r + 1].compartments[ci]
                    cur[lc.n_voxels - hi:] = nxt[nc.slab[0]:nc.slab[0] + nc.slab[2]]
                sims[r].setConcentrations(ci, cur)
    for ci, c in enumerate(p.compartments):
        owned = [np.array(sims[r].getConcentrations(ci))[
            locs[r].compartments[ci].slab[0]:
            locs[r].compartments[ci].n_voxels - locs[r].compartments[ci].slab[1]]
            for r in range(world)]
        assert np.array_equal(np.concatenate(owned), ref.getConcentrations(ci))


def test_listed_slab_module_compiles():
    lp = MG.local_problem(_cell_problem(), 1, 3, "x")
    sim = smeb.PixelSim(lp, compile_only=True)
    assert sim.errorMessage() == ""
    assert "smeb_stage_memb" in sim.kernelSource()


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
            # ordered sum: every rank gets the bits of "rank 0 + rank 1 + ..."
            w = rng.uniform(-1, 1, size=3) * 10.0 ** rng.integers(-8, 8)
            gathered = [torch.zeros(3, dtype=torch.float64) for _ in range(world)]
            dist.all_gather(gathered, torch.from_numpy(w.copy()))
            expect = np.zeros(3)
            for g in gathered:
                expect = expect + g.numpy()
            ok = ok and np.array_equal(comm.allreduce_sum(w), expect)
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


def _worker_robust(rank, world, name, q):
    """reductions of more than 4 values, a NaN on one rank only,
    back-to-back communicators on ONE name, and a rank that gives up"""
    sys.path.insert(0, ROOT)
    try:
        ok = True
        for generation in range(3):  # same name three times in a row
            comm = MG.Comm(name, rank, world)
            v = np.arange(11, dtype=np.float64) * (1.0 if rank == 0 else -1.0) + rank
            mx = comm.allreduce_max(v)
            expect = np.maximum.reduce([np.arange(11) * (1.0 if r == 0 else -1.0) + r
                                        for r in range(world)])
            ok = ok and np.array_equal(mx, expect)
            sm = comm.allreduce_sum(np.full(9, 0.1 * (rank + 1)))
            acc = np.zeros(9)
            for r in range(world):
                acc = acc + np.full(9, 0.1 * (r + 1))
            ok = ok and np.array_equal(sm, acc)
            # a NaN on one rank never wins, and every rank gets the same bits
            w = np.array([np.nan if rank == generation % world else 1.0 + rank, 2.0])
            got = comm.allreduce_max(w)
            others = [1.0 + r for r in range(world) if r != generation % world]
            ok = ok and got[0] == max(others) and got[1] == 2.0
            comm.barrier()
            comm.close()
        # a rank that gives up releases the others' waits with a common error
        comm = MG.Comm(name + "_abort", rank, world)
        comm.barrier()
        if rank == world - 1:
            comm.abort()
            failed = True
        else:
            try:
                comm.barrier()  # the last rank never comes
                failed = False
            except RuntimeError:
                failed = True
        comm.close()
        q.put((rank, ok and failed))
    except Exception as e:  # noqa: BLE001
        q.put((rank, f"exception: {e!r}"))


@pytest.mark.parametrize("world", [2, 3])
def test_communicator_generations_wide_reductions_and_abort(world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    name = f"smeb200_robust_{os.getpid()}_{world}"
    procs = [ctx.Process(target=_worker_robust, args=(r, world, name, q))
             for r in range(world)]
    for pr in procs:
        pr.start()
    results = [q.get(timeout=180) for _ in range(world)]
    for pr in procs:
        pr.join(timeout=60)
    assert sorted(results) == [(r, True) for r in range(world)]


def test_balanced_cuts_deal_voxels_not_planes():
    """voxel-list slabs: plane cuts by voxel count, every rank at least one plane"""
    rng = np.random.default_rng(3)
    nz = 40
    # an off-centre blob: nothing in the lowest 12 planes, a lot around plane 25
    z = np.clip(rng.normal(25.0, 4.0, size=20000).round().astype(int), 12, nz - 1)
    for world in (2, 3, 8):
        cuts = MG.balanced_cuts([z], nz, world)
        assert cuts[0] == 0 and cuts[-1] == nz and len(cuts) == world + 1
        assert all(b > a for a, b in zip(cuts, cuts[1:]))
        counts = [int(np.count_nonzero((z >= a) & (z < b))) for a, b in zip(cuts, cuts[1:])]
        assert sum(counts) == len(z)
        assert max(counts) < 2.2 * len(z) / world  # within a plane's worth of the mean
        even = [int(np.count_nonzero((z >= MG.slab_range(nz, r, world)[0])
                                     & (z < MG.slab_range(nz, r, world)[1])))
                for r in range(world)]
        assert max(counts) <= max(even)
    # more ranks than occupied planes still gives everybody a plane
    cuts = MG.balanced_cuts([np.array([5, 5, 5, 6])], 8, 8)
    assert cuts == list(range(9))
