"""z-slab decomposition on >= 2 GPUs of one box (one process per GPU) against
the single-GPU run of the same problem: bit-identical concentrations and
identical step sequences.  Skipped on a single-GPU box."""
import os
import sys

import numpy as np
import pytest

import spatial_model_editor_b200 as smeb
from spatial_model_editor_b200 import models as M

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _problem(integ, adaptive, dims):
    p = M.box_problem(M.GRAY_SCOTT, *dims, integrator=integ, voxel_size=(1.0, 2.0, 1.5))
    if adaptive:
        p.max_err_rel = 0.005
    else:
        p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
        p.max_timestep = 0.01
    return p


def _worker(rank, world, name, integ, adaptive, dims, t_end, q, fused=False):
    sys.path.insert(0, ROOT)
    if fused:  # halo exchange inside the tile kernel (opt-in)
        os.environ["SMEB200_FUSED_HALO"] = "1"
    from spatial_model_editor_b200.multigpu import SlabPixelSim
    try:
        p = _problem(integ, adaptive, dims)
        p.device = rank
        sim = SlabPixelSim(p, rank, world, comm_name=name)
        err0 = sim.errorMessage()
        n1 = sim.run(t_end / 2)
        n2 = sim.run(t_end / 2)
        q.put((rank, err0 + sim.errorMessage(), n1 + n2, sim.discardedSteps,
               sim.getConcentrations(0).copy(), sim.getDcdt(0).copy(), sim.haloExchanges()))
        sim.close()
    except Exception as e:  # noqa: BLE001
        q.put((rank, f"exception: {e!r}", 0, 0, None, None, 0))


@pytest.mark.parametrize("fused", [False, True], ids=["halo-stream", "in-kernel"])
@pytest.mark.parametrize("integ,adaptive", [(smeb.RK101, False), (smeb.RK212, False),
                                            (smeb.RK212, True), (smeb.RK323, True),
                                            (smeb.RK435, False)])
@pytest.mark.parametrize("dims", [(40, 70, 37), (64, 48, 1 + 8)])
def test_slabs_match_single_gpu(integ, adaptive, dims, fused):
    world = min(smeb.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    t_end = 0.2
    p = _problem(integ, adaptive, dims)
    p.device = 0
    ref = smeb.PixelSim(p)
    assert ref.errorMessage() == ""
    n_ref = ref.run(t_end / 2) + ref.run(t_end / 2)
    c_ref, d_ref = ref.getConcentrations(0).copy(), ref.getDcdt(0).copy()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    name = f"smeb200_slabtest_{os.getpid()}_{integ}_{int(adaptive)}_{dims[2]}_{int(fused)}"
    procs = [ctx.Process(target=_worker,
                         args=(r, world, name, integ, adaptive, dims, t_end, q, fused))
             for r in range(world)]
    for pr in procs:
        pr.start()
    res = sorted((q.get(timeout=300) for _ in range(world)), key=lambda r: r[0])
    for pr in procs:
        pr.join(timeout=60)
    for r in res:
        assert r[1] == "", r[1]
        assert r[2] == n_ref and r[3] == ref.discardedSteps
        assert r[6] > 0
    assert np.array_equal(np.concatenate([r[4] for r in res]), c_ref)
    assert np.array_equal(np.concatenate([r[5] for r in res]), d_ref)


# ---- voxel-list compartments and membranes cut at plane boundaries ----------------
def _listed_problem(kind, integ, adaptive):
    """kind "blob": the gray-scott-3d example (one irregular compartment);
    kind "cell": outer box / inner box in all three directions, brusselator kinetics
    in both, coupled by a membrane with faces in x, y AND z (pairs straddle slabs)"""
    from spatial_model_editor_b200.pde import ReactionDef
    if kind == "blob":
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from util import load_golden
        p, _, _ = load_golden("gray-scott-3d", integrator=integ)
    else:
        lab = np.ones((14, 20, 18), dtype=np.int32)
        lab[3:11, 5:15, 4:13] = 2
        mem = dict(a=0, b=1, species_a=["X_a", "Y_a"], species_b=["X_b", "Y_b"],
                   constants={"km": 0.3},
                   reactions=[ReactionDef("flux", "km * (X_a - X_b)",
                                          {"X_a": -1.0, "X_b": 1.0})])
        p = M.labelled_problem(lab, [dict(label=1, id="outer", kin=M.BRUSSELATOR),
                                     dict(label=2, id="inner", kin=M.BRUSSELATOR)], [mem],
                               voxel_size=(1.0, 1.2, 0.8), integrator=integ)
    if adaptive:
        p.max_err_rel = 0.005
    else:
        p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
        p.max_timestep = 0.01
    return p


def _listed_worker(rank, world, name, kind, integ, adaptive, t_end, q):
    sys.path.insert(0, ROOT)
    from spatial_model_editor_b200.multigpu import SlabPixelSim
    try:
        p = _listed_problem(kind, integ, adaptive)
        p.device = rank
        sim = SlabPixelSim(p, rank, world, comm_name=name)
        err0 = sim.errorMessage()
        n = sim.run(t_end / 2) + sim.run(t_end / 2)
        nc = len(p.compartments)
        q.put((rank, err0 + sim.errorMessage(), n, sim.discardedSteps,
               [sim.getConcentrations(i).copy() for i in range(nc)],
               [sim.getDcdt(i).copy() for i in range(nc)], sim.haloExchanges(),
               [sim.getAvgMinMax(i) for i in range(nc)]))
        sim.close()
    except Exception as e:  # noqa: BLE001
        q.put((rank, f"exception: {e!r}", 0, 0, None, None, 0, None))


@pytest.mark.parametrize("kind", ["blob", "cell"])
@pytest.mark.parametrize("integ,adaptive", [(smeb.RK101, False), (smeb.RK212, True),
                                            (smeb.RK323, False), (smeb.RK435, True)])
def test_listed_slabs_match_single_gpu(kind, integ, adaptive):
    world = min(smeb.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    t_end = 0.2
    p = _listed_problem(kind, integ, adaptive)
    p.device = 0
    ref = smeb.PixelSim(p)
    assert ref.errorMessage() == ""
    n_ref = ref.run(t_end / 2) + ref.run(t_end / 2)
    nc = len(p.compartments)
    c_ref = [ref.getConcentrations(i).copy() for i in range(nc)]
    d_ref = [ref.getDcdt(i).copy() for i in range(nc)]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    name = f"smeb200_listtest_{os.getpid()}_{kind}_{integ}_{int(adaptive)}"
    procs = [ctx.Process(target=_listed_worker,
                         args=(r, world, name, kind, integ, adaptive, t_end, q))
             for r in range(world)]
    for pr in procs:
        pr.start()
    res = sorted((q.get(timeout=300) for _ in range(world)), key=lambda r: r[0])
    for pr in procs:
        pr.join(timeout=60)
    for r in res:
        assert r[1] == "", r[1]
        assert r[2] == n_ref and r[3] == ref.discardedSteps
        assert r[6] > 0
    for i in range(nc):
        assert np.array_equal(np.concatenate([r[4][i] for r in res]), c_ref[i])
        assert np.array_equal(np.concatenate([r[5][i] for r in res]), d_ref[i])
        # per-rank device statistics cover the owned voxels: min / max combine exactly
        mn = min(r[7][i][:, 1].min() for r in res if len(r[4][i]))
        assert mn == c_ref[i][:, :].min(axis=0).min()


# ---- non-uniform (per-voxel) diffusion and cross-diffusion in slabs ----------------
def _diffusion_problem(feature, geom, integ):
    """feature "nonuniform" / "cross" / "cross-dynamic" (RK101 whose step follows the
    state-dependent forward-Euler bound, allreduced over the ranks every step);
    geom "box" or "list" (the 3-D single-compartment-diffusion example)"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from util import load_golden
    rng = np.random.default_rng(11)
    if geom == "box":
        p = M.box_problem(M.GRAY_SCOTT, 20, 24, 15, integrator=integ,
                          voxel_size=(1.0, 1.5, 2.0))
        p.voxel_size = (1.0, 1.5, 2.0)
        D = np.array(M.GRAY_SCOTT["D"])
    else:
        p, meta, _ = load_golden("single-compartment-diffusion-3d", integrator=integ)
        p.voxel_size = tuple(meta["voxel_size"])
        D = np.array(meta["compartments"][0]["diffusion_constants"])
    c = p.compartments[0]
    if feature == "nonuniform":
        c.diffusion_array = np.ascontiguousarray(
            D[None, :] * rng.uniform(0.0, 1.5, size=(c.n_voxels, c.n_species)))
    else:
        c.diffusion_raw = D
        k = "(6.0 * sme_var_0)" if feature == "cross-dynamic" else "(0.3 + 0.5 * sme_var_0)"
        c.cross_diffusion = [(1, 0, k), (0, 1, "(0.05 * sme_var_1)")]
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    if feature != "cross-dynamic":
        p.max_timestep = 0.01
    return p


def _diffusion_worker(rank, world, name, feature, geom, integ, t_end, q):
    sys.path.insert(0, ROOT)
    from spatial_model_editor_b200.multigpu import SlabPixelSim
    try:
        p = _diffusion_problem(feature, geom, integ)
        p.device = rank
        sim = SlabPixelSim(p, rank, world, comm_name=name)
        err0 = sim.errorMessage()
        dt0 = sim.maxStableTimestep
        n = sim.run(t_end / 2) + sim.run(t_end / 2)
        q.put((rank, err0 + sim.errorMessage(), n, dt0, sim.maxStableTimestep,
               sim.getConcentrations(0).copy(), sim.getDcdt(0).copy()))
        sim.close()
    except Exception as e:  # noqa: BLE001
        q.put((rank, f"exception: {e!r}", 0, 0.0, 0.0, None, None))


@pytest.mark.parametrize("geom", ["box", "list"])
@pytest.mark.parametrize("feature,integ", [("nonuniform", smeb.RK212), ("nonuniform", smeb.RK101),
                                           ("cross", smeb.RK323), ("cross-dynamic", smeb.RK101)])
def test_diffusion_features_in_slabs_match_single_gpu(feature, integ, geom):
    world = min(smeb.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    t_end = 1.0 if feature == "cross-dynamic" else 0.1
    p = _diffusion_problem(feature, geom, integ)
    p.device = 0
    ref = smeb.PixelSim(p)
    assert ref.errorMessage() == ""
    dt0 = ref.maxStableTimestep
    n_ref = ref.run(t_end / 2) + ref.run(t_end / 2)
    assert n_ref >= 10
    c_ref, d_ref = ref.getConcentrations(0).copy(), ref.getDcdt(0).copy()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    name = f"smeb200_difftest_{os.getpid()}_{feature}_{geom}_{integ}"
    procs = [ctx.Process(target=_diffusion_worker,
                         args=(r, world, name, feature, geom, integ, t_end, q))
             for r in range(world)]
    for pr in procs:
        pr.start()
    res = sorted((q.get(timeout=300) for _ in range(world)), key=lambda r: r[0])
    for pr in procs:
        pr.join(timeout=60)
    for r in res:
        assert r[1] == "", r[1]
        assert r[2] == n_ref
        # the forward-Euler bound is the global one on every rank
        assert r[3] == dt0 and r[4] == ref.maxStableTimestep
    assert np.array_equal(np.concatenate([r[5] for r in res]), c_ref)
    assert np.array_equal(np.concatenate([r[6] for r in res]), d_ref)


# ---- non-spatial species in slabs: the spatial mean is summed over the ranks -------
def _non_spatial_problem(geom, integ):
    """geom "box": gray-scott with V non-spatial; "cell": the two-compartment cell with
    the membrane, Y non-spatial in both compartments (SimCompartment::
    spatiallyAverageDcdt, pixelsim_impl.cpp:150-163)"""
    if geom == "box":
        p = M.box_problem(M.GRAY_SCOTT, 24, 20, 13, integrator=integ,
                          voxel_size=(1.0, 1.5, 2.0))
        p.compartments[0].non_spatial = [1]
    else:
        p = _listed_problem("cell", integ, False)
        for c in p.compartments:
            c.non_spatial = [1]
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    p.max_timestep = 0.01
    return p


def _non_spatial_worker(rank, world, name, geom, integ, t_end, q):
    sys.path.insert(0, ROOT)
    from spatial_model_editor_b200.multigpu import SlabPixelSim
    try:
        p = _non_spatial_problem(geom, integ)
        p.device = rank
        sim = SlabPixelSim(p, rank, world, comm_name=name)
        err0 = sim.errorMessage()
        n = sim.run(t_end / 2) + sim.run(t_end / 2)
        nc = len(p.compartments)
        q.put((rank, err0 + sim.errorMessage(), n,
               [sim.getConcentrations(i).copy() for i in range(nc)],
               [sim.getDcdt(i).copy() for i in range(nc)]))
        sim.close()
    except Exception as e:  # noqa: BLE001
        q.put((rank, f"exception: {e!r}", 0, None, None))


@pytest.mark.parametrize("geom", ["box", "cell"])
@pytest.mark.parametrize("integ", [smeb.RK101, smeb.RK212, smeb.RK435])
def test_non_spatial_species_in_slabs(geom, integ):
    """each rank sums dc/dt of the non-spatial species over its own voxels; the sums
    are added in rank order and divided by the compartment's voxel count.  The
    association differs from the single-GPU tree, so the bar is rounding (1e-10
    relative), not bit identity; the non-spatial species stays uniform over all
    slabs"""
    world = min(smeb.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from util import assert_parity
    t_end = 0.1
    p = _non_spatial_problem(geom, integ)
    p.device = 0
    ref = smeb.PixelSim(p)
    assert ref.errorMessage() == ""
    n_ref = ref.run(t_end / 2) + ref.run(t_end / 2)
    nc = len(p.compartments)
    c_ref = [ref.getConcentrations(i).copy() for i in range(nc)]
    d_ref = [ref.getDcdt(i).copy() for i in range(nc)]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    name = f"smeb200_nstest_{os.getpid()}_{geom}_{integ}"
    procs = [ctx.Process(target=_non_spatial_worker, args=(r, world, name, geom, integ, t_end, q))
             for r in range(world)]
    for pr in procs:
        pr.start()
    res = sorted((q.get(timeout=300) for _ in range(world)), key=lambda r: r[0])
    for pr in procs:
        pr.join(timeout=60)
    for r in res:
        assert r[1] == "", r[1]
        assert r[2] == n_ref
    for i in range(nc):
        c = np.concatenate([r[3][i] for r in res])
        d = np.concatenate([r[4][i] for r in res])
        assert_parity(c, c_ref[i], rel=1e-10, what=f"{geom} conc[{i}]")
        assert_parity(d, d_ref[i], rel=1e-10, what=f"{geom} dcdt[{i}]")
        # one value of dc/dt for the non-spatial species across all slabs
        assert np.ptp(d[:, 1]) == 0.0


# ---- a stop decided on ONE rank is a decision of all ranks ------------------------
def _stop_worker(rank, world, name, mode, q):
    """mode "timeout": every rank passes a deadline (the clocks differ);
    "callback": only the last rank's callback asks to stop, after 5 steps;
    "flag": only rank 0 gets setStopRequested, from another thread, mid-run"""
    sys.path.insert(0, ROOT)
    import threading
    import time
    from spatial_model_editor_b200.multigpu import SlabPixelSim
    try:
        p = _problem(smeb.RK212, False, (64, 64, 16 * world))
        p.device = rank
        sim = SlabPixelSim(p, rank, world, comm_name=name)
        err0 = sim.errorMessage()
        sim.nextTimestep = 0.01
        t0 = time.perf_counter()
        if mode == "timeout":
            n = sim.run(1000.0, 300.0)
        elif mode == "callback":
            calls = [0]

            def cb():
                calls[0] += 1
                return rank == world - 1 and calls[0] >= 5

            n = sim.run(1000.0, -1.0, cb)
        else:
            if rank == 0:
                threading.Timer(0.3, lambda: sim.setStopRequested(True)).start()
            n = sim.run(1000.0)
        el = time.perf_counter() - t0
        msg = sim.errorMessage()
        # the ranks are still in step: the states of the slabs fit together
        c = sim.getConcentrations(0).copy()
        q.put((rank, err0, msg, n, el, c))
        sim.close()
    except Exception as e:  # noqa: BLE001
        q.put((rank, f"exception: {e!r}", "", 0, 0.0, None))


@pytest.mark.parametrize("mode", ["timeout", "callback", "flag"])
def test_stop_is_collective_in_slabs(mode):
    world = min(smeb.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    name = f"smeb200_stoptest_{os.getpid()}_{mode}"
    procs = [ctx.Process(target=_stop_worker, args=(r, world, name, mode, q))
             for r in range(world)]
    for pr in procs:
        pr.start()
    res = sorted((q.get(timeout=300) for _ in range(world)), key=lambda r: r[0])
    for pr in procs:
        pr.join(timeout=60)
    for r in res:
        assert r[1] == "", r[1]
        assert r[2] == "Simulation stopped early"
        assert r[4] < 60.0  # nobody waited for a neighbour that had stopped
    steps = {r[3] for r in res}
    assert len(steps) == 1 and steps.pop() >= (5 if mode == "callback" else 1)
    # the same number of steps on one GPU gives the same state
    p = _problem(smeb.RK212, False, (64, 64, 16 * world))
    p.device = 0
    ref = smeb.PixelSim(p)
    ref.nextTimestep = 0.01
    n = res[0][3]
    calls = [0]

    def after_n():
        calls[0] += 1
        return calls[0] >= n

    assert ref.run(1000.0, -1.0, after_n) == n  # the same sequence of step sizes
    assert np.array_equal(np.concatenate([r[5] for r in res]), ref.getConcentrations(0))


# ---- zero-storage (algebraic) species: the relaxation's iterates cross the slabs ----
def _zero_storage_problem(geom, integ):
    from spatial_model_editor_b200.pde import ReactionDef  # noqa: F401
    if geom == "box":
        p = M.box_problem(M.ABTOC, 24, 32, 12, integrator=integ)
        zs = 2
    else:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from util import load_golden
        p, _, _ = load_golden("single-compartment-diffusion-3d", integrator=integ)
        zs = 0
    c = p.compartments[0]
    inv = np.ones(c.n_species)
    inv[zs] = 0.0
    c.inv_storage = inv
    p.max_err_abs, p.max_err_rel = smeb.DBL_MAX, 0.01  # tolerances of the relaxation
    p.max_timestep = 1e-3
    return p


def _zero_storage_worker(rank, world, name, geom, integ, q):
    sys.path.insert(0, ROOT)
    from spatial_model_editor_b200.multigpu import SlabPixelSim
    try:
        p = _zero_storage_problem(geom, integ)
        p.device = rank
        sim = SlabPixelSim(p, rank, world, comm_name=name)
        err0 = sim.errorMessage()
        if integ != smeb.RK101:
            sim.nextTimestep = 1e-3
        n = sim.run(3e-3) + sim.run(2e-3)
        q.put((rank, err0 + sim.errorMessage(), n, sim.getConcentrations(0).copy(),
               sim.haloExchanges()))
        sim.close()
    except Exception as e:  # noqa: BLE001
        q.put((rank, f"exception: {e!r}", 0, None, 0))


@pytest.mark.parametrize("geom", ["box", "list"])
@pytest.mark.parametrize("integ", [smeb.RK101, smeb.RK212, smeb.RK323])
def test_zero_storage_species_in_slabs_match_single_gpu(geom, integ):
    """PixelSim::solveZeroStorageConstraints (pixelsim.cpp:23-90) across z-slabs: the
    pseudo-time iterates exchange their boundary planes, the residual and error norms
    are max-reduced, every rank takes the single-GPU iteration — bit-identical state"""
    world = min(smeb.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    p = _zero_storage_problem(geom, integ)
    p.device = 0
    ref = smeb.PixelSim(p)
    assert ref.errorMessage() == ""
    if integ != smeb.RK101:
        ref.nextTimestep = 1e-3
    n_ref = ref.run(3e-3) + ref.run(2e-3)
    assert n_ref >= 5 and ref.errorMessage() == ""
    c_ref = ref.getConcentrations(0).copy()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    name = f"smeb200_zstest_{os.getpid()}_{geom}_{integ}"
    procs = [ctx.Process(target=_zero_storage_worker, args=(r, world, name, geom, integ, q))
             for r in range(world)]
    for pr in procs:
        pr.start()
    res = sorted((q.get(timeout=300) for _ in range(world)), key=lambda r: r[0])
    for pr in procs:
        pr.join(timeout=60)
    for r in res:
        assert r[1] == "", r[1]
        assert r[2] == n_ref
        assert r[4] >= 2 * n_ref  # the relaxation exchanged halos of its own
    assert np.array_equal(np.concatenate([r[3] for r in res]), c_ref)
