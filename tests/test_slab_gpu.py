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


def _worker(rank, world, name, integ, adaptive, dims, t_end, q):
    sys.path.insert(0, ROOT)
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


@pytest.mark.parametrize("integ,adaptive", [(smeb.RK101, False), (smeb.RK212, False),
                                            (smeb.RK212, True), (smeb.RK323, True),
                                            (smeb.RK435, False)])
@pytest.mark.parametrize("dims", [(40, 70, 37), (64, 48, 1 + 8)])
def test_slabs_match_single_gpu(integ, adaptive, dims):
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
    name = f"smeb200_slabtest_{os.getpid()}_{integ}_{int(adaptive)}_{dims[2]}"
    procs = [ctx.Process(target=_worker, args=(r, world, name, integ, adaptive, dims, t_end, q))
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
