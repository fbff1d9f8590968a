"""Parameter-scan ensemble (BASELINE config 5) on the GPU: every member equals
the CPU oracle run with that parameter value, and members run concurrently
through one shared compiled module."""
import copy
import time

import numpy as np
import pytest

import spatial_model_editor_b200 as smeb
from spatial_model_editor_b200 import ensemble as EN
from oracle.oracle_sim import OraclePixelSim
from util import assert_parity, load_golden

pytestmark = pytest.mark.gpu


def test_ensemble_members_match_oracle():
    p, _, _ = load_golden("liver-simplified-scan")
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    probe = smeb.PixelSim(copy.copy(p))
    p.max_timestep = 0.5 * probe.maxStableTimestep
    probe.close()
    vals = EN.scan_values(p.params[0], 8)
    t_end = 40 * p.max_timestep
    res = EN.Ensemble(p, vals, device=0, concurrency=4).run(t_end)
    assert [r.member for r in res] == list(range(8))
    for r in res:
        assert r.error == "" and r.steps >= 40
        q = copy.copy(p)
        q.params = list(r.params)
        cpu = OraclePixelSim(q)
        assert cpu.run(t_end) == r.steps
        for ci in range(2):
            assert_parity(r.concentrations[ci], cpu.getConcentrations(ci),
                          what=f"member {r.member} comp {ci}")


def test_set_params_between_runs():
    """smeb200_set_params: the same simulator re-run with another value"""
    p, _, _ = load_golden("liver-simplified-scan", integrator=smeb.RK212)
    p.device = 0
    sim = smeb.PixelSim(p)
    c0 = [sim.getConcentrations(ci).copy() for ci in range(2)]
    sim.run(0.001)
    a = sim.getConcentrations(0).copy()
    for ci in range(2):
        sim.setConcentrations(ci, c0[ci])
    sim.nextTimestep = 1e-7
    sim.setParams([p.params[0] * 5.0])
    sim.run(0.001)
    b = sim.getConcentrations(0).copy()
    q = copy.copy(p)
    q.params = [p.params[0] * 5.0]
    ref = smeb.PixelSim(q)
    ref.run(0.001)
    assert np.array_equal(b, ref.getConcentrations(0))
    assert not np.array_equal(a, b)


def test_module_cache_makes_members_cheap():
    p, _, _ = load_golden("liver-simplified-scan")
    p.device = 0
    t0 = time.perf_counter()
    smeb.PixelSim(p).close()
    first = time.perf_counter() - t0
    p.params = [1.234]
    t0 = time.perf_counter()
    smeb.PixelSim(p).close()
    second = time.perf_counter() - t0
    assert second < 0.5 * first or second < 0.05


def _scan_problem(integrator=None):
    p, _, _ = load_golden("liver-simplified-scan", integrator=integrator)
    p.device = 0
    return p


@pytest.mark.parametrize("integ", [smeb.RK101, smeb.RK212, smeb.RK323, smeb.RK435])
def test_batched_members_match_oracle_fixed_step(integ):
    """smeb200_ensemble_*: members as gridDim.y of one launch per stage; every
    member equals the CPU oracle run with its parameter value (<= 1e-12)"""
    p = _scan_problem(integ)
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    probe = smeb.PixelSim(copy.copy(p))
    p.max_timestep = 0.5 * probe.maxStableTimestep
    probe.close()
    vals = EN.scan_values(p.params[0], 6)
    t_end = 25 * p.max_timestep
    ens = EN.BatchedEnsemble(p, vals, device=0)
    assert ens.errorMessage() == ""
    res = ens.run(t_end)
    assert ens.launchCount() > 0
    for r in res:
        assert r.error == "" and r.steps >= 25
        q = copy.copy(p)
        q.params = list(r.params)
        cpu = OraclePixelSim(q)
        assert cpu.run(t_end) == r.steps
        for ci in range(2):
            assert_parity(r.concentrations[ci], cpu.getConcentrations(ci),
                          what=f"member {r.member} comp {ci}")
    ens.close()


def test_batched_members_adaptive_take_their_own_steps():
    """adaptive RK212 (the model's settings): every member has its own controller —
    the same accepted / discarded steps and the same state as that member simulated
    alone through smeb200_run (bit-identical: same kernels' arithmetic, same
    controller formulas), members finishing at different times"""
    p = _scan_problem(smeb.RK212)
    vals = EN.scan_values(p.params[0], 8)
    t_end = 0.05
    ens = EN.BatchedEnsemble(p, vals, device=0)
    res = ens.run(t_end)
    counts = set()
    for r in res:
        q = copy.copy(p)
        q.params = list(r.params)
        one = smeb.PixelSim(q)
        n = one.run(t_end)
        assert (n, one.discardedSteps) == (r.steps, r.discarded)
        counts.add((n, one.discardedSteps))
        for ci in range(2):
            assert np.array_equal(r.concentrations[ci], one.getConcentrations(ci))
        one.close()
    assert ens.batchedSteps() >= max(c[0] + c[1] for c in counts)
    # far fewer launches than members x steps x kernels
    assert ens.launchCount() <= 8 * ens.batchedSteps()
    ens.close()


def test_batched_ensemble_continues_and_restarts():
    """a second run continues from the members' states and step sizes; a member can
    be restarted from a given state"""
    p = _scan_problem(smeb.RK212)
    vals = EN.scan_values(p.params[0], 3)
    ens = EN.BatchedEnsemble(p, vals, device=0)
    a = ens.run(0.01)
    b = ens.run(0.01)
    for k, (ra, rb) in enumerate(zip(a, b)):
        q = copy.copy(p)
        q.params = list(ra.params)
        one = smeb.PixelSim(q)
        assert one.run(0.01) == ra.steps
        assert one.run(0.01) == rb.steps
        assert np.array_equal(rb.concentrations[0], one.getConcentrations(0))
        one.close()
    ens.close()
