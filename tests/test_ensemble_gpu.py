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
