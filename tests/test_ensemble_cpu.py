"""host logic of the parameter-scan ensembles (BASELINE config 5) on CPU"""
import numpy as np
import pytest

import spatial_model_editor_b200 as smeb
from spatial_model_editor_b200 import ensemble as EN
from oracle.oracle_sim import OraclePixelSim
from util import load_golden


def test_scan_values_and_sharding():
    v = EN.scan_values(0.9, 64)
    assert v.shape == (64,) and np.all(v >= 0.09) and np.all(v <= 9.0)
    assert np.array_equal(v, EN.scan_values(0.9, 64))
    owned = [EN.members_of_rank(64, r, 8) for r in range(8)]
    assert sorted(sum(owned, [])) == list(range(64))
    assert all(len(o) == 8 for o in owned)
    assert EN.members_of_rank(5, 3, 8) == [3] and EN.members_of_rank(5, 7, 8) == []


def test_runtime_parameter_module_is_shared():
    """all members generate the same module source: one NVRTC compile"""
    p, meta, _ = load_golden("liver-simplified-scan")
    assert len(p.params) == 1 and abs(p.params[0] - 4.97223415385606) < 1e-12
    srcs = set()
    for v in (0.09, 0.9, 9.0):
        p.params = [v]
        sim = smeb.PixelSim(p, compile_only=True)
        assert sim.errorMessage() == ""
        srcs.add(sim.kernelSource())
    assert len(srcs) == 1 and "sme_param_0" in srcs.pop()


def test_ensemble_driver_with_the_oracle_as_backend():
    """the driver's sharding / threading / result order, with the CPU oracle
    standing in for the simulator; a runtime parameter equals inlining its value"""
    p, _, _ = load_golden("liver-simplified-scan")
    meta_base = p.params[0]
    vals = EN.scan_values(p.params[0], 6)
    res = []
    for rank in range(2):
        ens = EN.Ensemble(p, vals, rank=rank, world=2, concurrency=2,
                          sim_factory=OraclePixelSim)
        res += ens.run(0.002)
    res.sort(key=lambda r: r.member)
    assert [r.member for r in res] == list(range(6))
    assert all(r.error == "" and r.steps > 0 for r in res)
    base, _, _ = load_golden("liver-simplified")
    k = [i for i, e in enumerate(p.compartments[0].expressions) if "sme_param_0" in e]
    assert k
    # member with the base value reproduces the fixture with the constant inlined
    p.params = [meta_base]
    a, b = OraclePixelSim(p), OraclePixelSim(base)
    assert a.run(0.002) == b.run(0.002)
    assert np.array_equal(a.getConcentrations(0), b.getConcentrations(0))
    # different parameter values give different trajectories
    assert not np.array_equal(res[0].concentrations[0], res[1].concentrations[0])


def test_batched_ensemble_without_a_device_carries_the_error():
    """smeb200_ensemble_create never throws: without a GPU the object carries the
    reference-style "no device / driver" message and nothing is computed on the CPU"""
    import spatial_model_editor_b200 as smeb
    if smeb.device_count() > 0:
        pytest.skip("a GPU is present")
    p, _, _ = load_golden("liver-simplified-scan")
    ens = EN.BatchedEnsemble(p, EN.scan_values(p.params[0], 4))
    msg = ens.errorMessage()
    assert msg.startswith("No CUDA devices were found") or \
        msg.startswith("Failed to initialize the CUDA driver") or \
        msg.startswith("Failed to query CUDA devices")
    with pytest.raises(RuntimeError):
        ens.run(0.1)
    assert ens.launchCount() == 0
    ens.close()


def test_members_are_dealt_round_robin_and_groups_cover_them():
    assert EN.members_of_rank(10, 1, 4) == [1, 5, 9]
    mine = EN.members_of_rank(64, 3, 8)
    groups = [mine[g::2] for g in range(2)]
    assert sorted(groups[0] + groups[1]) == mine and len(mine) == 8
