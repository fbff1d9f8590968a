"""GPU-vs-oracle parity through the C-ABI (the parity tests proper).

Protocol of the reference's own GPU-vs-CPU suite (cudapixelsim_t.cpp:416-801):
identical random initial state U(0.05, 1.25), compare dc/dt, concentrations and
the lower-order state after one step and after short runs — with the fixed-step
tolerance of BASELINE.json (1e-12 relative) instead of the reference's 1e-11.
"""
import numpy as np
import pytest

import spatial_model_editor_b200 as smeb
from spatial_model_editor_b200 import models as M
from spatial_model_editor_b200.pde import ReactionDef
from oracle.oracle_sim import OraclePixelSim
from util import assert_parity

pytestmark = pytest.mark.gpu

INTEGRATORS = [smeb.RK101, smeb.RK212, smeb.RK323, smeb.RK435]


def make_pair(problem):
    problem.device = 0
    gpu = smeb.PixelSim(problem)
    assert gpu.errorMessage() == "", gpu.errorMessage()
    cpu = OraclePixelSim(problem)
    assert cpu.errorMessage() == ""
    return gpu, cpu


def compare_state(gpu, cpu, problem, what, lower=True):
    for ci, comp in enumerate(problem.compartments):
        assert_parity(gpu.getConcentrations(ci), cpu.getConcentrations(ci),
                      what=f"{what} conc[{comp.id}]")
        assert_parity(gpu.getDcdt(ci), cpu.getDcdt(ci), what=f"{what} dcdt[{comp.id}]")
        if lower and comp.n_voxels > 0:
            rng = np.random.default_rng(7)
            for _ in range(16):
                s = int(rng.integers(comp.n_species))
                px = int(rng.integers(comp.n_voxels))
                a = gpu.getLowerOrderConcentration(ci, s, px)
                b = cpu.getLowerOrderConcentration(ci, s, px)
                assert abs(a - b) <= 1e-12 * max(abs(b), 1e-2), (what, a, b)


def fixed_step(problem, dt):
    """the reference benchmark's fixed-step configuration
    (benchmark/benchmark.cpp:214-216)"""
    problem.max_err_abs = smeb.DBL_MAX
    problem.max_err_rel = smeb.DBL_MAX
    problem.max_timestep = dt
    return problem


def _variant(problem, v):
    problem.kernel_variant = v
    return problem


GEOMETRIES = {
    "box3d": lambda integ: M.box_problem(M.GRAY_SCOTT, 19, 23, 7, integrator=integ,
                                         voxel_size=(2.0, 2.0, 2.0)),
    "box2d": lambda integ: M.box_problem(M.GRAY_SCOTT, 64, 48, 1, integrator=integ),
    "box3d_direct": lambda integ: _variant(M.box_problem(
        M.GRAY_SCOTT, 19, 23, 7, integrator=integ, voxel_size=(2.0, 2.0, 2.0)), 1),
    "box2d_direct": lambda integ: _variant(M.box_problem(
        M.GRAY_SCOTT, 64, 48, 1, integrator=integ), 1),
    # several tiles in x and y, several z chunks, ragged tile edges
    "box3d_tiles": lambda integ: _variant(M.box_problem(
        M.GRAY_SCOTT, 44, 150, 70, integrator=integ, voxel_size=(1.0, 2.0, 1.5)), 2),
    # 2-D: several y strips, several x chunks of bricks, ragged edges
    "box2d_tiles": lambda integ: _variant(M.box_problem(
        M.BRUSSELATOR, 300, 200, 1, integrator=integ), 2),
    "box_as_list": lambda integ: M.box_problem(M.BRUSSELATOR, 17, 9, 4, integrator=integ,
                                               as_voxel_list=True),
    "abtoc_disc": lambda integ: M.labelled_problem(
        M.disc_labels(40, 40), [dict(label=1, id="comp", kin=M.ABTOC)], integrator=integ),
    "membranes2d": lambda integ: M.two_compartment_brusselator(33, 29, 1, integrator=integ),
    "membranes3d": lambda integ: M.two_compartment_brusselator(12, 13, 6, integrator=integ,
                                                               voxel_size=(1.0, 0.5, 2.0)),
}


@pytest.mark.parametrize("geom", list(GEOMETRIES))
@pytest.mark.parametrize("integ", INTEGRATORS)
def test_single_step(geom, integ):
    """one step of size nextTimestep = 1e-7 (pixelsim_base.hpp:45) then a larger one"""
    p = fixed_step(GEOMETRIES[geom](integ), 1e-3)
    gpu, cpu = make_pair(p)
    assert gpu.run(1e-7) == cpu.run(1e-7) == 1
    compare_state(gpu, cpu, p, f"{geom} first step")
    assert gpu.run(1e-3) == cpu.run(1e-3)
    compare_state(gpu, cpu, p, f"{geom} second run")


@pytest.mark.parametrize("geom", list(GEOMETRIES))
@pytest.mark.parametrize("integ", INTEGRATORS)
def test_fixed_step_run(geom, integ):
    p = fixed_step(GEOMETRIES[geom](integ), 0.01)
    gpu, cpu = make_pair(p)
    ng, nc = gpu.run(0.5), cpu.run(0.5)
    assert ng == nc and ng >= 50
    compare_state(gpu, cpu, p, f"{geom} fixed-step run")
    assert gpu.discardedSteps == cpu.discardedSteps == 0


@pytest.mark.parametrize("geom", ["box3d", "box3d_direct", "box2d_tiles", "abtoc_disc", "membranes2d"])
@pytest.mark.parametrize("integ", [smeb.RK212, smeb.RK323, smeb.RK435])
def test_adaptive_run(geom, integ):
    """adaptive runs: same accepted/discarded step counts and a final state well
    within the solver's relative tolerance (here the step sequences coincide, so
    the states agree far tighter)"""
    p = GEOMETRIES[geom](integ)
    p.max_err_rel = 0.005
    gpu, cpu = make_pair(p)
    ng, nc = gpu.run(1.0), cpu.run(1.0)
    assert ng == nc
    assert gpu.discardedSteps == cpu.discardedSteps
    assert abs(gpu.nextTimestep - cpu.nextTimestep) <= 1e-9 * cpu.nextTimestep
    for ci in range(len(p.compartments)):
        assert_parity(gpu.getConcentrations(ci), cpu.getConcentrations(ci), rel=1e-9,
                      what=f"{geom} adaptive")
    # continue (second image interval), nextTimestep carried over
    assert gpu.run(0.5) == cpu.run(0.5)
    for ci in range(len(p.compartments)):
        assert_parity(gpu.getConcentrations(ci), cpu.getConcentrations(ci), rel=1e-9,
                      what=f"{geom} adaptive 2")


def test_negative_clamp():
    """a decay reaction overshooting below zero must be clamped after accepted
    steps exactly like pixelsim_impl.cpp:643-665"""
    kin = dict(species=["A", "B"], D=[1.0, 0.3], constants={},
               reactions=[ReactionDef("d", "40 * (A + 0.2)", {"A": -1.0, "B": 0.5})])
    for integ in INTEGRATORS:
        p = fixed_step(M.box_problem(kin, 9, 8, 3, integrator=integ), 0.02)
        gpu, cpu = make_pair(p)
        assert gpu.run(0.3) == cpu.run(0.3)
        b = cpu.getConcentrations(0)
        assert (b[:, 0] == 0.0).any()
        compare_state(gpu, cpu, p, f"clamp integ {integ}")
        assert (gpu.getConcentrations(0) >= 0.0).all()


def test_restart_concentrations():
    """SimulationData restart path: setConcentrations then run"""
    p = fixed_step(M.box_problem(M.GRAY_SCOTT, 10, 11, 3), 0.01)
    gpu, cpu = make_pair(p)
    gpu.run(0.1)
    cpu.run(0.1)
    new = M.random_state(p.compartments[0].n_voxels, 2, 4242)
    gpu.setConcentrations(0, new)
    cpu.setConcentrations(0, new)
    assert_parity(gpu.getConcentrations(0), new)
    assert gpu.run(0.1) == cpu.run(0.1)
    compare_state(gpu, cpu, p, "restart")


def test_species_counts():
    """odd and larger species counts (scalar vs vector load paths; liver has 8)"""
    for S in (1, 3, 5, 8):
        names = [f"s{i}" for i in range(S)]
        reacs = [ReactionDef(f"r{i}", f"0.3 * s{i} * s{(i + 1) % S} / (0.5 + s{i})",
                             {names[i]: -1.0, names[(i + 1) % S]: 1.0}) for i in range(S)]
        kin = dict(species=names, D=[0.1 * (i + 1) for i in range(S)], constants={},
                   reactions=reacs)
        p = fixed_step(M.box_problem(kin, 7, 6, 5, integrator=smeb.RK323), 0.02)
        gpu, cpu = make_pair(p)
        assert gpu.run(0.2) == cpu.run(0.2)
        compare_state(gpu, cpu, p, f"S={S}")


def test_stop_callback_and_timeout():
    """K9: stop callback -> 'Simulation stopped early' (pixelsim_t.cpp:12-20)"""
    p = M.box_problem(M.GRAY_SCOTT, 16, 16, 1)
    p.device = 0
    gpu = smeb.PixelSim(p)
    n = gpu.run(100.0, -1.0, lambda: True)
    assert n == 1
    assert gpu.errorMessage() == "Simulation stopped early"
    assert gpu.getStopRequested()
    gpu2 = smeb.PixelSim(p)
    n = gpu2.run(1e9, 50.0, None)
    assert gpu2.errorMessage() == "Simulation stopped early"
    assert n >= 1


def test_adaptive_failure_error_image():
    """K8: k1 = 1e40 makes the step size collapse -> message + error image of
    the geometry size (simulate_t.cpp:367-380)"""
    kin = dict(M.ABTOC)
    kin["reactions"] = [ReactionDef("r1", "A * B * k1", {"A": -1.0, "B": -1.0, "C": 1.0},
                                    {"k1": 1e40})]
    lab = M.disc_labels(30, 30)
    p = M.labelled_problem(lab, [dict(label=1, id="comp", kin=kin)])
    gpu, cpu = make_pair(p)
    ng, nc = gpu.run(1.0), cpu.run(1.0)
    assert ng == nc
    assert cpu.errorMessage().startswith("Failed to solve model to required accuracy")
    assert gpu.errorMessage() == cpu.errorMessage()
    gi, ci = gpu.errorImages(), cpu.errorImages()
    assert gi.shape == ci.shape == (1, 30, 30)
    assert (gi == ci).all() and (gi != 0).any()


def test_ctor_and_run_on_different_threads():
    """cudapixelsim_t.cpp:515-540"""
    import threading
    p = fixed_step(M.box_problem(M.GRAY_SCOTT, 12, 12, 2), 0.01)
    gpu, cpu = make_pair(p)
    out = {}
    t = threading.Thread(target=lambda: out.setdefault("n", gpu.run(0.1)))
    t.start()
    t.join()
    assert out["n"] == cpu.run(0.1)
    compare_state(gpu, cpu, p, "threads")


# ---- CUDA-graph replay of a step (small models) vs plain launches -----------------
@pytest.mark.parametrize("name", ["liver-simplified", "very-simple-model", "ABtoC"])
@pytest.mark.parametrize("integ", [smeb.RK101, smeb.RK212, smeb.RK435])
def test_graph_replay_is_bit_identical_to_plain_launches(name, integ, monkeypatch):
    """one cudaGraphLaunch per step (step size read from device memory) gives the
    bits of the kernel-by-kernel path, same step sequence, same on-demand outputs"""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from util import load_golden
    out = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("SMEB200_GRAPHS", mode)
        p, _, _ = load_golden(name, integrator=integ)
        p.device = 0
        sim = smeb.PixelSim(p)
        assert sim.errorMessage() == ""
        n = sim.run(0.05) + sim.run(0.05)
        out[mode] = (n, sim.discardedSteps, sim.launchCount(),
                     [sim.getConcentrations(i).copy() for i in range(len(p.compartments))],
                     [sim.getDcdt(i).copy() for i in range(len(p.compartments))],
                     sim.getLowerOrderConcentration(0, 0, 3))
    g, pl = out["1"], out["0"]
    assert g[0] == pl[0] and g[1] == pl[1] and g[2] == pl[2]
    for a, b in zip(g[3] + g[4], pl[3] + pl[4]):
        assert np.array_equal(a, b)
    assert g[5] == pl[5]


# ---- unbounded tolerances: error norms checked two steps late ----------------------
@pytest.mark.parametrize("geom", ["box3d_tiles", "box2d_tiles", "abtoc_disc", "membranes3d"])
@pytest.mark.parametrize("integ", [smeb.RK212, smeb.RK323, smeb.RK435])
def test_lagged_error_check_is_bit_identical(geom, integ, monkeypatch):
    """maxErr = {DBL_MAX, DBL_MAX} (benchmark.cpp:214-216): the controller accepts
    every finite error norm and sets the next step to dtMax (pixelsim.cpp:231-233),
    so the stream is not stalled after each step and the norms are checked two
    steps late.  Same steps, same bits as the step-by-step controller, same state
    as the CPU solver; getters, restarts and short intervals in between"""
    monkeypatch.setenv("SMEB200_GRAPHS", "0")  # small models would replay graphs
    out = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("SMEB200_LAGGED_ERROR", mode)
        p = fixed_step(GEOMETRIES[geom](integ), 0.01)
        gpu, cpu = make_pair(p)
        n = [gpu.run(1e-7), gpu.run(0.3), gpu.run(0.005), gpu.run(0.2)]
        assert n == [cpu.run(1e-7), cpu.run(0.3), cpu.run(0.005), cpu.run(0.2)]
        assert gpu.errorMessage() == "" and gpu.discardedSteps == 0
        assert gpu.nextTimestep == cpu.nextTimestep
        compare_state(gpu, cpu, p, f"{geom} lagged={mode}")
        out[mode] = [gpu.getConcentrations(i).copy() for i in range(len(p.compartments))]
    for a, b in zip(out["1"], out["0"]):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("lagged", ["1", "0"])
def test_unbounded_tolerances_blow_up_is_reported(lagged, monkeypatch):
    """a state that overflows makes the error norm infinite: errFactor = 0, the next
    step collapses and the run ends with the reference's message
    (pixelsim.cpp:231-251) — also when the norm is looked at two steps late"""
    monkeypatch.setenv("SMEB200_GRAPHS", "0")
    monkeypatch.setenv("SMEB200_LAGGED_ERROR", lagged)
    grow = dict(species=["A", "B"], D=[0.1, 0.1], constants={},
                reactions=[ReactionDef("r", "1e30 * A * A", {"A": 1.0})])
    p = fixed_step(M.box_problem(grow, 32, 16, 1, integrator=smeb.RK212), 0.05)
    gpu, cpu = make_pair(p)
    cpu.run(5.0)
    assert cpu.errorMessage().startswith("Failed to solve model to required accuracy")
    gpu.run(5.0)
    assert gpu.errorMessage().startswith("Failed to solve model to required accuracy")
    assert gpu.errorImages().size == cpu.errorImages().size > 0


def test_lagged_error_check_with_stop_and_timeout(monkeypatch):
    """stop callback and deadline still end the run after whole steps"""
    monkeypatch.setenv("SMEB200_GRAPHS", "0")
    p = fixed_step(M.box_problem(M.GRAY_SCOTT, 64, 48, 1, integrator=smeb.RK212), 0.01)
    gpu, cpu = make_pair(p)
    def stop_after(n):
        calls = {"n": 0}

        def cb():
            calls["n"] += 1
            return calls["n"] >= n
        return cb

    assert gpu.run(10.0, -1.0, stop_after(5)) == 5
    assert gpu.errorMessage() == "Simulation stopped early"
    gpu.setStopRequested(False)
    assert cpu.run(10.0, -1.0, stop_after(5)) == 5
    cpu.setStopRequested(False)
    compare_state(gpu, cpu, p, "stopped after 5 steps")
    assert gpu.run(1e9, 50.0) >= 1
    assert gpu.errorMessage() == "Simulation stopped early"


def test_symengine_pow_text_equals_products():
    """pow(x, 2) as SymEngine prints it vs x * x as the Python printer writes it: the
    same bits on the device, and the oracle's (product form) within 1e-12"""
    import copy
    p = M.box_problem(M.GRAY_SCOTT, 40, 36, 10, integrator=smeb.RK212,
                      max_err_rel=smeb.DBL_MAX, max_err_abs=smeb.DBL_MAX, max_timestep=0.05)
    p.device = 0
    q = copy.deepcopy(p)
    c = q.compartments[0]
    c.expressions = [e.replace("(sme_var_1 * sme_var_1)", "pow(sme_var_1, 2)")
                     for e in c.expressions]
    assert all("pow(" in e for e in c.expressions)
    a, b = smeb.PixelSim(p), smeb.PixelSim(q)
    assert a.errorMessage() == b.errorMessage() == ""
    assert a.run(1.0) == b.run(1.0)
    assert np.array_equal(a.getConcentrations(0), b.getConcentrations(0))
    cpu = OraclePixelSim(p)
    cpu.run(1.0)
    assert_parity(b.getConcentrations(0), cpu.getConcentrations(0), what="pow text")
