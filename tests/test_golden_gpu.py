"""GPU (C-ABI) vs CPU oracle on the reference's own example models
(tests/golden/*.npz): the protocol of cudapixelsim_t.cpp:542-801 — every model x
{RK101, RK212, RK323, RK435} one step and a short fixed-step run, three models
adaptive — at the fixed-step bar of 1e-12 relative."""
import numpy as np
import pytest

import spatial_model_editor_b200 as smeb
from oracle.oracle_sim import OraclePixelSim
from util import assert_parity, load_golden

pytestmark = pytest.mark.gpu

MODELS = ["single-pixel-chain", "very-simple-model", "ABtoC", "gray-scott", "gray-scott-3d",
          "brusselator-model", "liver-simplified", "single-compartment-diffusion",
          # the rest of allExampleModels (cudapixelsim_t.cpp:50-64): all 14 are here
          "single-compartment-diffusion-3d", "circadian-clock", "liver-cells",
          "very-simple-model-3d", "FitzhughNagumo3D", "SelKov3D", "CalciumWavePropagation3D"]
INTEGRATORS = [smeb.RK101, smeb.RK212, smeb.RK323, smeb.RK435]


def pair(name, integ, dt=None, rel=None):
    p, _, _ = load_golden(name, integrator=integ)
    if dt is not None:
        p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
        p.max_timestep = dt
    if rel is not None:
        p.max_err_rel = rel
    p.device = 0
    gpu = smeb.PixelSim(p)
    assert gpu.errorMessage() == "", gpu.errorMessage()
    return p, gpu, OraclePixelSim(p)


def compare(p, gpu, cpu, what, rel=1e-12):
    for ci, c in enumerate(p.compartments):
        assert_parity(gpu.getConcentrations(ci), cpu.getConcentrations(ci), rel=rel,
                      what=f"{what} conc[{c.id}]")
        assert_parity(gpu.getDcdt(ci), cpu.getDcdt(ci), rel=rel, what=f"{what} dcdt[{c.id}]")


@pytest.mark.parametrize("name", MODELS)
@pytest.mark.parametrize("integ", INTEGRATORS)
def test_models_fixed_step(name, integ):
    # a step below the model's diffusion stability bound (liver-simplified has
    # 0.3-unit voxels: 1/(2 sum D/dx^2) is well below 1e-3)
    p, gpu, cpu = pair(name, integ, dt=1e-3)
    dt = min(1e-3, 0.5 * gpu.maxStableTimestep)
    if dt < 1e-3:
        p, gpu, cpu = pair(name, integ, dt=dt)
    assert abs(gpu.maxStableTimestep - cpu.maxStableTimestep) <= 1e-15 * cpu.maxStableTimestep
    assert gpu.run(1e-7) == cpu.run(1e-7) == 1
    compare(p, gpu, cpu, f"{name} first step")
    ng, nc = gpu.run(20 * dt), cpu.run(20 * dt)
    assert ng == nc and ng >= 20
    compare(p, gpu, cpu, f"{name} run")
    assert gpu.errorMessage() == cpu.errorMessage() == ""


@pytest.mark.parametrize("name", ["very-simple-model", "ABtoC", "liver-simplified",
                                  "gray-scott-3d"])
@pytest.mark.parametrize("integ", [smeb.RK212, smeb.RK323, smeb.RK435])
def test_models_adaptive(name, integ):
    """adaptive runs with the models' own tolerances: same step counts, state
    within the solver tolerance (here far tighter because the sequences agree)"""
    p, gpu, cpu = pair(name, integ)
    t = 0.2
    ng, nc = gpu.run(t), cpu.run(t)
    assert ng == nc and gpu.discardedSteps == cpu.discardedSteps
    compare(p, gpu, cpu, f"{name} adaptive", rel=1e-9)
    for ci, c in enumerate(p.compartments):
        rng = np.random.default_rng(3)
        for _ in range(8):
            s, px = int(rng.integers(c.n_species)), int(rng.integers(c.n_voxels))
            a = gpu.getLowerOrderConcentration(ci, s, px)
            b = cpu.getLowerOrderConcentration(ci, s, px)
            assert abs(a - b) <= 1e-9 * max(abs(b), 1e-3)


@pytest.mark.parametrize("index,name", list(enumerate(MODELS[1:])))
def test_example_model_sweep_random_states(index, name):
    """cudapixelsim_t.cpp:542-575 "example-model sweep matches CPU for all example
    models": every species of every compartment restarted from U(0.05, 1.25)
    (randomizeReactiveSpecies, seed 1000 + index), RK101 with maxTimestep = dt, one
    interval of one step; state and dc/dt at 1e-12 relative (the reference: abs
    1e-13 / rel 1e-11).  dt = min(1e-3, half the stability bound) instead of the
    reference's 1e-6, so that the step visibly moves the state"""
    p, _, _ = load_golden(name, integrator=smeb.RK101)
    rng = np.random.default_rng(1000 + index)
    for c in p.compartments:
        c.conc0 = np.ascontiguousarray(rng.uniform(0.05, 1.25, size=c.conc0.shape))
    p.device = 0
    cpu = OraclePixelSim(p)
    dt = min(1e-3, 0.5 * cpu.maxStableTimestep)
    p.max_timestep = dt
    gpu, cpu = smeb.PixelSim(p), OraclePixelSim(p)
    assert gpu.errorMessage() == cpu.errorMessage() == ""
    assert gpu.run(dt) == cpu.run(dt) == 1
    compare(p, gpu, cpu, f"{name} random state")
    assert gpu.errorMessage() == cpu.errorMessage() == ""


def test_k1_chain_on_gpu():
    """simulate_t.cpp:127-147 through the CUDA path: closed-form Euler step"""
    dt = 0.134521234
    p, gpu, cpu = pair("single-pixel-chain", smeb.RK101, dt=dt)
    assert gpu.run(dt) == 1
    assert abs(gpu.getConcentrations(1)[0, 0] - 0.1 * dt) <= 1e-14 * 0.1 * dt
    assert abs(gpu.getDcdt(1)[0, 0] - 0.1) <= 1e-15


def test_adaptive_failure_on_gpu():
    """simulate_t.cpp:367-380 through the CUDA path"""
    p, _, _ = load_golden("very-simple-model")
    m = p.membranes[0]
    m.expressions = [e.replace("(0.1 * 1.0)", "(1e+40 * 1.0)") for e in m.expressions]
    p.device = 0
    gpu, cpu = smeb.PixelSim(p), OraclePixelSim(p)
    ng, nc = gpu.run(1.0), cpu.run(1.0)
    assert gpu.errorMessage() == cpu.errorMessage() != ""
    assert ng == nc
    assert gpu.errorImages().shape == (1, 100, 100)
    assert np.array_equal(gpu.errorImages(), cpu.errorImages())


def test_k7_voxel_size_invariance_on_gpu():
    """simulate_t.cpp:515-680 through the CUDA path, with the device-side
    avg / min / max that replaces calculateAvgMinMax"""
    from test_golden_cpu import k7_problems
    stats = []
    for p in k7_problems():
        p.device = 0
        sim = smeb.PixelSim(p)
        assert sim.run(0.20) == 20 and sim.errorMessage() == ""
        stats.append([sim.getAvgMinMax(ci) for ci in range(3)])
    for a, b in zip(*stats):
        assert np.all(np.abs(a - b) <= np.maximum(1e-13, 1e-8 * np.maximum(np.abs(a), np.abs(b))))
