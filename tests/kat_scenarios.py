"""More of the reference's own known-answer tests for the Pixel path, restated once
and run twice: on the CPU oracle (tests/test_kat_more_cpu.py, pins the oracle) and
through the CUDA path (tests/test_kat_more_gpu.py).

Every scenario takes ``make``: a callable ``Problem -> simulator`` with the
``BaseSim`` surface (run, getConcentrations, getDcdt, errorMessage,
getConcentrationPadding).  The expected values are the assertions of the cited
reference tests (simulate_t.cpp, pixelsim_t.cpp), not outputs of this repository.
"""
import time

import numpy as np

import spatial_model_editor_b200 as smeb
from util import load_golden

DBL_MAX = smeb.DBL_MAX


def _small_disc(integrator=None):
    p, meta, labels = load_golden("small-single-compartment-diffusion", integrator=integrator)
    p.voxel_size = tuple(meta["voxel_size"])
    return p, meta, labels


def _avg_min_max(c):
    return c.mean(axis=0), c.min(axis=0), c.max(axis=0)


def uniform_steady_state(make):
    """simulate_t.cpp:1158-1202: small-single-compartment-diffusion, maxErr =
    {DBL_MAX, 0.001}, one interval of 50: both species end up uniform,
    |min - avg| / avg and |max - avg| / avg < 1e-3"""
    p, _, _ = _small_disc()
    p.max_err_abs, p.max_err_rel = DBL_MAX, 0.001
    sim = make(p)
    assert sim.run(50.0) >= 1
    assert sim.errorMessage() == ""
    avg, lo, hi = _avg_min_max(sim.getConcentrations(0))
    assert np.all(np.abs((lo - avg) / avg) < 1e-3)
    assert np.all(np.abs((hi - avg) / avg) < 1e-3)
    # zero-flux boundaries: the averages are the initial ones
    assert np.all(np.abs(avg - p.compartments[0].conc0.mean(axis=0)) <= 1e-9 * avg)
    return sim


def zero_storage_uniform(make, tol=1e-12, t=5.0):
    """simulate_t.cpp:1204-1237: storage(slow) = 0 with D > 0 — the constraint
    0 = div(D grad c) makes "slow" spatially uniform: maxErr = {1e-12, 1e-12},
    one interval of 5, avg > 1e-12, |min - avg| / avg and |max - avg| / avg < 1e-6.
    (3.5 M steps at those tolerances: the CUDA test runs maxErr = 1e-7 over 0.5,
    7 k steps — the residual non-uniformity is 2.7 x maxErr on the CPU solver.)"""
    p, _, _ = _small_disc()
    c = p.compartments[0]
    inv = np.ones(c.n_species)
    inv[0] = 0.0
    c.inv_storage = inv
    p.max_err_abs, p.max_err_rel = tol, tol
    sim = make(p)
    assert sim.errorMessage() == ""
    assert sim.run(t) >= 1
    assert sim.errorMessage() == ""
    avg, lo, hi = _avg_min_max(sim.getConcentrations(0))
    assert avg[0] > 1e-12
    assert abs((lo[0] - avg[0]) / avg[0]) < 1e-6
    assert abs((hi[0] - avg[0]) / avg[0]) < 1e-6
    return sim


def varying_diffusion_steady_state(make):
    """simulate_t.cpp:1239-1303: D(x) = 1 + x / w for both species (sampled field):
    uniform to 1e-3 after one interval of 50 and, for the Pixel solver, the total
    amount is conserved to 1e-8 relative"""
    p, _, labels = _small_disc()
    c = p.compartments[0]
    w = labels.shape[2]
    d = 1.0 + c.voxels[:, 0].astype(np.float64) / float(w)
    c.diffusion_array = np.ascontiguousarray(np.repeat(d[:, None], c.n_species, axis=1))
    p.max_err_abs, p.max_err_rel = DBL_MAX, 0.001
    mass0 = c.conc0.sum(axis=0)
    sim = make(p)
    assert sim.run(50.0) >= 1
    assert sim.errorMessage() == ""
    out = sim.getConcentrations(0)
    avg, lo, hi = _avg_min_max(out)
    assert np.all(np.abs((lo - avg) / avg) < 1e-3)
    assert np.all(np.abs((hi - avg) / avg) < 1e-3)
    assert np.all(np.abs((out.sum(axis=0) - mass0) / mass0) < 1e-8)
    return sim


def storage_scales_dcdt(make):
    """pixelsim_t.cpp:131-159: ABtoC, RK101, maxTimestep 1e-3, run(1e-3): with
    storage(C) = 2 the dc/dt of A and B is unchanged and that of C is halved"""
    sims = []
    for storage_c in (1.0, 2.0):
        p, _, _ = load_golden("ABtoC", integrator=smeb.RK101)
        p.max_timestep = 1e-3
        if storage_c != 1.0:
            p.compartments[0].inv_storage = np.array([1.0, 1.0, 1.0 / storage_c])
        sim = make(p)
        assert sim.errorMessage() == ""
        sim.run(1e-3)
        sims.append(sim)
    d0, d1 = sims[0].getDcdt(0), sims[1].getDcdt(0)
    assert d0.shape == d1.shape and d0.shape[1] == 3
    tol = 1e-12  # dbl_approx (catch_wrapper.hpp)
    assert np.all(np.abs(d1[:, :2] - d0[:, :2]) <= tol * np.abs(d0[:, :2]))
    assert np.all(np.abs(d1[:, 2] - 0.5 * d0[:, 2]) <= tol * np.abs(d0[:, 2]))
    assert np.any(d0 != 0.0)
    return sims


def _disc(integrator):
    p, meta, _ = load_golden("single-compartment-diffusion", integrator=integrator)
    p.voxel_size = tuple(meta["voxel_size"])
    raw = np.array(meta["compartments"][0]["diffusion_constants"])
    return p, meta, raw


def cross_diffusion_vanishes_for_uniform_source(make):
    """pixelsim_t.cpp:160-187: "slow" uniform at 1 -> grad(slow) = 0, so the
    cross-diffusion term fast <- slow with constant 2 changes nothing: after RK101
    run(1e-3) both models hold the same concentrations"""
    out = []
    for cross in (False, True):
        p, meta, raw = _disc(smeb.RK101)
        c = p.compartments[0]
        c.conc0 = np.ascontiguousarray(c.conc0.copy())
        c.conc0[:, 0] = 1.0
        if cross:
            c.diffusion_raw = raw
            c.cross_diffusion = [(1, 0, "2.0")]
        p.max_timestep = 1e-3
        sim = make(p)
        assert sim.errorMessage() == ""
        sim.run(1e-3)
        out.append(np.array(sim.getConcentrations(0)))  # the view dies with sim
    assert np.all(np.abs(out[0] - out[1]) <= 1e-12 * np.abs(out[0]))
    return out


def pad_with_coordinates(p, meta, ny):
    """what the reference does when an expression depends on x, y or z
    (pixelsim.cpp:282-322, pixelsim_impl.cpp:405-423): the state of every voxel is
    followed by its centre coordinates, which never change"""
    c = p.compartments[0]
    dx, dy, dz = meta["voxel_size"]
    ox, oy, oz = meta["origin"]
    v = c.voxels.astype(np.float64)
    x = ox + dx * (v[:, 0] + 0.5)
    y = oy + dy * ((ny - 1 - v[:, 1]) + 0.5)
    z = oz + dz * (v[:, 2] + 0.5)
    n = c.n_species
    c.n_primary = n
    c.species = list(c.species) + ["x", "y", "sme_coord_z"]
    c.expressions = list(c.expressions) + ["0.0", "0.0", "0.0"]
    c.diffusion = np.vstack([c.diffusion, np.zeros((3, 3))])
    c.conc0 = np.ascontiguousarray(np.column_stack([c.conc0, x, y, z]))
    return n


def voxel_centre_coordinates(make):
    """pixelsim_t.cpp:217-258: cross-diffusion constant "1 + x" pads the state with
    x, y, z (padding 3); each equals origin + (pixel + 0.5) * voxel size for an
    integer pixel inside the image — also after a step"""
    p, meta, raw = _disc(smeb.RK101)
    labels_shape = (1, 100, 100)
    n = pad_with_coordinates(p, meta, labels_shape[1])
    c = p.compartments[0]
    c.diffusion_raw = np.concatenate([raw, np.zeros(3)])
    c.cross_diffusion = [(0, 1, f"(1.0 + sme_var_{n})")]
    p.max_timestep = 1e-3
    sim = make(p)
    assert sim.errorMessage() == ""
    assert sim.getConcentrationPadding() == 3
    dx, dy, dz = meta["voxel_size"]
    ox, oy, oz = meta["origin"]
    for _ in range(2):
        conc = sim.getConcentrations(0)
        assert conc.shape[1] == 2 + 3
        for k, (o, d, dim) in enumerate(((ox, dx, labels_shape[2]), (oy, dy, labels_shape[1]),
                                         (oz, dz, labels_shape[0]))):
            i = (conc[:, 2 + k] - o) / d - 0.5
            assert np.all(np.abs(i - np.round(i)) <= 1e-9 * np.maximum(np.abs(i), 1.0))
            assert np.all(i >= 0.0) and np.all(i < dim)
        sim.run(1e-3)
        assert sim.errorMessage() == ""
    return sim


def cross_diffusion_z_only_3d(make):
    """pixelsim_t.cpp:259-272: cross-diffusion constant "1 + z" on the 3-d example:
    padding 3, one RK101 run of 1e-3 without error"""
    p, meta, _ = load_golden("single-compartment-diffusion-3d", integrator=smeb.RK101)
    p.voxel_size = tuple(meta["voxel_size"])
    raw = np.array(meta["compartments"][0]["diffusion_constants"])
    n = pad_with_coordinates(p, meta, 50)
    c = p.compartments[0]
    c.diffusion_raw = np.concatenate([raw, np.zeros(3)])
    c.cross_diffusion = [(0, 1, f"(1.0 + sme_var_{n + 2})")]
    p.max_timestep = 1e-3
    sim = make(p)
    assert sim.errorMessage() == ""
    assert sim.getConcentrationPadding() == 3
    sim.run(1e-3)
    assert sim.errorMessage() == ""
    return sim


def rk101_capped_by_cross_diffusion(make):
    """pixelsim_t.cpp:273-299: D = 0 for both species, cross-diffusion 80 each way,
    RK101 with maxTimestep 1e-2: run(1e-2) takes more than one step (the
    forward-Euler bound of the cross terms caps it) and nothing goes negative"""
    p, meta, _ = _disc(smeb.RK101)
    c = p.compartments[0]
    c.diffusion = np.zeros_like(c.diffusion)
    c.diffusion_raw = np.zeros(2)
    c.cross_diffusion = [(0, 1, "80.0"), (1, 0, "80.0")]
    p.max_timestep = 1e-2
    sim = make(p)
    assert sim.errorMessage() == ""
    n = sim.run(1e-2)
    assert sim.errorMessage() == ""
    assert n > 1
    assert sim.getConcentrations(0).min() >= 0.0
    return sim


def timeout(make):
    """simulate_t.cpp:1576-1594: brusselator with a 200 ms timeout — one very long
    interval does not complete ("Simulation stopped early"), many tiny ones make
    progress before the deadline"""
    p, _, _ = load_golden("brusselator-model")
    sim = make(p)
    t0 = time.perf_counter()
    sim.run(1e9, 200.0)
    assert time.perf_counter() - t0 < 5.0
    assert sim.errorMessage() == "Simulation stopped early"
    p2, _, _ = load_golden("brusselator-model")
    sim2 = make(p2)
    assert sim2.run(1e-12, 200.0) >= 1
    assert sim2.errorMessage() == ""
    return sim2


def stop_then_continue(make):
    """simulate_t.cpp:2746-2768 (issue 544): a stop request ends a huge interval
    early (at least one step is taken first, pixelsim.cpp:437-463); after the caller
    clears the flag (Simulation::doMultipleTimesteps, simulate.cpp:448-453; run()
    itself clears the message, pixelsim.cpp:428) three more intervals of 0.01 run
    without error"""
    p, _, _ = load_golden("very-simple-model")
    sim = make(p)
    sim.setStopRequested(True)
    assert sim.run(5000.0) >= 1
    assert sim.errorMessage() == "Simulation stopped early"
    assert sim.getStopRequested()
    sim.setStopRequested(False)
    steps = [sim.run(0.01) for _ in range(3)]
    assert sim.errorMessage() == ""
    assert all(s >= 1 for s in steps)
    return sim


def example_3d_models_tiny_step(make, names=("gray-scott-3d", "single-compartment-diffusion-3d",
                                             "very-simple-model-3d", "FitzhughNagumo3D",
                                             "SelKov3D")):
    """simulate_t.cpp:2954-2976 "Simulate example 3d models: do a tiny step, don't
    crash": one interval of 0.01 with the model's own options completes"""
    sims = []
    for name in names:
        p, _, _ = load_golden(name)
        sim = make(p)
        assert sim.errorMessage() == "", name
        assert sim.run(0.01) >= 1, name
        assert sim.errorMessage() == "", name
        for ci in range(len(p.compartments)):
            assert np.all(np.isfinite(sim.getConcentrations(ci))), name
        sims.append(sim)
    return sims


def do_timesteps(make, name, time, integrator=None, tolerances=None):
    """Simulation::doTimesteps(time) for one interval on a fixture with SBML events
    (simulate.cpp:397-445 and 227-270): an event inside the interval ends a sub-step
    at its time, the state is read back, the simulator is destroyed and re-created
    with the event's value substituted into the reaction expressions — so its
    controller starts again from nextTimestep = 1e-7 (pixelsim_base.hpp:45) — and
    restarted from that state (pixelsim.cpp:399-406); then the rest of the interval.
    Returns (final state per compartment, list of simulators used)."""
    def problem(stage):
        p, meta, _ = load_golden(name, integrator=integrator)
        if tolerances is not None:
            p.max_err_abs, p.max_err_rel = tolerances
        if stage is not None:
            for c, ex in zip(p.compartments, stage["expressions"]):
                c.expressions = list(ex)
            for m, ex in zip(p.membranes, stage["membrane_expressions"]):
                m.expressions = list(ex)
        return p, meta

    p, meta = problem(None)
    sims = [make(p)]
    t = 0.0
    for stage in meta["events"]:
        # events at (to 1e-12 of the interval) the current time are applied before
        # stepping; events after the end of the interval stay queued
        if (t + (time - t) - stage["time"]) / time <= 0.1e-12:
            break
        if stage["time"] > t:
            assert sims[-1].run(stage["time"] - t) >= 1
            assert sims[-1].errorMessage() == ""
            t = stage["time"]
        state = [np.array(sims[-1].getConcentrations(i))
                 for i in range(len(p.compartments))]
        p, _ = problem(stage)
        for c, s in zip(p.compartments, state):
            c.conc0 = s
        sims.append(make(p))
    if time > t:
        assert sims[-1].run(time - t) >= 1
        assert sims[-1].errorMessage() == ""
    return [np.array(sims[-1].getConcentrations(i)) for i in range(len(p.compartments))], sims


def brusselator_rk_orders_agree(make, make_accurate=None):
    """K5, simulate_t.cpp:1596-1647: on the brusselator example RK212, RK323 and RK435
    at maxErr = {DBL_MAX, 0.01} end within 1 % (signed, like the reference's
    assertion) of RK435 at 1e-6 after Simulation::doTimesteps(30).  The model file
    carries the event "time >= 25: k2 = 4" (brusselator-model.xml, listOfEvents), so
    the interval is a sub-step of 25, a re-created simulator with k2 = 4, and 5 more."""
    tol = (DBL_MAX, 1e-6)
    (c4,), _ = do_timesteps(make_accurate or make, "brusselator-model", 30.0,
                            integrator=smeb.RK435, tolerances=tol)
    worst = {}
    for integ in (smeb.RK212, smeb.RK323, smeb.RK435):
        (c,), sims = do_timesteps(make, "brusselator-model", 30.0, integrator=integ,
                                  tolerances=(DBL_MAX, 0.01))
        assert len(sims) == 2  # the event at t = 25 re-created the simulator once
        worst[integ] = float(np.max((c[:, 0] - c4[:, 0]) / (c4[:, 0] + 1e-20)))
        assert worst[integ] < 0.01, (integ, worst)
    return worst


def membrane_parallel_matches_serial(make_serial, make_parallel):
    """K6, simulate_t.cpp:382-439: very-simple-model, RK101, maxErr = DBL_MAX,
    maxTimestep 0.01, one interval of 1: the parallel execution takes the same
    steps as the serial one and agrees in every concentration and dc/dt to 1e-13
    (Catch::Approx margin 1e-13, epsilon 1e-13)"""
    sims = []
    for make in (make_serial, make_parallel):
        p, _, _ = load_golden("very-simple-model", integrator=smeb.RK101)
        p.max_err_abs = p.max_err_rel = DBL_MAX
        p.max_timestep = 0.01
        sim = make(p)
        assert sim.errorMessage() == ""
        sims.append((p, sim, sim.run(1.0)))
    (p, serial, n_serial), (_, parallel, n_parallel) = sims
    assert n_serial == n_parallel >= 100
    tol = 1e-13
    for ci in range(len(p.compartments)):
        for get in ("getConcentrations", "getDcdt"):
            a, b = getattr(serial, get)(ci), getattr(parallel, get)(ci)
            assert a.shape == b.shape
            assert np.all(np.abs(a - b) <= tol + tol * np.abs(a)), (ci, get)
    return sims
