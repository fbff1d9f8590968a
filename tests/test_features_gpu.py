"""GPU (C-ABI) vs CPU oracle for the CPU-solver semantics the reference GPU
backend rejects (cudapixelsim.cpp:789-809) and this library computes: reactions
depending on time / space (padded state), non-unit storage, non-spatial species —
on voxel-list fixtures and on box compartments (TMA tile kernel)."""
import numpy as np
import pytest

import spatial_model_editor_b200 as smeb
from spatial_model_editor_b200 import models as M
from spatial_model_editor_b200.pde import ReactionDef, build_rhs
from oracle.oracle_sim import OraclePixelSim
from util import assert_parity, load_golden

pytestmark = pytest.mark.gpu

INTEGRATORS = [smeb.RK101, smeb.RK212, smeb.RK323, smeb.RK435]
# non-spatial species: the spatial mean is a serial sum on the CPU and a tree on
# the GPU, so the averaged dc/dt agrees to rounding of an N-term sum
REL = {"txy": 1e-12, "txy-storage": 1e-12, "single-compartment-diffusion-3d-storage": 1e-12,
       "very-simple-model-non-spatial-species": 1e-11}


def both(p):
    p.device = 0
    gpu = smeb.PixelSim(p)
    assert gpu.errorMessage() == "", gpu.errorMessage()
    return gpu, OraclePixelSim(p)


def compare(p, gpu, cpu, what, rel=1e-12):
    for ci, c in enumerate(p.compartments):
        assert_parity(gpu.getConcentrations(ci), cpu.getConcentrations(ci), rel=rel,
                      what=f"{what} conc[{c.id}]")
        assert_parity(gpu.getDcdt(ci), cpu.getDcdt(ci), rel=rel, what=f"{what} dcdt[{c.id}]")


@pytest.mark.parametrize("name", sorted(REL))
@pytest.mark.parametrize("integ", INTEGRATORS)
def test_feature_fixtures_fixed_step(name, integ):
    p, _, _ = load_golden(name, integrator=integ)
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    p.max_timestep = 1e-3
    gpu, cpu = both(p)
    assert gpu.getConcentrationPadding() == p.padding
    assert abs(gpu.maxStableTimestep - cpu.maxStableTimestep) <= 1e-15 * cpu.maxStableTimestep
    assert gpu.run(1e-7) == cpu.run(1e-7) == 1
    compare(p, gpu, cpu, f"{name} first step", REL[name])
    ng, nc = gpu.run(20e-3), cpu.run(20e-3)
    assert ng == nc and ng >= 20
    compare(p, gpu, cpu, f"{name} run", REL[name])


@pytest.mark.parametrize("name", ["txy-storage", "very-simple-model-non-spatial-species",
                                  "single-compartment-diffusion-3d-storage"])
@pytest.mark.parametrize("integ", [smeb.RK212, smeb.RK323, smeb.RK435])
def test_feature_fixtures_adaptive(name, integ):
    p, _, _ = load_golden(name, integrator=integ)
    gpu, cpu = both(p)
    ng, nc = gpu.run(0.2), cpu.run(0.2)
    if "non-spatial" in name:
        # The first steps (dt = 1e-7 ...) have truncation errors below fp64
        # rounding, so the controller's step sequence is driven by the last bits
        # of the error norm; the spatial mean (serial sum on the CPU, tree on the
        # GPU) differs in exactly those bits, the sequences separate, and the bar
        # is the solver's own tolerance (BASELINE.json: "within the solver's
        # stated error tolerance for adaptive runs")
        assert abs(ng - nc) <= 3
        for ci in range(len(p.compartments)):
            a, b = gpu.getConcentrations(ci), cpu.getConcentrations(ci)
            assert np.max(np.abs(a - b) / (0.5 * (np.abs(a) + np.abs(b)) + 1e-14)) \
                <= p.max_err_rel
        return
    assert ng == nc and gpu.discardedSteps == cpu.discardedSteps
    compare(p, gpu, cpu, f"{name} adaptive", rel=1e-9)


def test_txy_closed_form_on_gpu():
    """simulate_t.cpp:2009-2207 restated: dA/dt = x, dB/dt = y, dC/dt = t, forward
    Euler from zero through the CUDA path"""
    dt = 1e-3
    p, _, _ = load_golden("txy", integrator=smeb.RK101)
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    p.max_timestep = dt
    p.device = 0
    gpu = smeb.PixelSim(p)
    v = p.compartments[0].voxels.astype(np.float64)
    x, y = v[:, 0] + 0.5, (99 - v[:, 1]) + 0.5
    assert gpu.run(dt) == 1
    c = gpu.getConcentrations(0)
    assert c.shape == (3149, 7) and gpu.getConcentrationPadding() == 4
    assert np.max(np.abs(c[:, 0] - dt * x)) < 1e-15
    assert np.max(np.abs(c[:, 1] - dt * y)) < 1e-15
    n = 1 + sum(gpu.run(dt) for _ in range(9))
    c = gpu.getConcentrations(0)
    assert np.allclose(c[:, 2], dt * dt * n * (n - 1) / 2, rtol=1e-13, atol=0)
    assert np.allclose(c[:, 3], n * dt, rtol=1e-14, atol=0)
    assert np.array_equal(c[:, 4], x) and np.array_equal(c[:, 5], y)


def box_with_features(nx, ny, nz, integ, time_space, storage, non_spatial, seed=11):
    """gray-scott-like kinetics on a full box (tile kernel), optionally with a
    source term depending on t, x, y, z (padded state), storage and a non-spatial
    species"""
    kin = M.GRAY_SCOTT
    ids = list(kin["species"])
    extra = ["time", "x", "y", "z"] if time_space else []
    reactions = list(kin["reactions"])
    if time_space:
        reactions.append(ReactionDef("src", "0.01 * (1 + 0.5 * time) * (x + 2 * y + 3 * z) / 64",
                                     {"V": 1.0}))
    exprs = build_rhs(ids, reactions, kin["constants"], extra_vars=extra) + (
        ["1.0", "0.0", "0.0", "0.0"] if time_space else [])
    p = M.box_problem(kin, nx, ny, nz, integrator=integ, seed=seed)
    c = p.compartments[0]
    n = c.n_voxels
    if time_space:
        z, x, y = np.meshgrid(np.arange(nz), np.arange(nx), np.arange(ny), indexing="ij")
        pad = np.stack([np.zeros(n), x.reshape(-1) + 0.5, (ny - 1 - y.reshape(-1)) + 0.5,
                        z.reshape(-1) + 0.5], axis=1)
        c.conc0 = np.ascontiguousarray(np.concatenate([c.conc0, pad], axis=1))
        c.species = ids + extra
        c.diffusion = np.concatenate([c.diffusion, np.zeros((4, 3))])
        c.n_primary = len(ids)
    c.expressions = exprs
    if storage:
        c.inv_storage = np.array([1.0 / 1.5, 1.0 / 0.75] + [1.0] * len(extra))
    if non_spatial:
        c.non_spatial = [1]
    return p


@pytest.mark.parametrize("shape", [(32, 64, 8), (24, 40, 1), (16, 30, 5)])
@pytest.mark.parametrize("integ", INTEGRATORS)
@pytest.mark.parametrize("features", [(True, False, False), (False, True, False),
                                      (False, False, True), (True, True, True)])
def test_box_features_fixed_step(shape, integ, features):
    """the same semantics through the box kernels (TMA tile kernel where the row
    pitch allows it)"""
    p = box_with_features(*shape, integ, *features)
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    p.max_timestep = 0.05
    gpu, cpu = both(p)
    rel = 1e-11 if features[2] else 1e-12
    assert gpu.run(1e-7) == cpu.run(1e-7) == 1
    compare(p, gpu, cpu, f"box {shape} {features} first step", rel)
    ng, nc = gpu.run(0.5), cpu.run(0.5)
    assert ng == nc and ng >= 10
    compare(p, gpu, cpu, f"box {shape} {features} run", rel)


@pytest.mark.parametrize("integ", [smeb.RK212, smeb.RK323, smeb.RK435])
def test_box_features_adaptive(integ):
    p = box_with_features(32, 64, 8, integ, True, True, False)
    gpu, cpu = both(p)
    ng, nc = gpu.run(1.0), cpu.run(1.0)
    assert ng == nc and gpu.discardedSteps == cpu.discardedSteps
    compare(p, gpu, cpu, "box adaptive", rel=1e-9)
    # with a non-spatial species the step sequences may separate (see above):
    # the states agree within the solver tolerance
    p = box_with_features(32, 64, 8, integ, True, True, True)
    gpu, cpu = both(p)
    ng, nc = gpu.run(1.0), cpu.run(1.0)
    assert abs(ng - nc) <= 3
    a, b = gpu.getConcentrations(0), cpu.getConcentrations(0)
    assert np.max(np.abs(a - b) / (0.5 * (np.abs(a) + np.abs(b)) + 1e-14)) <= p.max_err_rel


def zero_storage_problem(kind, integ):
    """storage 0 for one species: voxel list (the example disc / ABtoC), a full box
    (tile kernel), and a two-compartment cell with a membrane"""
    if kind in ("single-compartment-diffusion", "ABtoC"):
        p, _, _ = load_golden(kind, integrator=integ)
        zs = {"single-compartment-diffusion": 0, "ABtoC": 2}[kind]
    elif kind == "box":
        p = M.box_problem(M.ABTOC, 32, 64, 6, integrator=integ)
        zs = 2
    else:
        p = M.two_compartment_brusselator(24, 20, 3, integrator=integ)
        zs = 1
    for c in p.compartments[:1] if kind != "cell" else p.compartments:
        inv = np.ones(c.n_species)
        inv[zs] = 0.0
        c.inv_storage = inv
    return p


@pytest.mark.parametrize("kind", ["single-compartment-diffusion", "ABtoC", "box", "cell"])
@pytest.mark.parametrize("integ", INTEGRATORS)
def test_zero_storage_relaxation_fixed_step(kind, integ):
    """PixelSim::solveZeroStorageConstraints (pixelsim.cpp:23-90) before every
    stage: same iteration decisions (max norms are exact), same states"""
    p = zero_storage_problem(kind, integ)
    p.max_err_abs, p.max_err_rel = smeb.DBL_MAX, 0.01   # tolerances of the relaxation
    p.max_timestep = 1e-3
    gpu, cpu = both(p)
    assert abs(gpu.maxStableTimestep - cpu.maxStableTimestep) <= 1e-15 * cpu.maxStableTimestep
    if integ != smeb.RK101:
        gpu.nextTimestep = cpu.nextTimestep = 1e-3
    ng, nc = gpu.run(5e-3), cpu.run(5e-3)
    assert ng == nc and ng >= 1
    compare(p, gpu, cpu, f"zero storage {kind}", rel=1e-11)
    assert gpu.errorMessage() == cpu.errorMessage() == ""


def test_zero_storage_reference_checks_on_gpu():
    """pixelsim_t.cpp:50-131 through the CUDA path"""
    p = zero_storage_problem("ABtoC", smeb.RK101)
    p.max_timestep = 1e-3
    p.device = 0
    gpu = smeb.PixelSim(p)
    assert gpu.errorMessage() == ""
    gpu.run(1e-3)
    assert gpu.errorMessage() == ""
    assert np.all(gpu.getDcdt(0)[:, 2] == 0.0)
    p = zero_storage_problem("single-compartment-diffusion", smeb.RK101)
    p.max_timestep = 1e-3
    p.device = 0
    gpu = smeb.PixelSim(p)
    gpu.run(0.01)
    c = gpu.getConcentrations(0)[:, 0]
    assert c.max() > 1e-12 and (c.max() - c.min()) / c.max() < 0.75


# ---- non-uniform diffusion and cross-diffusion ------------------------------------
def nonuniform_problem(integ, box, seed=5):
    """voxel list (the example disc) or full box with smoothly varying per-voxel
    diffusion constants"""
    rng = np.random.default_rng(seed)
    if box is None:
        p, meta, _ = load_golden("single-compartment-diffusion", integrator=integ)
        p.voxel_size = tuple(meta["voxel_size"])
        D = np.array(meta["compartments"][0]["diffusion_constants"])
    else:
        p = M.box_problem(M.GRAY_SCOTT, *box, integrator=integ, voxel_size=(1.0, 1.5, 2.0))
        p.voxel_size = (1.0, 1.5, 2.0)
        D = np.array(M.GRAY_SCOTT["D"])
    c = p.compartments[0]
    c.diffusion_array = np.ascontiguousarray(
        D[None, :] * rng.uniform(0.0, 1.5, size=(c.n_voxels, c.n_species)))
    return p


@pytest.mark.parametrize("box", [None, (24, 40, 1), (16, 30, 5)])
@pytest.mark.parametrize("integ", INTEGRATORS)
def test_non_uniform_diffusion(box, integ):
    p = nonuniform_problem(integ, box)
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    p.max_timestep = 0.02
    gpu, cpu = both(p)
    assert abs(gpu.maxStableTimestep - cpu.maxStableTimestep) <= 1e-15 * cpu.maxStableTimestep
    assert gpu.run(1e-7) == cpu.run(1e-7) == 1
    compare(p, gpu, cpu, "non-uniform first step")
    ng, nc = gpu.run(0.4), cpu.run(0.4)
    assert ng == nc and ng >= 20
    compare(p, gpu, cpu, "non-uniform run")


def cross_problem(integ, box, expr="(0.3 + 0.5 * sme_var_0)"):
    if box is None:
        p, meta, _ = load_golden("single-compartment-diffusion", integrator=integ)
        p.voxel_size = tuple(meta["voxel_size"])
        raw = np.array(meta["compartments"][0]["diffusion_constants"])
    else:
        p = M.box_problem(M.GRAY_SCOTT, *box, integrator=integ, voxel_size=(1.0, 1.5, 2.0))
        p.voxel_size = (1.0, 1.5, 2.0)
        raw = np.array(M.GRAY_SCOTT["D"])
    c = p.compartments[0]
    c.diffusion_raw = raw
    c.cross_diffusion = [(1, 0, expr), (0, 1, "(0.05 * sme_var_1)")]
    return p


@pytest.mark.parametrize("box", [None, (24, 40, 1), (16, 30, 5)])
@pytest.mark.parametrize("integ", INTEGRATORS)
def test_cross_diffusion(box, integ):
    p = cross_problem(integ, box)
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    p.max_timestep = 0.01
    gpu, cpu = both(p)
    # the bound of the initial state (SimCompartment ctor)
    assert abs(gpu.maxStableTimestep - cpu.maxStableTimestep) <= 1e-15 * cpu.maxStableTimestep
    assert gpu.run(1e-7) == cpu.run(1e-7) == 1
    compare(p, gpu, cpu, "cross-diffusion first step")
    ng, nc = gpu.run(0.2), cpu.run(0.2)
    assert ng == nc and ng >= 20
    compare(p, gpu, cpu, "cross-diffusion run")


def test_cross_diffusion_rk101_follows_dynamic_bound():
    """RK101 steps are capped by the forward-Euler bound of the CURRENT state
    (pixelsim.cpp:122-129): same step counts and states as the CPU solver when
    the bound, not maxTimestep, decides the step"""
    p = cross_problem(smeb.RK101, None, expr="(6.0 * sme_var_0)")
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    gpu, cpu = both(p)
    ng, nc = gpu.run(2.0), cpu.run(2.0)
    assert ng == nc and ng > 20
    assert abs(gpu.maxStableTimestep - cpu.maxStableTimestep) <= 1e-14 * cpu.maxStableTimestep
    compare(p, gpu, cpu, "cross-diffusion RK101 dynamic bound", rel=1e-11)


def test_cross_diffusion_known_answer_on_gpu():
    """simulate_t.cpp:2780-2816 through the CUDA path"""
    p, meta, _ = load_golden("single-compartment-diffusion")
    p.voxel_size = tuple(meta["voxel_size"])
    c = p.compartments[0]
    c.diffusion = c.diffusion.copy()
    c.diffusion[1, :] = 0.0
    c.diffusion_raw = np.array([1.0, 0.0])
    c.cross_diffusion = [(1, 0, "1.0")]
    p.device = 0
    gpu = smeb.PixelSim(p)
    gpu.run(5.0)
    assert gpu.errorMessage() == ""
    out = gpu.getConcentrations(0)
    d = np.abs(out[:, 0] - out[:, 1])
    assert d.mean() < 1e-11
    assert (2.0 * d / (np.abs(out[:, 0] + out[:, 1]) + 1e-15)).mean() < 1e-7


@pytest.mark.parametrize("integ", [smeb.RK212, smeb.RK435])
def test_diffusion_variants_adaptive(integ):
    for p in (nonuniform_problem(integ, None), cross_problem(integ, (16, 30, 5))):
        gpu, cpu = both(p)
        ng, nc = gpu.run(0.5), cpu.run(0.5)
        assert ng == nc and gpu.discardedSteps == cpu.discardedSteps
        compare(p, gpu, cpu, "diffusion variants adaptive", rel=1e-9)


# ---- device-side result statistics -----------------
def avg_min_max_host(conc, n_primary):
    """calculateAvgMinMax (simulate.cpp:272-290) on the downloaded state"""
    out = np.zeros((n_primary, 3))
    for s in range(n_primary):
        col = conc[:, s]
        out[s] = [col.sum() / len(col), min(np.finfo(np.float64).max, col.min()),
                  max(0.0, col.max())]
    return out


@pytest.mark.parametrize("name", ["very-simple-model", "liver-simplified", "txy", "gray-scott-3d"])
def test_avg_min_max_on_device(name):
    p, _, _ = load_golden(name)
    p.device = 0
    gpu = smeb.PixelSim(p)
    for t in (0.0, 0.05):
        if t > 0:
            assert gpu.run(t) > 0
        for ci, c in enumerate(p.compartments):
            dev = gpu.getAvgMinMax(ci)
            host = avg_min_max_host(gpu.getConcentrations(ci), c.n_primary_species)
            assert dev.shape == host.shape
            assert np.array_equal(dev[:, 1:], host[:, 1:])      # min / max: exact
            assert np.allclose(dev[:, 0], host[:, 0], rtol=1e-13, atol=1e-300)


def test_avg_min_max_box():
    p = M.box_problem(M.GRAY_SCOTT, 64, 48, 20)
    p.device = 0
    gpu = smeb.PixelSim(p)
    gpu.run(0.3)
    dev = gpu.getAvgMinMax(0)
    host = avg_min_max_host(gpu.getConcentrations(0), 2)
    assert np.array_equal(dev[:, 1:], host[:, 1:])
    assert np.allclose(dev[:, 0], host[:, 0], rtol=1e-13, atol=0)


# ---- decimated species image -----------------------------
@pytest.mark.parametrize("stride", [(1, 1, 1), (2, 3, 4), (5, 1, 7), (64, 64, 64)])
def test_species_decimated_box(stride):
    nx, ny, nz = 37, 29, 11
    p = M.box_problem(M.GRAY_SCOTT, nx, ny, nz)
    p.device = 0
    gpu = smeb.PixelSim(p)
    gpu.run(0.2)
    full = gpu.getConcentrations(0).reshape(nz, nx, ny, 2)
    sx, sy, sz = stride
    for s in (0, 1):
        img = gpu.getSpeciesDecimated(0, s, stride)
        want = full[::sz, ::sx, ::sy, s]
        assert np.array_equal(img.reshape(want.shape), want)


@pytest.mark.parametrize("k", [1, 3, 1000])
def test_species_decimated_voxel_list(k):
    p, _, _ = load_golden("very-simple-model")
    p.device = 0
    gpu = smeb.PixelSim(p)
    gpu.run(0.01)
    for ci, c in enumerate(p.compartments):
        stride_values = gpu.getConcentrationPadding() + c.n_primary_species
        full = gpu.getConcentrations(ci).reshape(c.n_voxels, -1)
        assert full.shape[1] == stride_values or full.shape[1] == c.n_species
        for s in range(full.shape[1]):
            img = gpu.getSpeciesDecimated(ci, s, (k, 1, 1))
            assert np.array_equal(img.ravel(), full[::k, s])


def test_species_decimated_errors():
    p = M.box_problem(M.GRAY_SCOTT, 8, 8, 2)
    p.device = 0
    gpu = smeb.PixelSim(p)
    with pytest.raises(RuntimeError, match="species index"):
        gpu.getSpeciesDecimated(0, 2, (1, 1, 1))
    with pytest.raises(RuntimeError, match="positive"):
        gpu.getSpeciesDecimated(0, 0, (0, 1, 1))
    with pytest.raises(RuntimeError, match="compartment index"):
        gpu.getSpeciesDecimated(3, 0, (1, 1, 1))
