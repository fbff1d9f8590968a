"""CPU-solver semantics beyond the reference GPU backend (SURVEY.md section 8(f)
rank 1), pinned on the oracle: reactions that depend on time / space (state
padded with time, x, y, z), non-unit storage, non-spatial species.

Expected values are closed forms of what the reference computes
(pixelsim_impl.cpp:124-147,150-189,312-349,405-423; simulate_t.cpp:2009-2207,
1003-1156), not outputs of this repository.
"""
import math

import numpy as np
import pytest

import spatial_model_editor_b200 as smeb
from oracle.oracle_sim import OraclePixelSim
from util import load_golden

DT = 1e-3  # simulate_t.cpp:2013


def txy(name="txy", integrator=smeb.RK101):
    """simulate_t.cpp:2009-2020: RK101, maxErr = DBL_MAX, maxTimestep = dt"""
    p, meta, _ = load_golden(name, integrator=integrator)
    p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    p.max_timestep = DT
    return p, meta


def centres(comp, ny=100):
    """model_geometry.cpp:1038-1051 with origin 0 and unit voxels"""
    v = comp.voxels.astype(np.float64)
    return v[:, 0] + 0.5, (ny - 1 - v[:, 1]) + 0.5, v[:, 2] + 0.5


def test_txy_fixture_layout():
    """ReacExpr (pixelsim_impl.cpp:124-147): variables = species + time, x, y, z;
    expressions end with 1, 0, 0, 0; padding initialised to 0 and the voxel centres
    (pixelsim_impl.cpp:405-423)"""
    p, _ = txy()
    c = p.compartments[0]
    assert c.species == ["A", "B", "C", "time", "x", "y", "sme_coord_z"]
    assert c.n_primary_species == 3 and p.padding == 4
    assert c.expressions[3:] == ["1.0", "0.0", "0.0", "0.0"]
    assert np.all(c.diffusion[3:] == 0.0)
    x, y, z = centres(c)
    assert np.all(c.conc0[:, :4] == 0.0)
    assert np.array_equal(c.conc0[:, 4], x)
    assert np.array_equal(c.conc0[:, 5], y)
    assert np.array_equal(c.conc0[:, 6], z)
    # dA/dt = x, dB/dt = y, dC/dt = time
    assert c.expressions[0].replace(" ", "") in ("sme_var_4", "(sme_var_4)")
    assert c.expressions[1].replace(" ", "") in ("sme_var_5", "(sme_var_5)")
    assert c.expressions[2].replace(" ", "") in ("sme_var_3", "(sme_var_3)")


@pytest.mark.parametrize("name,sa,sc", [("txy", 1.0, 1.0), ("txy-storage", 0.5, 2.0)])
def test_txy_euler_closed_form(name, sa, sc):
    """forward Euler from zero: after one step A = dt x / storage_A, B = dt y (the
    stencil of a zero field vanishes), C = 0; after n steps C = dt^2 n (n - 1) / 2 /
    storage_C and the time component is n dt; x, y, z never change"""
    p, _ = txy(name)
    sim = OraclePixelSim(p)
    assert sim.getConcentrationPadding() == 4
    x, y, z = centres(p.compartments[0])
    assert sim.run(DT) == 1
    c = sim.getConcentrations(0)
    assert c.shape == (3149, 7)
    assert np.max(np.abs(c[:, 0] - DT * x / sa)) < 1e-15
    assert np.max(np.abs(c[:, 1] - DT * y)) < 1e-15
    assert np.all(c[:, 2] == 0.0)
    assert np.all(c[:, 3] == DT)
    n = 1
    for _ in range(9):
        n += sim.run(DT)
    c = sim.getConcentrations(0)
    assert n == 10
    assert np.allclose(c[:, 2], DT * DT * n * (n - 1) / 2 / sc, rtol=1e-13, atol=0)
    assert np.allclose(c[:, 3], n * DT, rtol=1e-14, atol=0)
    assert np.array_equal(c[:, 4], x) and np.array_equal(c[:, 5], y)
    assert np.array_equal(c[:, 6], z)
    # dc/dt of the last stage: dC/dt = time at the start of that step / storage
    d = sim.getDcdt(0)
    assert np.allclose(d[:, 2], (n - 1) * DT / sc, rtol=1e-13, atol=0)
    assert np.all(d[:, 3] == 1.0) and np.all(d[:, 4:] == 0.0)


@pytest.mark.parametrize("integrator", [smeb.RK212, smeb.RK323, smeb.RK435])
def test_txy_higher_order_integrators_are_exact_for_c(integrator):
    """dC/dt = t is integrated exactly by every scheme of order >= 2: C = t^2 / 2"""
    p, _ = txy(integrator=integrator)
    sim = OraclePixelSim(p)
    sim.nextTimestep = DT
    assert sim.run(20 * DT) == 20
    c = sim.getConcentrations(0)
    t = 20 * DT
    assert np.allclose(c[:, 3], t, rtol=1e-13, atol=0)
    assert np.allclose(c[:, 2], t * t / 2, rtol=1e-12, atol=0)


def test_clamp_skips_padding():
    """clampNegativeConcentrations touches the model species only
    (pixelsim_impl.cpp:643-654): negative coordinates survive"""
    p, _ = txy()
    c = p.compartments[0]
    c.conc0 = c.conc0.copy()
    c.conc0[:, 4] -= 60.0  # as if the geometry origin were at x = -60
    c.conc0[:, 0] = 1.0
    sim = OraclePixelSim(p)
    sim.run(DT)
    out = sim.getConcentrations(0)
    assert np.array_equal(out[:, 4], c.conc0[:, 4]) and out[:, 4].min() < 0
    # dA/dt = x < 0 drives A below its start, but never below zero
    for _ in range(40):
        sim.run(50 * DT)
    assert sim.getConcentrations(0)[:, 0].min() >= 0.0


def test_storage_rescales_time():
    """dc/dt *= 1 / storage (pixelsim_impl.cpp:165-189): with storage s the state at
    time t is the unit-storage state at t / s (pure diffusion + fixed step);
    forward-Euler bound uses D / storage (pixelsim_impl.cpp:249-255)"""
    ps, _, _ = load_golden("single-compartment-diffusion-3d-storage", integrator=smeb.RK101)
    pu, meta, _ = load_golden("single-compartment-diffusion-3d", integrator=smeb.RK101)
    assert ps.compartments[0].inv_storage.tolist() == [0.5, 1.0 / 3.0]
    for p in (ps, pu):
        p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
    # species 0: storage 2 -> dt_s = 2 dt_u gives identical Euler sequences
    pu.max_timestep, ps.max_timestep = 0.05, 0.1
    su, ss = OraclePixelSim(pu), OraclePixelSim(ps)
    D = meta["compartments"][0]["diffusion_constants"]
    dx = meta["voxel_size"][0]
    bound = lambda d: 1.0 / (2.0 * 3.0 * d / (dx * dx))
    assert math.isclose(su.maxStableTimestep, min(bound(D[0]), bound(D[1])), rel_tol=1e-14)
    assert math.isclose(ss.maxStableTimestep, min(bound(D[0] / 2), bound(D[1] / 3)),
                        rel_tol=1e-14)
    assert su.run(0.25) == 5 and ss.run(0.5) == 5
    a, b = su.getConcentrations(0)[:, 0], ss.getConcentrations(0)[:, 0]
    assert np.max(np.abs(a - b)) <= 1e-13 * np.max(np.abs(a))


def test_storage_heat_kernel_3d():
    """simulate_t.cpp:1003-1156 "non-unit storage" (2, 3): after two intervals of
    10 the Gaussian follows the heat kernel at t / storage inside r < 16 (max 4 %,
    mean 1.5 %), total amount conserved to 1e-9 after each interval (cubic 2x2x2
    voxels here; the reference also stretches them)"""
    p, meta, _ = load_golden("single-compartment-diffusion-3d-storage")
    p.max_err_abs, p.max_err_rel = smeb.DBL_MAX, 0.01
    comp = p.compartments[0]
    D = meta["compartments"][0]["diffusion_constants"]
    dx, dy, dz = meta["voxel_size"]
    ox, oy, oz = meta["origin"]
    nz, ny, nx = 50, 50, 50
    v = comp.voxels.astype(np.float64)
    x = ox + dx * (v[:, 0] + 0.5)
    y = oy + dy * ((ny - 1 - v[:, 1]) + 0.5)
    z = oz + dz * (v[:, 2] + 0.5)
    r2 = x * x + y * y + z * z
    total0 = comp.conc0.sum(axis=0)
    sim = OraclePixelSim(p, threads=0)
    t = 0.0
    for _ in range(2):
        sim.run(10.0)
        t += 10.0
        c = sim.getConcentrations(0)
        for s in range(2):
            assert abs(c[:, s].sum() - total0[s]) / total0[s] < 1e-9
    inside = r2 < 16 * 16
    for s, storage in enumerate((2.0, 3.0)):
        t0 = 36.0 / 4.0 / D[s]
        te = t / storage
        ana = (t0 / (te + t0)) ** 1.5 * np.exp(-r2 / (4.0 * D[s] * (te + t0)))
        rel = np.abs(c[:, s] - ana)[inside] / ana[inside]
        assert rel.max() < 0.04 and rel.mean() < 0.015, (s, rel.max(), rel.mean())


def test_non_spatial_species_are_averaged():
    """SimCompartment::spatiallyAverageDcdt (pixelsim_impl.cpp:150-163): dc/dt of a
    non-spatial species is the spatial mean of the full dc/dt (reactions +
    diffusion + membrane fluxes) of that species; other species are untouched"""
    pn, _, _ = load_golden("very-simple-model-non-spatial-species", integrator=smeb.RK101)
    pf, _, _ = load_golden("very-simple-model", integrator=smeb.RK101)
    assert [list(c.non_spatial) for c in pn.compartments] == [[], [0], [1]]
    rng = np.random.default_rng(7)
    for p in (pn, pf):
        p.max_err_abs = p.max_err_rel = smeb.DBL_MAX
        p.max_timestep = 1e-3
    for cn, cf in zip(pn.compartments, pf.compartments):
        cn.conc0 = rng.uniform(0.05, 1.25, size=cn.conc0.shape)
        cf.conc0 = cn.conc0.copy()
    sn, sf = OraclePixelSim(pn), OraclePixelSim(pf)
    sn.run(1e-3)
    sf.run(1e-3)
    for ci, comp in enumerate(pn.compartments):
        dn, df = sn.getDcdt(ci), sf.getDcdt(ci)
        for s in range(comp.n_species):
            if s in comp.non_spatial:
                assert np.all(dn[:, s] == dn[0, s])
                assert math.isclose(dn[0, s], df[:, s].mean(), rel_tol=1e-12)
            else:
                assert np.array_equal(dn[:, s], df[:, s])
    # a non-spatial species that starts uniform stays uniform
    pn2, _, _ = load_golden("very-simple-model-non-spatial-species")
    s2 = OraclePixelSim(pn2)
    s2.run(0.5)
    c = s2.getConcentrations(1)
    assert np.ptp(c[:, 0]) <= 1e-12 * abs(c[0, 0])


# ---- non-uniform diffusion and cross-diffusion (SURVEY.md section 8(a) rows a7', a13) ----
def diffusion_disc(integrator=smeb.RK101):
    p, meta, _ = load_golden("single-compartment-diffusion", integrator=integrator)
    p.voxel_size = tuple(meta["voxel_size"])
    return p, meta


def with_array_diffusion(p, meta, scale=None):
    """Field::setDiffusionConstant(array): per-voxel constants, here the uniform
    value at every voxel (optionally modulated)"""
    c = p.compartments[0]
    D = np.array(meta["compartments"][0]["diffusion_constants"])
    arr = np.repeat(D[None, :], c.n_voxels, axis=0)
    if scale is not None:
        arr = arr * scale[:, None]
    c.diffusion_array = np.ascontiguousarray(arr)
    return p


def test_k4_uniform_matches_array_diffusion():
    """pixelsim_t.cpp:21-49: the uniform stencil and the face-averaged array stencil
    agree to 1e-14 after one RK101 run of 1e-3 (first step 1e-3, maxTimestep 1e-3)"""
    pu, meta = diffusion_disc()
    pa, _ = diffusion_disc()
    with_array_diffusion(pa, meta)
    for p in (pu, pa):
        p.max_timestep = 1e-3
    su, sa = OraclePixelSim(pu), OraclePixelSim(pa)
    assert su.run(1e-3) == sa.run(1e-3) == 1
    cu, ca = su.getConcentrations(0), sa.getConcentrations(0)
    assert np.all(np.abs(cu - ca) <= 1e-14 * np.maximum(np.abs(cu), np.abs(ca)))
    assert math.isclose(su.maxStableTimestep, sa.maxStableTimestep, rel_tol=1e-15)


def test_array_diffusion_zero_region_blocks_flux():
    """simulate_t.cpp:1304-1398 restated: with D = 0 in the left half the field there
    does not move (the face average with a D = 1 neighbour is 1/2, so only the
    interface column changes); the total amount is conserved"""
    p, meta = diffusion_disc(smeb.RK212)
    c = p.compartments[0]
    left = c.voxels[:, 0] < 50
    with_array_diffusion(p, meta, scale=np.where(left, 0.0, 1.0))
    p.max_err_abs, p.max_err_rel = smeb.DBL_MAX, 0.001
    total0 = c.conc0.sum(axis=0)
    sim = OraclePixelSim(p)
    sim.run(5.0)
    out = sim.getConcentrations(0)
    interior = c.voxels[:, 0] + 2 < 50
    assert np.max(np.abs(out[interior] - c.conc0[interior])) < 1e-8
    assert np.max(np.abs(out[~left] - c.conc0[~left])) > 1e-3
    assert np.all(np.abs(out.sum(axis=0) - total0) <= 1e-9 * total0)


def cross_diffusion_disc(integrator=None):
    """simulate_t.cpp:2780-2790: D_fast = 0, cross-diffusion constant 1 from slow"""
    p, meta, _ = load_golden("single-compartment-diffusion", integrator=integrator)
    p.voxel_size = tuple(meta["voxel_size"])
    c = p.compartments[0]
    c.diffusion = c.diffusion.copy()
    c.diffusion[1, :] = 0.0
    c.diffusion_raw = np.array([meta["compartments"][0]["diffusion_constants"][0], 0.0])
    c.cross_diffusion = [(1, 0, "1.0")]
    return p, meta


def test_cross_diffusion_equivalent_to_diffusion():
    """simulate_t.cpp:2780-2816: "fast" (D = 0, cross-diffusion 1 from "slow", D_slow =
    1, equal initial fields) follows "slow": average |slow - fast| < 1e-11, average
    relative difference < 1e-7 at t = 0 and t = 5"""
    p, meta = cross_diffusion_disc()
    assert meta["compartments"][0]["diffusion_constants"][0] == 1.0
    sim = OraclePixelSim(p)
    for t in (0.0, 5.0):
        if t > 0:
            sim.run(t)
        c = sim.getConcentrations(0)
        d = np.abs(c[:, 0] - c[:, 1])
        assert d.mean() < 1e-11
        assert (2.0 * d / (np.abs(c[:, 0] + c[:, 1]) + 1e-15)).mean() < 1e-7
    assert sim.errorMessage() == ""


def test_cross_diffusion_dynamic_stability_bound():
    """updateCrossDiffusionMaxStableTimestep (pixelsim_impl.cpp:506-543): bound from
    max(D_ii) + sum of max |D_ij| per target; with a state-dependent coefficient it
    follows the state"""
    p, meta = cross_diffusion_disc(smeb.RK101)
    sim = OraclePixelSim(p)
    # slow: D = 1; fast: 0 + |1| -> 1 / (2 * (1 + 1)) in 2-D unit voxels, z term included
    assert math.isclose(sim.maxStableTimestep, 1.0 / (2.0 * 3.0), rel_tol=1e-15)
    p2, _ = cross_diffusion_disc(smeb.RK101)
    p2.compartments[0].cross_diffusion = [(1, 0, "(4.0 * sme_var_0)")]
    s2 = OraclePixelSim(p2)
    cmax = p2.compartments[0].conc0[:, 0].max()
    assert math.isclose(s2.maxStableTimestep, 1.0 / (2.0 * 3.0 * 4.0 * cmax), rel_tol=1e-14)
    s2.run(0.5)  # RK101 steps are capped by the bound of the current state
    cmax2 = s2.getConcentrations(0)[:, 0].max()
    assert cmax2 < cmax
    assert s2.maxStableTimestep > 1.0 / (2.0 * 3.0 * 4.0 * cmax)


# ---- zero-storage species: algebraic constraints by pseudo-time relaxation --------
def zero_storage_problem(name, species_index, integrator=smeb.RK101, d_zero=False):
    p, meta, _ = load_golden(name, integrator=integrator)
    c = p.compartments[0]
    inv = np.ones(c.n_species)
    inv[species_index] = 0.0      # Species::setStorage(id, 0)
    c.inv_storage = inv
    if d_zero:
        c.diffusion = c.diffusion.copy()
        c.diffusion[species_index, :] = 0.0
    p.max_timestep = 1e-3
    return p


@pytest.mark.parametrize("d_zero", [False, True])
def test_zero_storage_accepted_and_dcdt_zero(d_zero):
    """pixelsim_t.cpp:50-91: ABtoC with storage(C) = 0 (and D_C = 0): one RK101 run of
    1e-3 without error, dc/dt of C is zero after applyStorage"""
    p = zero_storage_problem("ABtoC", 2, d_zero=d_zero)
    sim = OraclePixelSim(p)
    assert sim.errorMessage() == ""
    sim.run(1e-3)
    assert sim.errorMessage() == ""
    assert np.all(sim.getDcdt(0)[:, 2] == 0.0)
    # A and B are ordinary species: they did change
    assert np.any(sim.getConcentrations(0)[:, 0] != p.compartments[0].conc0[:, 0])


def test_zero_storage_constraint_solver_reduces_residual():
    """pixelsim_t.cpp:92-131: single-compartment-diffusion with storage(slow) = 0:
    after run(0.01) "slow" is more uniform (relative range < 0.75); the other
    species evolves as without the constraint"""
    p = zero_storage_problem("single-compartment-diffusion", 0)
    sim = OraclePixelSim(p)
    sim.run(0.01)
    assert sim.errorMessage() == ""
    c = sim.getConcentrations(0)
    lo, hi = c[:, 0].min(), c[:, 0].max()
    c0 = p.compartments[0].conc0[:, 0]
    assert (c0.max() - c0.min()) / c0.max() > 0.99
    assert hi > 1e-12 and (hi - lo) / hi < 0.75
    # maxStableTimestep ignores the zero-storage species (pixelsim_impl.cpp:238-255)
    pu, meta, _ = load_golden("single-compartment-diffusion", integrator=smeb.RK101)
    D_fast = meta["compartments"][0]["diffusion_constants"][1]
    assert math.isclose(sim.maxStableTimestep, 1.0 / (2.0 * 3.0 * D_fast), rel_tol=1e-14)


@pytest.mark.parametrize("integrator", [smeb.RK212, smeb.RK323, smeb.RK435])
def test_zero_storage_adaptive_integrators_run(integrator):
    """the relaxation runs before every stage (pixelsim.cpp:146,155,196); the
    zero-storage species is excluded from the error norm (pixelsim_impl.cpp:778-781)"""
    p = zero_storage_problem("single-compartment-diffusion", 0, integrator)
    p.max_timestep = smeb.DBL_MAX
    sim = OraclePixelSim(p)
    n = sim.run(0.05)
    assert sim.errorMessage() == "" and n > 3
    c = sim.getConcentrations(0)
    assert np.all(np.isfinite(c))
    if integrator == smeb.RK212:
        # RK212 keeps the relaxed values (s3 is taken after the relaxation,
        # pixelsim.cpp:146-154); RK323 / RK435 mix them with the pre-relaxation
        # state again (doRKInit runs before the first relaxation, pixelsim.cpp:168)
        assert (c[:, 0].max() - c[:, 0].min()) / c[:, 0].max() < 0.75
