"""The baseline in baselines/refcuda_port.cu (the reference GPU backend's unfused
kernel sequence, restated) must compute the same RK212 steps as the oracle and as
the product path: otherwise timing it beside ours would compare different work."""
import os

import numpy as np
import pytest

import spatial_model_editor_b200 as smeb
from spatial_model_editor_b200 import models

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _state(nx, ny, nz, seed=1000):
    rng = np.random.RandomState(seed)  # U(0.05, 1.25) like cudapixelsim_t.cpp:121-132
    return rng.uniform(0.05, 1.25, size=(nx * ny * nz, 2))


@pytest.mark.parametrize("shape", [(24, 20, 12), (64, 64, 1), (7, 5, 3)])
def test_port_matches_oracle_and_product(shape):
    if not os.path.exists(os.path.join(ROOT, "baselines", "librefcuda_port.so")):
        pytest.fail("baselines/librefcuda_port.so missing: run __graft_entry__.build()")
    from baselines.refcuda_port import RefCudaPort
    from oracle.oracle_sim import OraclePixelSim
    nx, ny, nz = shape
    dt, n = 0.05, 12
    c0 = _state(nx, ny, nz)
    p = models.box_problem(models.GRAY_SCOTT, nx, ny, nz, conc0=c0, voxel_size=(2.0, 2.0, 2.0),
                           integrator=smeb.RK212, max_err_rel=smeb.DBL_MAX,
                           max_err_abs=smeb.DBL_MAX, max_timestep=dt)
    p.device = 0
    port = RefCudaPort(nx, ny, nz, c0)
    _, err = port.steps(dt, n)
    a = port.get()
    gpu = smeb.PixelSim(p)
    assert not gpu.errorMessage()
    gpu.nextTimestep = dt
    assert gpu.run(n * dt) == n
    b = gpu.getConcentrations(0).reshape(a.shape)
    cpu = OraclePixelSim(p)
    cpu.nextTimestep = dt
    assert cpu.run(n * dt) == n
    c = cpu.getConcentrations(0).reshape(a.shape)
    # the port is compiled with fused multiply-adds (like the reference's NVRTC
    # default); ours and the oracle without: agreement to rounding, not bit-exact
    assert np.max(np.abs(a - c) / np.maximum(np.abs(c), 1e-300)) <= 1e-11
    assert np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)) <= 1e-11
    assert err[0] > 0 and err[1] > 0
    assert port.launches() == 9 * n
    port.close()
