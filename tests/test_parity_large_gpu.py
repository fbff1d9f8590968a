"""Parity at BASELINE.json's own sizes: the CUDA path against the CPU oracle (all
host threads) on the shapes the bench runs — many z-chunks per launch, thousands of
CTAs, byte offsets beyond 2^31, the 2-D brick march, run tables of millions of
voxels.  A few fixed steps each, <= 1e-12 relative; the adaptive case must take the
same accepted and discarded steps as the oracle."""
import os
import sys

import numpy as np
import pytest

import spatial_model_editor_b200 as smeb
from spatial_model_editor_b200 import models as M
from oracle.oracle_sim import OraclePixelSim
from util import assert_parity

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _threads():
    return max(1, len(os.sched_getaffinity(0)))


def _available_gib():
    try:
        import psutil
        return psutil.virtual_memory().available / 2 ** 30
    except Exception:  # noqa: BLE001
        return 0.0


def _compare_fixed(p, dt, steps, rel=1e-12):
    p.device = 0
    gpu = smeb.PixelSim(p)
    assert gpu.errorMessage() == ""
    cpu = OraclePixelSim(p, threads=_threads())
    gpu.nextTimestep = cpu.nextTimestep = dt
    assert gpu.run(steps * dt) == steps
    assert cpu.run(steps * dt) == steps
    worst = 0.0
    for i in range(len(p.compartments)):
        worst = max(worst, assert_parity(gpu.getConcentrations(i), cpu.getConcentrations(i),
                                         rel=rel, what=f"compartment {i}"))
    launches = gpu.launchCount()
    gpu.close()
    return worst, launches


def test_c4_512_cubed_rk212_matches_oracle():
    """BASELINE config 4 at full size (512 x 512 x 512, S = 2: 2 GiB per buffer, 8
    z-chunks x 256 tiles = 2048 CTAs per launch); 128 planes if the host is short of
    memory for the oracle's tables"""
    import bench
    nz = 512 if _available_gib() > 64.0 else 128
    p = bench.make_problem(512, 512, nz, smeb.RK212)
    _, launches = _compare_fixed(p, bench.DT, 3)
    assert launches in (6, 7)  # 3 steps x 2 stage kernels (+ the one-warp flush of the
    # last step's error norms)


def test_c2_4096_squared_rk212_matches_oracle():
    """BASELINE config 2: gray-scott on the 4096 x 4096 full box (2-D brick march)"""
    import bench
    p = bench.make_problem(4096, 4096, 1, smeb.RK212)
    _compare_fixed(p, bench.DT, 5)


def test_rk435_256_cubed_matches_oracle():
    """the five-stage integrator (running sum s2 streamed through HBM) at 256^3"""
    import bench
    p = bench.make_problem(256, 256, 256, smeb.RK435)
    _compare_fixed(p, bench.DT, 3)


def test_c3_two_compartment_brusselator_2048_rk323_adaptive():
    """BASELINE config 3 at 2048^2: two voxel-list compartments (run-compressed
    neighbour codes) coupled by a membrane, RK323 adaptive at relTol 0.005 from the
    reference's random-state protocol: same accepted / discarded step counts, state
    within 1e-9 (dozens of adaptive steps)"""
    p = M.two_compartment_brusselator(2048, 2048, integrator=smeb.RK323, max_err_rel=0.005)
    p.device = 0
    gpu = smeb.PixelSim(p)
    assert gpu.errorMessage() == ""
    assert gpu.neighbourBytesPerVoxel(0) == 4
    cpu = OraclePixelSim(p, threads=_threads())
    t = 1.0
    ng, nc = gpu.run(t), cpu.run(t)
    assert ng == nc and ng > 5, (ng, nc)
    assert gpu.discardedSteps == cpu.discardedSteps
    for i in range(2):
        assert_parity(gpu.getConcentrations(i), cpu.getConcentrations(i), rel=1e-9,
                      what=f"compartment {i}")
    # and fixed steps at 1e-12
    p2 = M.two_compartment_brusselator(2048, 2048, integrator=smeb.RK323,
                                       max_err_rel=smeb.DBL_MAX, max_err_abs=smeb.DBL_MAX,
                                       max_timestep=0.01)
    _compare_fixed(p2, 0.01, 3)


def test_c2_disc_4096_voxel_list_matches_oracle():
    """gray-scott on the example disc upscaled to 4096^2 (N = 5.28 M voxel list)"""
    import argparse
    import bench
    args = argparse.Namespace(size=4096, adaptive=False, dt=0.25, workload="c2disc")
    p, _ = bench.make_listed_problem(args, smeb.RK212, 0)
    _compare_fixed(p, 0.25, 3)


@pytest.mark.parametrize("workload,size,integ,dt", [("c3", 1024, smeb.RK323, 0.01),
                                                    ("c2disc", 2048, smeb.RK212, 0.25),
                                                    ("c3", 1024, smeb.RK435, 0.01)])
def test_segment_staged_kernel_matches_oracle(workload, size, integ, dt):
    """kernel_variant = 4: the voxel-list stage kernel that stages pieces of runs in
    shared memory with bulk copies (smeb_stage_seg) — same bits as the oracle's, and
    the same bits as the direct kernel's"""
    import argparse
    import bench
    args = argparse.Namespace(size=size, adaptive=False, dt=dt, workload=workload)
    p, _ = bench.make_listed_problem(args, integ, 0)
    p.kernel_variant = 4
    gpu = smeb.PixelSim(p)
    assert gpu.errorMessage() == ""
    assert "smeb_stage_seg" in gpu.dominantKernel()
    gpu.close()
    _compare_fixed(p, dt, 3)
    q, _ = bench.make_listed_problem(args, integ, 0)
    q.device = 0
    a, b = smeb.PixelSim(p), smeb.PixelSim(q)
    assert "smeb_stage_nn" in b.dominantKernel()
    a.nextTimestep = b.nextTimestep = dt
    assert a.run(4 * dt) == b.run(4 * dt) == 4
    for i in range(len(p.compartments)):
        assert np.array_equal(a.getConcentrations(i), b.getConcentrations(i))
        assert np.array_equal(a.getDcdt(i), b.getDcdt(i))
