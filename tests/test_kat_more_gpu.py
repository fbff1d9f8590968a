"""The reference's known-answer tests of tests/kat_scenarios.py through the CUDA
path (C-ABI), each simulator paired with the CPU oracle on the same problem: every
run() must take the same number of steps and every state / dc/dt read must agree."""
import functools

import pytest

import kat_scenarios as K
import spatial_model_editor_b200 as smeb
from oracle.oracle_sim import OraclePixelSim
from util import assert_parity

pytestmark = pytest.mark.gpu


class Paired:
    """the CUDA simulator with the oracle beside it"""

    def __init__(self, p, rel=1e-12, same_steps=True):
        p.device = 0
        self.gpu = smeb.PixelSim(p)
        self.cpu = OraclePixelSim(p)
        self.rel, self.same_steps = rel, same_steps
        self.launches = 0

    def run(self, *args):
        before = self.gpu.launchCount()
        ng, nc = self.gpu.run(*args), self.cpu.run(*args)
        self.launches += self.gpu.launchCount() - before
        if self.same_steps:
            assert ng == nc, (ng, nc)
            assert self.gpu.discardedSteps == self.cpu.discardedSteps
        return ng

    def _both(self, name, *args):
        a, b = getattr(self.gpu, name)(*args), getattr(self.cpu, name)(*args)
        if self.same_steps:
            assert_parity(a, b, rel=self.rel, what=name)
        return a

    def getConcentrations(self, i):
        return self._both("getConcentrations", i)

    def getDcdt(self, i):
        return self._both("getDcdt", i)

    def errorMessage(self):
        assert self.gpu.errorMessage() == self.cpu.errorMessage()
        return self.gpu.errorMessage()

    def getConcentrationPadding(self):
        return self.gpu.getConcentrationPadding()

    def getStopRequested(self):
        return self.gpu.getStopRequested()

    def setStopRequested(self, stop):
        self.gpu.setStopRequested(stop)
        self.cpu.setStopRequested(stop)


FIXED = functools.partial(Paired, rel=1e-12)      # RK101 / few steps
ADAPTIVE = functools.partial(Paired, rel=1e-9)    # hundreds of adaptive steps
WALL_CLOCK = functools.partial(Paired, same_steps=False)  # the deadline decides the steps

CASES = [
    (K.uniform_steady_state, ADAPTIVE),
    (functools.partial(K.zero_storage_uniform, tol=1e-7, t=0.5), ADAPTIVE),
    (K.varying_diffusion_steady_state, ADAPTIVE),
    (K.storage_scales_dcdt, FIXED),
    (K.cross_diffusion_vanishes_for_uniform_source, FIXED),
    (K.voxel_centre_coordinates, FIXED),
    (K.cross_diffusion_z_only_3d, FIXED),
    (K.rk101_capped_by_cross_diffusion, FIXED),
    (K.timeout, WALL_CLOCK),
    (K.stop_then_continue, ADAPTIVE),
    (K.example_3d_models_tiny_step, ADAPTIVE),
    # K5 with the model's event at t = 25 (simulator re-created with k2 = 4)
    (K.brusselator_rk_orders_agree, ADAPTIVE),
]


def _name(case):
    f = case[0]
    return getattr(f, "__name__", None) or f.func.__name__


@pytest.mark.parametrize("case", CASES, ids=_name)
def test_cuda_path(case):
    scenario, make = case
    made = []

    def tracked(p):
        made.append(make(p))
        return made[-1]

    scenario(tracked)
    assert made and all(m.launches > 0 for m in made)


def test_k6_membranes_cuda_matches_serial_cpu():
    """K6 with the CUDA path in the role of the multi-threaded solver: one thread per
    membrane voxel, fluxes accumulated in the reference's order without atomics"""
    def cuda(p):
        p.device = 0
        return smeb.PixelSim(p)

    K.membrane_parallel_matches_serial(lambda p: OraclePixelSim(p, threads=1), cuda)
