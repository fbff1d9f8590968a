"""The CPU oracle pinned on more of the reference's own known-answer tests for the
Pixel path (scenarios and citations in tests/kat_scenarios.py)."""
import pytest

import kat_scenarios as K
from oracle.oracle_sim import OraclePixelSim

SCENARIOS = [K.uniform_steady_state, K.zero_storage_uniform, K.varying_diffusion_steady_state,
             K.storage_scales_dcdt, K.cross_diffusion_vanishes_for_uniform_source,
             K.voxel_centre_coordinates, K.cross_diffusion_z_only_3d,
             K.rk101_capped_by_cross_diffusion, K.timeout, K.stop_then_continue,
             K.example_3d_models_tiny_step]


@pytest.mark.parametrize("scenario", SCENARIOS, ids=lambda f: f.__name__)
def test_oracle(scenario):
    scenario(OraclePixelSim)


def test_k6_membranes_parallel_matches_serial():
    """the oracle with 1 thread (the reference default) against 2 OpenMP threads"""
    K.membrane_parallel_matches_serial(lambda p: OraclePixelSim(p, threads=1),
                                       lambda p: OraclePixelSim(p, threads=2))
