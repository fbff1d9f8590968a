"""smeb200_avg_min_max_no_alloc (include/smeb200.h): the entry point that allocates
nothing, checked through the C-ABI against smeb200_get_avg_min_max on two calls in
a row (they share one file-scope scratch array on the device).  Collected last on
purpose."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import spatial_model_editor_b200 as smeb
from spatial_model_editor_b200 import models as M
from spatial_model_editor_b200.pixelsim import LIB_PATH
from util import load_golden

MAX_SPECIES, MAX_VALUES, MAX_VOXELS = 16, 64, 2 ** 31


def no_alloc_fn():
    smeb.load_library()
    fn = C.CDLL(LIB_PATH).smeb200_avg_min_max_no_alloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p,
                   C.c_void_p, C.c_char_p, C.c_size_t]
    return fn


def test_sizes_beyond_the_maximum_are_rejected():
    """argument checks come before any CUDA call: no device needed"""
    fn = no_alloc_fn()
    err = C.create_string_buffer(256)
    for n, s, sp, word in [(10, 4, MAX_SPECIES + 1, "n_primary_species"),
                           (10, 4, 0, "n_primary_species"),
                           (10, MAX_VALUES + 1, 2, "values_per_voxel"),
                           (10, 1, 2, "values_per_voxel"),
                           (MAX_VOXELS + 1, 2, 2, "n_voxels")]:
        assert fn(None, n, s, sp, None, None, err, len(err)) == 1
        assert word in err.value.decode(), err.value
    # the message buffer is optional
    assert fn(None, 10, 4, MAX_SPECIES + 1, None, None, None, 0) == 1


def test_source_calls_no_allocating_routine():
    """the guarantee is structural: outside comments, csrc/no_alloc.cu names no
    allocator and resolves its scratch array with cudaGetSymbolAddress"""
    src = open(os.path.join(os.path.dirname(LIB_PATH), "csrc", "no_alloc.cu")).read()
    code = re.sub(r"//[^\n]*|/\*.*?\*/", "", src, flags=re.S)
    for word in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cuMemAlloc", "malloc(",
                 "new ", "std::vector", "std::string", "cudaMemset", "cudaMemcpy"):
        assert word not in code, word
    assert "static __device__ __align__(256) double g_scratch" in code
    assert "cudaGetSymbolAddress(&p, g_scratch)" in code


@pytest.mark.gpu
def test_no_alloc_matches_get_avg_min_max_twice_in_a_row():
    import torch
    fn = no_alloc_fn()
    err = C.create_string_buffer(256)
    torch.cuda.set_device(0)

    # (simulator, compartment): a box (TMA tile kernels), voxel lists with t,x,y
    # padding behind the species (values_per_voxel > n_primary_species)
    cases = []
    p = M.box_problem(M.GRAY_SCOTT, 64, 48, 20)
    p.device = 0
    box = smeb.PixelSim(p)
    assert box.errorMessage() == "", box.errorMessage()
    assert box.run(0.3) > 0
    cases.append((box, p, 0))
    q, _, _ = load_golden("txy")
    q.device = 0
    txy = smeb.PixelSim(q)
    assert txy.errorMessage() == "", txy.errorMessage()
    assert txy.run(0.05) > 0
    cases += [(txy, q, ci) for ci in range(len(q.compartments))]

    torch.cuda.set_device(0)
    # set-up only: resolves the scratch array's address on this device
    assert fn(None, 0, 2, 2, None, None, err, len(err)) == 0, err.value
    stream = torch.cuda.current_stream().cuda_stream

    def call(sim, prob, ci):
        host = np.array(sim.getConcentrations(ci), dtype=np.float64, order="C")  # a copy
        n, s = host.shape
        sp = prob.compartments[ci].n_primary_species or s
        d_state = torch.from_numpy(host).cuda()
        d_out = torch.full((3 * sp,), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        rc = fn(d_state.data_ptr(), n, s, sp, d_out.data_ptr(), stream, err, len(err))
        assert rc == 0, err.value
        torch.cuda.synchronize()
        got = d_out.cpu().numpy().reshape(sp, 3)
        want = sim.getAvgMinMax(ci)
        assert got.shape == want.shape
        assert np.array_equal(got[:, 1:], want[:, 1:])                 # min / max: exact
        assert np.allclose(got[:, 0], want[:, 0], rtol=1e-12, atol=0)  # averages

    # two passes over all cases: consecutive calls differ in size and species
    # count, so each one meets the previous call's contents in the scratch array
    for _ in range(2):
        for sim, prob, ci in cases:
            call(sim, prob, ci)
    # the same input twice in a row
    call(box, p, 0)
    call(box, p, 0)
