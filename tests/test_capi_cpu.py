"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a
GPU, exports every symbol include/smeb200.h declares, NVRTC-compiles model
modules for sm_100a, and — with no device — refuses to compute (no fallback)."""
import ctypes
import os
import re

import pytest

import spatial_model_editor_b200 as smeb
from spatial_model_editor_b200 import models as M

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "smeb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(smeb200_[a-z0-9_]+)\s*\(", text)))


def test_exports_every_declared_symbol():
    lib = ctypes.CDLL(os.path.join(ROOT, "spatial_model_editor_b200", "libsmeb200.so"))
    names = declared_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), f"libsmeb200.so does not export {name}"


def test_struct_layout_matches_header(tmp_path):
    """ctypes mirrors vs the real header, measured with the C compiler"""
    import subprocess
    from spatial_model_editor_b200.problem import CCompartment, CMembrane, CProblem
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "smeb200.h"\n'
                   'int main(void){printf("%zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(smeb200_compartment),sizeof(smeb200_membrane),'
                   'sizeof(smeb200_problem),offsetof(smeb200_problem,max_err_abs),'
                   'offsetof(smeb200_problem,device),'
                   'offsetof(smeb200_membrane,flux_length));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    assert got == [ctypes.sizeof(CCompartment), ctypes.sizeof(CMembrane),
                   ctypes.sizeof(CProblem), CProblem.max_err_abs.offset,
                   CProblem.device.offset, CMembrane.flux_length.offset]


@pytest.mark.parametrize("integ", [smeb.RK101, smeb.RK212, smeb.RK323, smeb.RK435])
def test_nvrtc_compiles_model_modules(integ):
    for p in (M.box_problem(M.GRAY_SCOTT, 8, 8, 2, integrator=integ),
              M.two_compartment_brusselator(12, 12, 1, integrator=integ),
              M.labelled_problem(M.disc_labels(20, 20),
                                 [dict(label=1, id="comp", kin=M.ABTOC)], integrator=integ)):
        sim = smeb.PixelSim(p, compile_only=True)
        assert sim.errorMessage() == ""
        src = sim.kernelSource()
        # generated functors use the reference's relabelled variable names
        # (cudapixelsim_t.cpp:284-297 checks the same)
        assert "sme_var_0" in src and "struct Reac0" in src
        assert sim.run(1.0) == 0


def test_bad_expression_reports_nvrtc_error():
    p = M.box_problem(M.GRAY_SCOTT, 4, 4, 1)
    p.compartments[0].expressions[0] = "sme_var_0 +* 2"
    sim = smeb.PixelSim(p, compile_only=True)
    assert "NVRTC" in sim.errorMessage()


def test_unsupported_feature_messages():
    """same substrings the reference tests look for (cudapixelsim_t.cpp:389-411)"""
    from spatial_model_editor_b200 import problem as P
    for bit, text in [(P.HAS_CROSS_DIFFUSION, "cross-diffusion"),
                      (P.HAS_NON_UNIT_STORAGE, "unit storage"),
                      (P.HAS_NON_UNIFORM_DIFFUSION, "non-uniform diffusion"),
                      (P.HAS_NON_SPATIAL_SPECIES, "non-spatial")]:
        p = M.box_problem(M.GRAY_SCOTT, 4, 4, 1)
        p.unsupported = bit
        sim = smeb.PixelSim(p)
        assert text in sim.errorMessage()
        assert sim.run(1.0) == 0


def test_no_gpu_means_error_not_fallback():
    if smeb.device_count() > 0:
        pytest.skip("a GPU is present")
    sim = smeb.PixelSim(M.box_problem(M.GRAY_SCOTT, 4, 4, 1))
    msg = sim.errorMessage()
    # prefixes the reference's tests treat as "GPU runtime unavailable"
    # (cudapixelsim_t.cpp:150-156)
    assert msg.startswith(("Failed to initialize the CUDA driver", "No CUDA devices were found",
                           "Failed to query CUDA devices"))
    assert sim.run(1.0) == 0
