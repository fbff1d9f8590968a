"""adapter/b200pixelsim.hpp type-checked against the reference's own headers
(tests/adapter_check/README.md): every sme:: member the adapter calls must exist
with a compatible signature, and the class must be a complete BaseSim."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
HERE = os.path.join(ROOT, "tests", "adapter_check")
REF_INCLUDES = ["core/simulate/src", "core/simulate/include", "core/model/include",
                "core/common/include", "core/mesh/include", "core/model/src", "core/common/src"]

pytestmark = pytest.mark.skipif(
    not os.path.isdir(os.path.join(REF, "core/simulate/src")) or shutil.which("g++") is None,
    reason="needs the reference headers under /root/reference and g++")


def syntax_check(adapter_dir):
    cmd = ["g++", "-std=c++20", "-fsyntax-only", "-I", os.path.join(HERE, "stubs"),
           "-I", adapter_dir, "-I", os.path.join(ROOT, "include")]
    for inc in REF_INCLUDES:
        cmd += ["-I", os.path.join(REF, inc)]
    cmd.append(os.path.join(HERE, "drive_adapter.cpp"))
    return subprocess.run(cmd, capture_output=True, text=True)


def test_adapter_type_checks_against_reference_headers():
    r = syntax_check(os.path.join(ROOT, "adapter"))
    errors = [line for line in r.stderr.splitlines() if "error" in line]
    assert r.returncode == 0 and not errors, "\n".join(errors[:20])


def test_check_is_sensitive(tmp_path):
    """negative control: a member that does not exist in geometry::Compartment fails"""
    src = open(os.path.join(ROOT, "adapter", "b200pixelsim.hpp")).read()
    assert "comp->nVoxels()" in src
    (tmp_path / "b200pixelsim.hpp").write_text(src.replace("comp->nVoxels()", "comp->nVoxelsX()"))
    r = syntax_check(str(tmp_path))
    assert r.returncode != 0 and "nVoxelsX" in r.stderr
