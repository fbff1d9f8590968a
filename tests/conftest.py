import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


@pytest.fixture(scope="session", autouse=True)
def built_libraries():
    """the product library and the oracle must exist; build them if not"""
    lib = os.path.join(ROOT, "spatial_model_editor_b200", "libsmeb200.so")
    if not os.path.exists(lib):
        subprocess.check_call(
            ["make", "-C", os.path.join(ROOT, "spatial_model_editor_b200", "csrc"),
             "../libsmeb200.so"])
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"],
                          stdout=subprocess.DEVNULL)
    yield
