"""TEST INFRASTRUCTURE — Python handle on the CPU oracle (liboracle.so).

Same method names as the product's ``PixelSim`` so parity tests read
``gpu.getConcentrations(i)`` vs ``cpu.getConcentrations(i)`` like the
reference's own cudapixelsim_t.cpp.  Only tests/, __graft_entry__.smoke() and
bench.py (cpu_baseline / --impl reference) may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

from spatial_model_editor_b200.pixelsim import CSim, _bind
from spatial_model_editor_b200.problem import Problem

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "pixel_oracle.cpp")
    if force or not os.path.exists(LIB_PATH) or \
            os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"],
                              stdout=subprocess.DEVNULL)
    return LIB_PATH


def load():
    global _lib
    if _lib is None:
        build()
        _lib = _bind(C.CDLL(LIB_PATH), "oracle_")
    return _lib


class OraclePixelSim(CSim):
    def __init__(self, problem: Problem, threads: int = 1):
        super().__init__(load(), "oracle_", problem)
        self._f("set_threads")(self._h, threads)

    def setThreads(self, n: int) -> None:
        """1 = serial (reference default enableMultiThreading=false);
        0 = all cores; n = n OpenMP threads"""
        self._f("set_threads")(self._h, n)

    @staticmethod
    def maxThreads() -> int:
        return int(load().oracle_max_threads())

    def setCurrentErrormessage(self, msg: str) -> None:
        raise NotImplementedError
