"""Python mirror of the reference simulator interface over the C-ABI.

``PixelSim`` has the methods of ``sme::simulate::BaseSim`` / ``PixelSim``
(basesim.hpp:15-55, pixelsim.hpp:52-88) with the same names, argument meaning
and error behaviour; every call goes through ``libsmeb200.so`` (ctypes).  There
is no CPU fallback: if the CUDA library cannot be loaded this module raises, and
on a machine without a GPU the simulator carries the reference's
"No CUDA devices were found" error and ``run`` returns 0.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Callable, Optional

import numpy as np

from .problem import CProblem, Problem

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsmeb200.so")

_STOP_CB = C.CFUNCTYPE(C.c_int, C.c_void_p)
_ALLREDUCE_CB = C.CFUNCTYPE(None, C.POINTER(C.c_double), C.c_int, C.c_void_p)


def _bind(lib, prefix: str):
    """declare the argument/return types of the <prefix>_* entry points"""
    vp = C.c_void_p
    dpp = C.POINTER(C.POINTER(C.c_double))
    szp = C.POINTER(C.c_size_t)

    def fn(name, res, args, optional=False):
        try:
            f = getattr(lib, prefix + name)
        except AttributeError:
            if optional:
                return
            raise
        f.restype = res
        f.argtypes = args

    fn("create", C.c_int, [C.POINTER(CProblem), C.POINTER(vp)])
    fn("destroy", None, [vp])
    fn("run", C.c_size_t, [vp, C.c_double, C.c_double, _STOP_CB, vp])
    fn("get_concentrations", C.c_int, [vp, C.c_size_t, dpp, szp])
    fn("get_dcdt", C.c_int, [vp, C.c_size_t, dpp, szp])
    fn("get_lower_order", C.c_int,
       [vp, C.c_size_t, C.c_size_t, C.c_size_t, C.POINTER(C.c_double)])
    fn("set_concentrations", C.c_int, [vp, C.c_size_t, C.POINTER(C.c_double), C.c_size_t])
    fn("error", C.c_char_p, [vp])
    fn("error_images", C.c_int, [vp, C.POINTER(C.POINTER(C.c_uint32)), szp])
    fn("get_stop_requested", C.c_int, [vp])
    fn("set_stop_requested", None, [vp, C.c_int])
    fn("get_next_timestep", C.c_double, [vp])
    fn("set_next_timestep", None, [vp, C.c_double])
    fn("get_discarded_steps", C.c_size_t, [vp])
    fn("get_max_stable_timestep", C.c_double, [vp])
    # product-only entry points
    fn("device_count", C.c_int, [], optional=True)
    fn("compile_only", C.c_int, [C.POINTER(CProblem), C.POINTER(vp)], optional=True)
    fn("get_concentration_padding", C.c_size_t, [vp], optional=True)
    fn("set_error", None, [vp, C.c_char_p], optional=True)
    fn("neighbour_bytes_per_voxel", C.c_size_t, [vp, C.c_size_t], optional=True)
    fn("neighbour_runs", C.c_size_t, [vp, C.c_size_t], optional=True)
    fn("get_species_decimated", C.c_int,
       [vp, C.c_size_t, C.c_size_t, C.POINTER(C.c_uint32),
        C.POINTER(C.POINTER(C.c_double)), C.POINTER(C.c_size_t)], optional=True)
    fn("get_avg_min_max", C.c_int, [vp, C.c_size_t, C.POINTER(C.c_double), C.c_size_t],
       optional=True)
    fn("bench_steps", C.c_double, [vp, C.c_double, C.c_size_t], optional=True)
    fn("launch_count", C.c_uint64, [vp], optional=True)
    fn("last_run_ms", C.c_double, [vp], optional=True)
    fn("set_kernel_timing", None, [vp, C.c_int], optional=True)
    fn("get_kernel_timing", C.c_int,
       [vp, C.POINTER(C.c_double), C.POINTER(C.c_uint64)], optional=True)
    fn("kernel_source", C.c_char_p, [vp], optional=True)
    fn("dominant_kernel", C.c_char_p, [vp], optional=True)
    fn("set_allreduce_max", None, [vp, _ALLREDUCE_CB, vp], optional=True)
    fn("halo_exchanges", C.c_uint64, [vp], optional=True)
    fn("timed_launch_planes", C.c_size_t, [vp], optional=True)
    fn("set_params", C.c_int, [vp, C.POINTER(C.c_double), C.c_size_t], optional=True)
    fn("set_async_download", None, [vp, C.c_int], optional=True)
    fn("set_host_mirror", C.c_int, [vp, C.c_size_t, C.POINTER(C.c_double), C.c_size_t],
       optional=True)
    fn("comm_create", C.c_int, [C.c_char_p, C.c_int, C.c_int, C.POINTER(vp)], optional=True)
    fn("comm_allreduce_max", C.c_int, [vp, C.POINTER(C.c_double), C.c_int], optional=True)
    fn("comm_allreduce_sum", C.c_int, [vp, C.POINTER(C.c_double), C.c_int], optional=True)
    fn("comm_barrier", C.c_int, [vp], optional=True)
    fn("comm_abort", None, [vp], optional=True)
    fn("comm_destroy", None, [vp], optional=True)
    # member-batched ensembles
    u32 = C.c_uint32
    fn("ensemble_create", C.c_int,
       [C.POINTER(CProblem), u32, C.POINTER(C.c_double), C.POINTER(vp)], optional=True)
    fn("ensemble_destroy", None, [vp], optional=True)
    fn("ensemble_run", C.c_size_t, [vp, C.c_double, C.c_double, _STOP_CB, vp], optional=True)
    fn("ensemble_get_concentrations", C.c_int, [vp, u32, C.c_size_t, dpp, szp], optional=True)
    fn("ensemble_set_concentrations", C.c_int,
       [vp, u32, C.c_size_t, C.POINTER(C.c_double), C.c_size_t], optional=True)
    fn("ensemble_get_steps", C.c_size_t, [vp, u32], optional=True)
    fn("ensemble_get_discarded_steps", C.c_size_t, [vp, u32], optional=True)
    fn("ensemble_get_next_timestep", C.c_double, [vp, u32], optional=True)
    fn("ensemble_set_next_timestep", None, [vp, u32, C.c_double], optional=True)
    fn("ensemble_error", C.c_char_p, [vp, C.c_int], optional=True)
    fn("ensemble_set_stop_requested", None, [vp, C.c_int], optional=True)
    fn("ensemble_launch_count", C.c_uint64, [vp], optional=True)
    fn("ensemble_batched_steps", C.c_uint64, [vp], optional=True)
    fn("ensemble_last_run_ms", C.c_double, [vp], optional=True)
    # oracle-only
    fn("set_threads", None, [vp, C.c_int], optional=True)
    fn("max_threads", C.c_int, [], optional=True)
    return lib


_lib = None


def load_library():
    """Load libsmeb200.so; fails loudly if the CUDA extension is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with "
                "`python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        _lib = _bind(C.CDLL(LIB_PATH), "smeb200_")
    return _lib


def device_count() -> int:
    return int(load_library().smeb200_device_count())


class CSim:
    """A simulator object behind a C interface with a given symbol prefix."""

    def __init__(self, lib, prefix: str, problem: Problem, create: str = "create"):
        self._lib = lib
        self._prefix = prefix
        self._problem = problem
        self._h = C.c_void_p()
        cproblem, keep = problem.to_c()
        self._f(create)(C.byref(cproblem), C.byref(self._h))
        del keep
        if not self._h:
            raise RuntimeError("simulator construction failed")
        self._cb_keep = None

    def _f(self, name):
        return getattr(self._lib, self._prefix + name)

    def close(self):
        if getattr(self, "_h", None):
            self._f("destroy")(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- BaseSim interface (basesim.hpp:24-54) ----
    def run(self, time: float, timeout_ms: float = -1.0,
            stopRunningCallback: Optional[Callable[[], bool]] = None) -> int:
        if stopRunningCallback is None:
            cb = _STOP_CB()
        else:
            cb = _STOP_CB(lambda _u: 1 if stopRunningCallback() else 0)
        return int(self._f("run")(self._h, time, timeout_ms, cb, None))

    def _vec(self, getter: str, compartmentIndex: int) -> np.ndarray:
        p = C.POINTER(C.c_double)()
        n = C.c_size_t()
        rc = self._f(getter)(self._h, compartmentIndex, C.byref(p), C.byref(n))
        if rc != 0:
            raise RuntimeError(self.errorMessage())
        comp = self._problem.compartments[compartmentIndex]
        if n.value == 0:
            return np.zeros((0, comp.n_species))
        a = np.ctypeslib.as_array(p, shape=(n.value,))
        return a.reshape(-1, max(comp.n_species, 1))

    def getConcentrations(self, compartmentIndex: int) -> np.ndarray:
        """view [N, S] of the host mirror, valid until the next run()"""
        return self._vec("get_concentrations", compartmentIndex)

    def getConcentrationPadding(self) -> int:
        """number of t,x,y,z components behind the species of every voxel
        (pixelsim.cpp:478)"""
        try:
            return int(self._f("get_concentration_padding")(self._h))
        except AttributeError:  # the oracle has no such entry point
            return self._problem.padding

    def errorMessage(self) -> str:
        return self._f("error")(self._h).decode()

    def errorImages(self) -> np.ndarray:
        p = C.POINTER(C.c_uint32)()
        n = C.c_size_t()
        self._f("error_images")(self._h, C.byref(p), C.byref(n))
        nx, ny, nz = self._problem.image_size
        if n.value == 0:
            return np.zeros((0, ny, nx), dtype=np.uint32)
        return np.ctypeslib.as_array(p, shape=(n.value,)).reshape(nz, ny, nx).copy()

    def setCurrentErrormessage(self, msg: str) -> None:
        self._f("set_error")(self._h, msg.encode())

    def getStopRequested(self) -> bool:
        return bool(self._f("get_stop_requested")(self._h))

    def setStopRequested(self, stop: bool) -> None:
        self._f("set_stop_requested")(self._h, 1 if stop else 0)

    # ---- PixelSim extras reached by dynamic_cast (simulate.cpp:89-115) ----
    def getDcdt(self, compartmentIndex: int) -> np.ndarray:
        return self._vec("get_dcdt", compartmentIndex)

    def getLowerOrderConcentration(self, compartmentIndex: int, speciesIndex: int,
                                   pixelIndex: int) -> float:
        v = C.c_double()
        rc = self._f("get_lower_order")(self._h, compartmentIndex, speciesIndex,
                                        pixelIndex, C.byref(v))
        if rc != 0:
            raise RuntimeError(self.errorMessage())
        return v.value

    def setConcentrations(self, compartmentIndex: int, conc: np.ndarray) -> None:
        a = np.ascontiguousarray(conc, dtype=np.float64)
        rc = self._f("set_concentrations")(
            self._h, compartmentIndex, a.ctypes.data_as(C.POINTER(C.c_double)), a.size)
        if rc != 0:
            raise RuntimeError(self.errorMessage())

    # ---- controller state (pixelsim_base.hpp:40-49) ----
    @property
    def nextTimestep(self) -> float:
        return float(self._f("get_next_timestep")(self._h))

    @nextTimestep.setter
    def nextTimestep(self, dt: float) -> None:
        self._f("set_next_timestep")(self._h, dt)

    @property
    def discardedSteps(self) -> int:
        return int(self._f("get_discarded_steps")(self._h))

    @property
    def maxStableTimestep(self) -> float:
        return float(self._f("get_max_stable_timestep")(self._h))


class PixelSim(CSim):
    """The B200 Pixel simulator (drop-in for PixelSim / CudaPixelSim)."""

    def __init__(self, problem: Problem, compile_only: bool = False):
        super().__init__(load_library(), "smeb200_", problem,
                         "compile_only" if compile_only else "create")

    def benchSteps(self, dt: float, n_steps: int) -> float:
        ms = float(self._f("bench_steps")(self._h, dt, n_steps))
        if ms < 0:
            raise RuntimeError(self.errorMessage())
        return ms

    def lastRunMs(self) -> float:
        return float(self._f("last_run_ms")(self._h))

    def launchCount(self) -> int:
        return int(self._f("launch_count")(self._h))

    def setKernelTiming(self, enable) -> None:
        """False / 0: off; True / 1: every step; n > 1: every n-th step"""
        self._f("set_kernel_timing")(self._h, int(enable))

    def kernelTiming(self):
        ms = C.c_double()
        n = C.c_uint64()
        self._f("get_kernel_timing")(self._h, C.byref(ms), C.byref(n))
        return ms.value, int(n.value)

    def setParams(self, values) -> None:
        """new values of the runtime model parameters (sme_param_k)"""
        a = np.ascontiguousarray(values, dtype=np.float64)
        rc = self._f("set_params")(self._h, a.ctypes.data_as(C.POINTER(C.c_double)), a.size)
        if rc != 0:
            raise RuntimeError(self.errorMessage())

    def getAvgMinMax(self, compartmentIndex: int) -> np.ndarray:
        """[n_primary_species, 3] average, minimum, maximum per species, reduced on
        the device (calculateAvgMinMax, simulate.cpp:272-290)"""
        sp = self._problem.compartments[compartmentIndex].n_primary_species
        out = np.zeros((sp, 3), dtype=np.float64)
        rc = self._f("get_avg_min_max")(self._h, compartmentIndex,
                                        out.ctypes.data_as(C.POINTER(C.c_double)), out.size)
        if rc != 0:
            raise RuntimeError(self.errorMessage())
        return out

    def getSpeciesDecimated(self, compartmentIndex: int, species: int,
                            stride=(1, 1, 1)) -> np.ndarray:
        """one stored species of every (sx, sy, sz)-th voxel, gathered on the device:
        box compartments -> array [mz, mx, my]; voxel lists -> every sx-th voxel"""
        st = (C.c_uint32 * 3)(*[int(v) for v in stride])
        data = C.POINTER(C.c_double)()
        dims = (C.c_size_t * 3)()
        rc = self._f("get_species_decimated")(self._h, compartmentIndex, species, st,
                                              C.byref(data), dims)
        if rc != 0:
            raise RuntimeError(self.errorMessage())
        mx, my, mz = int(dims[0]), int(dims[1]), int(dims[2])
        n = mx * my * mz
        if n == 0:
            return np.zeros((0,), dtype=np.float64)
        a = np.ctypeslib.as_array(data, shape=(n,)).copy()
        return a.reshape(mz, mx, my) if (my > 1 or mz > 1) else a

    def setAsyncDownload(self, enable: bool) -> None:
        """run() ends by starting the D2H of the state on a transfer stream;
        getConcentrations waits for it (cudapixelsim.cpp:1486-1516)"""
        self._f("set_async_download")(self._h, 1 if enable else 0)

    def setHostMirror(self, compartmentIndex: int, buffer: np.ndarray) -> None:
        """``buffer`` (float64, C-contiguous, [N, S]) becomes the host mirror that
        getConcentrations returns; it must outlive the simulator"""
        assert buffer.dtype == np.float64 and buffer.flags.c_contiguous
        rc = self._f("set_host_mirror")(self._h, compartmentIndex,
                                        buffer.ctypes.data_as(C.POINTER(C.c_double)),
                                        buffer.size)
        if rc != 0:
            raise RuntimeError(self.errorMessage())
        self._mirrors = getattr(self, "_mirrors", {})
        self._mirrors[compartmentIndex] = buffer

    def neighbourBytesPerVoxel(self, compartmentIndex: int) -> int:
        """0 (box), 4 (run-compressed neighbour codes) or 24 (plain table)"""
        return int(self._f("neighbour_bytes_per_voxel")(self._h, compartmentIndex))

    def neighbourRuns(self, compartmentIndex: int) -> int:
        return int(self._f("neighbour_runs")(self._h, compartmentIndex))

    def dominantKernel(self) -> str:
        return self._f("dominant_kernel")(self._h).decode()

    def kernelSource(self) -> str:
        return self._f("kernel_source")(self._h).decode()

    def setAllreduceMax(self, fn: Optional[Callable[[np.ndarray], None]]) -> None:
        """fn(values) must max-reduce the float64 array in place over ranks"""
        if fn is None:
            self._cb_keep = _ALLREDUCE_CB()
        else:
            def tramp(p, n, _u):
                fn(np.ctypeslib.as_array(p, shape=(n,)))
            self._cb_keep = _ALLREDUCE_CB(tramp)
        self._f("set_allreduce_max")(self._h, self._cb_keep, None)
