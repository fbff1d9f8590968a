"""ctypes wrapper of baselines/librefcuda_port.so (BASELINE, not product): the
unfused kernel sequence of the reference's CUDA backend (cudapixelsim.cpp:150-349,
1425-1474, 1549-1555) restated for gray-scott on a full box, RK212."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def load():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "librefcuda_port.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} is missing: run `make -C baselines`")
        lib = ctypes.CDLL(path)
        dp = ctypes.POINTER(ctypes.c_double)
        lib.refport_create.restype = ctypes.c_void_p
        lib.refport_create.argtypes = [ctypes.c_int, ctypes.c_uint32, ctypes.c_uint32,
                                       ctypes.c_uint32, dp, dp, ctypes.c_double,
                                       ctypes.c_double, dp]
        lib.refport_steps.restype = ctypes.c_double
        lib.refport_steps.argtypes = [ctypes.c_void_p, ctypes.c_double, ctypes.c_uint64, dp]
        lib.refport_get.argtypes = [ctypes.c_void_p, dp]
        lib.refport_set.argtypes = [ctypes.c_void_p, dp]
        lib.refport_launches.restype = ctypes.c_ulonglong
        lib.refport_launches.argtypes = [ctypes.c_void_p]
        lib.refport_destroy.argtypes = [ctypes.c_void_p]
        _lib = lib
    return _lib


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


class RefCudaPort:
    def __init__(self, nx, ny, nz, conc0, D=(0.2097, 0.105), voxel_size=(2.0, 2.0, 2.0),
                 f=0.037, k=0.06, device=0):
        lib = load()
        self.shape = (nx * ny * nz, 2)
        conc0 = np.ascontiguousarray(conc0, dtype=np.float64).reshape(self.shape)
        D = np.asarray(D, dtype=np.float64)
        h = np.asarray(voxel_size, dtype=np.float64)
        self._h = lib.refport_create(device, nx, ny, nz, _dp(D), _dp(h), f, k, _dp(conc0))
        if not self._h:
            raise RuntimeError("refport_create failed")

    def steps(self, dt, n):
        """n fixed RK212 steps; returns (device ms, (abs err, rel err) of the last)"""
        err = np.zeros(2)
        ms = load().refport_steps(self._h, dt, n, _dp(err))
        if ms < 0:
            raise RuntimeError("refport_steps failed")
        return ms, err

    def get(self):
        out = np.empty(self.shape)
        if load().refport_get(self._h, _dp(out)):
            raise RuntimeError("refport_get failed")
        return out

    def set(self, conc):
        conc = np.ascontiguousarray(conc, dtype=np.float64).reshape(self.shape)
        if load().refport_set(self._h, _dp(conc)):
            raise RuntimeError("refport_set failed")

    def launches(self):
        return int(load().refport_launches(self._h))

    def close(self):
        if self._h:
            load().refport_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
