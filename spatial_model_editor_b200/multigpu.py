"""z-slab decomposition of a box compartment over the GPUs of one node, one
process per GPU (the reference has no multi-GPU path).

The voxel order is z-outermost (geometry.cpp:55-72), so a slab is a contiguous
range of the voxel list.  Rank r owns planes ``slab_range(nz, r, world)`` and
its local array carries one ghost plane towards each neighbour.  The C library
does the rest: boundary planes are copied into the neighbours' ghost planes over
NVLink after every Runge-Kutta stage (CUDA IPC peer memory, stream-ordered flag
handshake) and the two error norms are max-reduced over ranks once per adaptive
step through a node-local shared-memory communicator, so that every rank takes
the same step-size decisions.  There is no data-path collective beyond the halo
planes; ``torch.distributed`` is only used by callers to agree on the
communicator's name and to gather results.
"""
from __future__ import annotations

import copy
import ctypes as C
import os
from typing import Callable, Optional, Tuple

import numpy as np

from .pixelsim import PixelSim, load_library
from .problem import Compartment, Problem


def slab_range(nz: int, rank: int, world: int) -> Tuple[int, int]:
    """owned planes [z0, z1) of ``rank``: as even as possible, low ranks first"""
    base, extra = divmod(nz, world)
    z0 = rank * base + min(rank, extra)
    return z0, z0 + base + (1 if rank < extra else 0)


def ghost_range(nz: int, rank: int, world: int) -> Tuple[int, int, int, int]:
    """(g0, g1, lo, hi): planes [g0, g1) of the local array incl. ghosts and the
    number of ghost planes below / above"""
    z0, z1 = slab_range(nz, rank, world)
    lo = 1 if rank > 0 else 0
    hi = 1 if rank < world - 1 else 0
    return z0 - lo, z1 + hi, lo, hi


def default_comm_name(tag: str = "") -> str:
    """a name all ranks of one torchrun job agree on without communicating"""
    job = os.environ.get("TORCHELASTIC_RUN_ID", "") + os.environ.get("MASTER_PORT", "")
    return f"smeb200_{os.getuid()}_{job}_{tag}"


def local_problem(problem: Problem, rank: int, world: int, comm_name: str,
                  make_slab: Optional[Callable[[int, int], np.ndarray]] = None) -> Problem:
    """the Problem rank ``rank`` hands to the C library.  ``problem`` describes the
    GLOBAL box; its ``conc0`` may be a placeholder when ``make_slab(g0, g1)``
    supplies the local initial state (planes g0..g1, ghosts included) — at 512^3
    per GPU nobody wants to build the global array on every rank."""
    if problem.membranes:
        raise ValueError("z-slab decomposition does not support membranes yet")
    out = copy.copy(problem)
    out.compartments = []
    for c in problem.compartments:
        if c.box is None:
            raise ValueError("z-slab decomposition needs box compartments")
        nx, ny, nz = c.box
        g0, g1, _, _ = ghost_range(nz, rank, world)
        plane = nx * ny
        if make_slab is not None:
            conc = np.ascontiguousarray(make_slab(g0, g1), dtype=np.float64)
            conc = conc.reshape((g1 - g0) * plane, c.n_species)
        else:
            conc = np.ascontiguousarray(c.conc0[g0 * plane:g1 * plane])
        lc = Compartment(id=c.id, species=c.species, expressions=c.expressions,
                         diffusion=c.diffusion, conc0=conc, box=(nx, ny, g1 - g0))
        out.compartments.append(lc)
    out.rank, out.n_ranks, out.comm_name = rank, world, comm_name
    return out


class SlabPixelSim(PixelSim):
    """PixelSim of one rank's slab.  ``getConcentrations`` returns the OWNED
    planes only; ``gather`` assembles the global array on every rank with a
    caller-supplied all-gather (e.g. ``torch.distributed.all_gather_object``)."""

    def __init__(self, problem: Problem, rank: int, world: int,
                 make_slab: Optional[Callable[[int, int], np.ndarray]] = None,
                 comm_name: Optional[str] = None, variant: int = 0):
        self.rank, self.world = rank, world
        self.global_problem = problem
        lp = local_problem(problem, rank, world, comm_name or default_comm_name(),
                           make_slab)
        lp.kernel_variant = variant
        super().__init__(lp)

    def _owned(self, ci: int, a: np.ndarray) -> np.ndarray:
        nx, ny, nz = self.global_problem.compartments[ci].box
        _, _, lo, hi = ghost_range(nz, self.rank, self.world)
        plane = nx * ny
        return a[lo * plane: a.shape[0] - hi * plane]

    def getConcentrationsWithGhosts(self, compartmentIndex: int) -> np.ndarray:
        return super().getConcentrations(compartmentIndex)

    def getConcentrations(self, compartmentIndex: int) -> np.ndarray:
        return self._owned(compartmentIndex, super().getConcentrations(compartmentIndex))

    def getDcdt(self, compartmentIndex: int) -> np.ndarray:
        return self._owned(compartmentIndex, super().getDcdt(compartmentIndex))

    def setConcentrations(self, compartmentIndex: int, conc: np.ndarray) -> None:
        """conc: the local array INCLUDING ghost planes"""
        super().setConcentrations(compartmentIndex, conc)

    def haloExchanges(self) -> int:
        return int(self._lib.smeb200_halo_exchanges(self._h))


class Comm:
    """the node-local communicator on its own (CPU-usable; tests)"""

    def __init__(self, name: str, rank: int, world: int):
        self._lib = load_library()
        self._h = C.c_void_p()
        rc = self._lib.smeb200_comm_create(name.encode(), rank, world, C.byref(self._h))
        if rc != 0 or not self._h:
            raise RuntimeError("smeb200_comm_create failed")

    def allreduce_max(self, values: np.ndarray) -> np.ndarray:
        a = np.ascontiguousarray(values, dtype=np.float64).copy()
        rc = self._lib.smeb200_comm_allreduce_max(
            self._h, a.ctypes.data_as(C.POINTER(C.c_double)), a.size)
        if rc != 0:
            raise RuntimeError("smeb200_comm_allreduce_max failed")
        return a

    def barrier(self) -> None:
        if self._lib.smeb200_comm_barrier(self._h) != 0:
            raise RuntimeError("smeb200_comm_barrier failed")

    def close(self) -> None:
        if self._h:
            self._lib.smeb200_comm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
