"""z-slab decomposition over the GPUs of one node, one process per GPU
(the reference has no multi-GPU path): box compartments
as whole planes, voxel-list compartments (and the membranes between them) cut at
plane boundaries.

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
from typing import Callable, List, Optional, Sequence, Tuple

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


_comm_generation = 0


def default_comm_name(tag: str = "") -> str:
    """a name all ranks of one torchrun job agree on without communicating: the
    job's identity plus a counter of the communicators this process has named (all
    ranks create their decomposed simulators in the same order), so back-to-back
    simulators never meet on one segment"""
    global _comm_generation
    _comm_generation += 1
    job = os.environ.get("TORCHELASTIC_RUN_ID", "") + os.environ.get("MASTER_PORT", "")
    return f"smeb200_{os.getuid()}_{job}_{tag}_{_comm_generation}"


def balanced_cuts(voxel_z: Sequence[np.ndarray], nz: int, world: int) -> List[int]:
    """plane cuts [c_0 = 0, ..., c_world = nz] for voxel-list geometries: the planes
    are dealt so that every rank gets about the same number of VOXELS (all
    compartments together) and at least one plane — an irregular compartment does
    not fill its image evenly along z (the gray-scott-3d blob upscaled to 256^3:
    equal plane counts give the eight ranks 0 / 0 / 145 k / ... / 0 voxels)"""
    per_plane = np.zeros(nz, dtype=np.int64)
    for z in voxel_z:
        per_plane += np.bincount(np.asarray(z, dtype=np.int64), minlength=nz)[:nz]
    cum = np.concatenate([[0], np.cumsum(per_plane)])
    total = int(cum[-1])
    cuts = [0]
    for r in range(1, world):
        target = total * r / world
        c = int(np.searchsorted(cum, target, "left"))
        # the nearer of the two planes around the target
        if c > 0 and abs(cum[c - 1] - target) <= abs(cum[min(c, nz)] - target):
            c -= 1
        c = max(c, cuts[-1] + 1)            # at least one plane for the rank before
        c = min(c, nz - (world - r))        # and for every rank after
        cuts.append(c)
    cuts.append(nz)
    return cuts


def list_slab(voxels: np.ndarray, nz: int, rank: int, world: int,
              cuts: Optional[Sequence[int]] = None) -> Tuple[int, int, int, int]:
    """(g0, o0, o1, g1) for a voxel-list compartment: the local list is the global
    slice [g0, g1), the owned voxels are [o0, o1).  The voxel order is z-major
    (geometry.cpp:55-72), so the ghost voxels (the compartment's voxels in the
    planes just below / above the owned planes) are adjacent to the owned range."""
    z = voxels[:, 2]
    z0, z1 = (cuts[rank], cuts[rank + 1]) if cuts is not None else slab_range(nz, rank, world)
    o0, o1 = int(np.searchsorted(z, z0, "left")), int(np.searchsorted(z, z1, "left"))
    g0 = int(np.searchsorted(z, z0 - 1, "left")) if rank > 0 else o0
    g1 = int(np.searchsorted(z, z1 + 1, "left")) if rank < world - 1 else o1
    return g0, o0, o1, g1


def local_problem(problem: Problem, rank: int, world: int, comm_name: str,
                  make_slab: Optional[Callable[[int, int], np.ndarray]] = None) -> Problem:
    """the Problem rank ``rank`` hands to the C library.  ``problem`` describes the
    GLOBAL geometry.  Box compartments: ``conc0`` may be a placeholder when
    ``make_slab(g0, g1)`` supplies the local initial state (planes g0..g1, ghosts
    included) — at 512^3 per GPU nobody wants to build the global array on every
    rank.  Voxel-list compartments (with ``voxels``) are cut at plane boundaries:
    local list = [ghost voxels of the plane below | owned | ghost voxels of the
    plane above], local neighbour indices, and every membrane pair that has at
    least one owned voxel (the partner may be a ghost: each rank accumulates the
    fluxes of its own voxels only, no remote writes)."""
    out = copy.copy(problem)
    out.compartments = []
    nz_image = problem.image_size[2]
    slabs = []
    # only voxel lists: cut where the voxels are, not where the planes are
    cuts = None
    if all(c.nn is not None and c.voxels is not None for c in problem.compartments) \
            and nz_image >= world:
        cuts = balanced_cuts([c.voxels[:, 2] for c in problem.compartments], nz_image, world)
    for c in problem.compartments:
        if c.box is not None and c.nn is None:
            nx, ny, nz = c.box
            g0, g1, _, _ = ghost_range(nz, rank, world)
            plane = nx * ny
            if make_slab is not None:
                conc = np.ascontiguousarray(make_slab(g0, g1), dtype=np.float64)
                conc = conc.reshape((g1 - g0) * plane, c.n_species)
            else:
                conc = np.ascontiguousarray(c.conc0[g0 * plane:g1 * plane])
            lc = copy.copy(c)
            lc.conc0, lc.box = conc, (nx, ny, g1 - g0)
            if c.diffusion_array is not None:
                lc.diffusion_array = np.ascontiguousarray(
                    c.diffusion_array[g0 * plane:g1 * plane])
            slabs.append(None)
        else:
            if c.voxels is None:
                raise ValueError("z-slab decomposition of a voxel-list compartment needs "
                                 "its voxel coordinates")
            g0, o0, o1, g1 = list_slab(c.voxels, nz_image, rank, world, cuts)
            z = c.voxels[:, 2]
            z0, z1 = (cuts[rank], cuts[rank + 1]) if cuts is not None else \
                slab_range(nz_image, rank, world)
            nn = c.nn[g0:g1].astype(np.int64) - g0
            own = np.arange(g1 - g0, dtype=np.int64)
            ghost = (own < o0 - g0) | (own >= o1 - g0)
            nn[ghost] = own[ghost, None]          # rows of ghost voxels are ignored
            assert nn.min(initial=0) >= 0 and nn.max(initial=0) < max(g1 - g0, 1)
            lc = copy.copy(c)
            lc.conc0 = np.ascontiguousarray(c.conc0[g0:g1])
            lc.nn = nn.astype(np.uint32)
            lc.voxels = np.ascontiguousarray(c.voxels[g0:g1])
            if c.diffusion_array is not None:
                lc.diffusion_array = np.ascontiguousarray(c.diffusion_array[g0:g1])
            lc.box = None
            first = int(np.count_nonzero(z[o0:o1] == z0)) if o1 > o0 else 0
            last = int(np.count_nonzero(z[o0:o1] == z1 - 1)) if o1 > o0 else 0
            lc.slab = (o0 - g0, g1 - o1, first, last)
            slabs.append((g0, o0, o1, g1))
        out.compartments.append(lc)
    out.membranes = []
    for m in problem.membranes:
        sa = slabs[m.comp_a] if m.comp_a >= 0 else None
        sb = slabs[m.comp_b] if m.comp_b >= 0 else None
        lm = copy.copy(m)
        lm.pairs = []
        for d in range(6):
            pr = np.asarray(m.pairs[d], dtype=np.int64).reshape(-1, 2)
            keep = np.zeros(len(pr), dtype=bool)
            if sa is not None:
                keep |= (pr[:, 0] >= sa[1]) & (pr[:, 0] < sa[2])
            if sb is not None:
                keep |= (pr[:, 1] >= sb[1]) & (pr[:, 1] < sb[2])
            pr = pr[keep].copy()
            if sa is not None:
                pr[:, 0] -= sa[0]
                assert np.all((pr[:, 0] >= 0) & (pr[:, 0] < sa[3] - sa[0]))
            if sb is not None:
                pr[:, 1] -= sb[0]
                assert np.all((pr[:, 1] >= 0) & (pr[:, 1] < sb[3] - sb[0]))
            lm.pairs.append(pr.astype(np.uint32))
        out.membranes.append(lm)
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
        c = self._problem.compartments[ci]
        if c.nn is not None:
            lo, hi = c.slab[0], c.slab[1]
            return a[lo: a.shape[0] - hi]
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

    def timedLaunchPlanes(self) -> int:
        """planes of the largest compartment the timed stage launch covers"""
        return int(self._lib.smeb200_timed_launch_planes(self._h))


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

    def allreduce_sum(self, values: np.ndarray) -> np.ndarray:
        """sum over the ranks, added in rank order (the same bits on every rank)"""
        a = np.ascontiguousarray(values, dtype=np.float64).copy()
        rc = self._lib.smeb200_comm_allreduce_sum(
            self._h, a.ctypes.data_as(C.POINTER(C.c_double)), a.size)
        if rc != 0:
            raise RuntimeError("smeb200_comm_allreduce_sum failed")
        return a

    def barrier(self) -> None:
        if self._lib.smeb200_comm_barrier(self._h) != 0:
            raise RuntimeError("smeb200_comm_barrier failed")

    def abort(self) -> None:
        """this rank gives up: every wait of the other ranks raises from now on"""
        self._lib.smeb200_comm_abort(self._h)

    def close(self) -> None:
        if self._h:
            self._lib.smeb200_comm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
