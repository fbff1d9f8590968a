"""Flat problem description: the contract between the "model world" (SME's
``sme::model::Model``) and the kernel world (``include/smeb200.h``).

Everything the reference's two Pixel backends pull out of ``Model`` in their
constructors (pixelsim.cpp:272-406, cudapixelsim.cpp:857-1150) ends up in these
plain arrays and strings.  ``Problem.to_c()`` produces the ``smeb200_problem``
POD that crosses the C-ABI.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

RK101, RK212, RK323, RK435 = 0, 1, 2, 3
INTEGRATOR_NAMES = {"rk101": RK101, "rk212": RK212, "rk323": RK323, "rk435": RK435}
STAGES = {RK101: 1, RK212: 2, RK323: 3, RK435: 5}
DBL_MAX = float(np.finfo(np.float64).max)

HAS_CROSS_DIFFUSION = 1
HAS_TIME_DEPENDENCE = 2
HAS_SPACE_DEPENDENCE = 32
HAS_NON_SPATIAL_SPECIES = 4
HAS_NON_UNIFORM_DIFFUSION = 8
HAS_NON_UNIT_STORAGE = 16

ABI_VERSION = 3


class CCompartment(C.Structure):
    _fields_ = [
        ("id", C.c_char_p),
        ("n_voxels", C.c_uint64),
        ("n_species", C.c_uint32),
        ("box_nx", C.c_int32),
        ("box_ny", C.c_int32),
        ("box_nz", C.c_int32),
        ("nn", C.POINTER(C.c_uint32)),
        ("voxels", C.POINTER(C.c_int32)),
        ("diffusion", C.POINTER(C.c_double)),
        ("conc0", C.POINTER(C.c_double)),
        ("expressions", C.POINTER(C.c_char_p)),
        ("species_names", C.POINTER(C.c_char_p)),
        ("n_primary_species", C.c_uint32),
        ("n_non_spatial", C.c_uint32),
        ("non_spatial", C.POINTER(C.c_uint32)),
        ("inv_storage", C.POINTER(C.c_double)),
        ("diffusion_array", C.POINTER(C.c_double)),
        ("n_cross_diffusion", C.c_uint32),
        ("cross_target", C.POINTER(C.c_uint32)),
        ("cross_source", C.POINTER(C.c_uint32)),
        ("cross_expressions", C.POINTER(C.c_char_p)),
        ("diffusion_raw", C.POINTER(C.c_double)),
        ("slab_ghost_lo", C.c_uint32),
        ("slab_ghost_hi", C.c_uint32),
        ("slab_first_plane", C.c_uint32),
        ("slab_last_plane", C.c_uint32),
    ]


class CMembrane(C.Structure):
    _fields_ = [
        ("id", C.c_char_p),
        ("comp_a", C.c_int32),
        ("comp_b", C.c_int32),
        ("n_pairs", C.c_uint64 * 6),
        ("pairs", C.POINTER(C.c_uint32) * 6),
        ("flux_length", C.c_double * 6),
        ("expressions", C.POINTER(C.c_char_p)),
    ]


class CProblem(C.Structure):
    _fields_ = [
        ("abi_version", C.c_uint32),
        ("n_compartments", C.c_uint32),
        ("compartments", C.POINTER(CCompartment)),
        ("n_membranes", C.c_uint32),
        ("membranes", C.POINTER(CMembrane)),
        ("integrator", C.c_int32),
        ("max_err_abs", C.c_double),
        ("max_err_rel", C.c_double),
        ("max_timestep", C.c_double),
        ("image_nx", C.c_int32),
        ("image_ny", C.c_int32),
        ("image_nz", C.c_int32),
        ("unsupported", C.c_uint32),
        ("device", C.c_int32),
        ("rank", C.c_int32),
        ("n_ranks", C.c_int32),
        ("kernel_variant", C.c_int32),
        ("n_params", C.c_uint32),
        ("comm_name", C.c_char_p),
        ("params", C.POINTER(C.c_double)),
        ("voxel_size", C.c_double * 3),
        ("reserved", C.c_uint32 * 2),
    ]


@dataclass
class Compartment:
    """geometry::Compartment + the SimCompartment inputs."""

    id: str
    species: List[str]                    # simulated species ids
    expressions: List[str]                # C expressions in sme_var_k, one per species
    diffusion: np.ndarray                 # [S, 3]  D/dx^2, D/dy^2, D/dz^2
    conc0: np.ndarray                     # [N, S]
    nn: Optional[np.ndarray] = None       # [N, 6] uint32 or None for a box
    box: Optional[Tuple[int, int, int]] = None  # (nx, ny, nz) when nn is None
    voxels: Optional[np.ndarray] = None   # [N, 3] int32 (x, y, z) image coordinates
    n_primary: Optional[int] = None       # model species; the rest is t,x,y,z padding
    non_spatial: Sequence[int] = ()       # species whose dc/dt is spatially averaged
    inv_storage: Optional[np.ndarray] = None  # [S] 1/storage, None = unit storage
    diffusion_array: Optional[np.ndarray] = None  # [N, S] raw D per voxel (non-uniform)
    cross_diffusion: Sequence[Tuple[int, int, str]] = ()  # (target, source, expression)
    diffusion_raw: Optional[np.ndarray] = None    # [S] raw D (needed with cross-diffusion)
    slab: Tuple[int, int, int, int] = (0, 0, 0, 0)  # ghost_lo, ghost_hi, first/last plane

    @property
    def n_primary_species(self) -> int:
        return self.n_species if self.n_primary is None else int(self.n_primary)

    @property
    def n_voxels(self) -> int:
        return int(self.conc0.shape[0])

    @property
    def n_species(self) -> int:
        return len(self.species)


@dataclass
class Membrane:
    """geometry::Membrane + the SimMembrane inputs."""

    id: str
    comp_a: int
    comp_b: int
    pairs: List[np.ndarray]               # 6 x [n_d, 2] uint32 (ixA, ixB)
    flux_length: Sequence[float]          # 6 values
    expressions: List[str]                # SA + SB expressions


@dataclass
class Problem:
    compartments: List[Compartment]
    membranes: List[Membrane] = field(default_factory=list)
    integrator: int = RK212
    max_err_abs: float = DBL_MAX          # PixelOptions defaults,
    max_err_rel: float = 0.005            # simulate_options.hpp:132-136
    max_timestep: float = DBL_MAX
    image_size: Tuple[int, int, int] = (0, 0, 0)
    unsupported: int = 0
    device: int = -1
    kernel_variant: int = 0
    rank: int = 0                         # z-slab decomposition (multigpu.py)
    n_ranks: int = 1
    comm_name: Optional[str] = None
    params: Sequence[float] = ()          # values of sme_param_k (runtime parameters)
    voxel_size: Tuple[float, float, float] = (1.0, 1.0, 1.0)

    def evals_per_step(self) -> int:
        """voxel*species RHS evaluations in one RK step (the bench metric);
        the t,x,y,z padding components do not count."""
        return STAGES[self.integrator] * sum(
            c.n_voxels * c.n_primary_species for c in self.compartments)

    @property
    def padding(self) -> int:
        """getConcentrationPadding (pixelsim.cpp:478)"""
        return max((c.n_species - c.n_primary_species for c in self.compartments),
                   default=0)

    def to_c(self):
        """Returns (CProblem, keepalive).  keepalive owns every buffer the POD
        points into and must outlive the create call."""
        keep = []

        def strs(items):
            arr = (C.c_char_p * max(len(items), 1))()
            for i, s in enumerate(items):
                arr[i] = s.encode()
            keep.append(arr)
            return C.cast(arr, C.POINTER(C.c_char_p))

        def arr(a, dtype, ctype):
            a = np.ascontiguousarray(a, dtype=dtype)
            keep.append(a)
            return a.ctypes.data_as(C.POINTER(ctype))

        comps = (CCompartment * max(len(self.compartments), 1))()
        for i, c in enumerate(self.compartments):
            cc = comps[i]
            cc.id = c.id.encode()
            cc.n_voxels = c.n_voxels
            cc.n_species = c.n_species
            if c.nn is None:
                assert c.box is not None
                cc.box_nx, cc.box_ny, cc.box_nz = c.box
                cc.nn = None
            else:
                assert c.nn.shape == (c.n_voxels, 6)
                cc.nn = arr(c.nn, np.uint32, C.c_uint32)
            cc.voxels = arr(c.voxels, np.int32, C.c_int32) if c.voxels is not None else None
            d = np.asarray(c.diffusion, dtype=np.float64).reshape(c.n_species, 3)
            cc.diffusion = arr(d, np.float64, C.c_double)
            assert c.conc0.shape == (c.n_voxels, c.n_species)
            cc.conc0 = arr(c.conc0, np.float64, C.c_double)
            assert len(c.expressions) == c.n_species
            cc.expressions = strs(c.expressions)
            cc.species_names = strs(c.species)
            cc.n_primary_species = c.n_primary_species
            cc.n_non_spatial = len(c.non_spatial)
            cc.non_spatial = arr(np.asarray(c.non_spatial, dtype=np.uint32), np.uint32,
                                 C.c_uint32) if len(c.non_spatial) else None
            cc.inv_storage = arr(np.asarray(c.inv_storage, dtype=np.float64).reshape(
                c.n_species), np.float64, C.c_double) if c.inv_storage is not None else None
            if c.diffusion_array is not None:
                assert c.diffusion_array.shape == (c.n_voxels, c.n_species)
                cc.diffusion_array = arr(c.diffusion_array, np.float64, C.c_double)
            (cc.slab_ghost_lo, cc.slab_ghost_hi, cc.slab_first_plane,
             cc.slab_last_plane) = c.slab
            cc.n_cross_diffusion = len(c.cross_diffusion)
            if len(c.cross_diffusion):
                cc.cross_target = arr([t for t, _, _ in c.cross_diffusion], np.uint32, C.c_uint32)
                cc.cross_source = arr([s_ for _, s_, _ in c.cross_diffusion], np.uint32,
                                      C.c_uint32)
                cc.cross_expressions = strs([e for _, _, e in c.cross_diffusion])
            if c.diffusion_raw is not None:
                cc.diffusion_raw = arr(np.asarray(c.diffusion_raw, dtype=np.float64).reshape(
                    c.n_species), np.float64, C.c_double)
        membs = (CMembrane * max(len(self.membranes), 1))()
        for i, m in enumerate(self.membranes):
            cm = membs[i]
            cm.id = m.id.encode()
            cm.comp_a = m.comp_a
            cm.comp_b = m.comp_b
            for d in range(6):
                p = np.ascontiguousarray(m.pairs[d], dtype=np.uint32).reshape(-1, 2)
                cm.n_pairs[d] = p.shape[0]
                cm.pairs[d] = arr(p, np.uint32, C.c_uint32)
                cm.flux_length[d] = float(m.flux_length[d])
            cm.expressions = strs(m.expressions)
        keep += [comps, membs]
        p = CProblem()
        p.abi_version = ABI_VERSION
        p.n_compartments = len(self.compartments)
        p.compartments = C.cast(comps, C.POINTER(CCompartment))
        p.n_membranes = len(self.membranes)
        p.membranes = C.cast(membs, C.POINTER(CMembrane))
        p.integrator = self.integrator
        p.max_err_abs = self.max_err_abs
        p.max_err_rel = self.max_err_rel
        p.max_timestep = self.max_timestep
        p.image_nx, p.image_ny, p.image_nz = self.image_size
        p.unsupported = self.unsupported
        p.device = self.device
        p.rank = self.rank
        p.n_ranks = self.n_ranks
        p.comm_name = self.comm_name.encode() if self.comm_name else None
        p.kernel_variant = self.kernel_variant
        p.voxel_size[0], p.voxel_size[1], p.voxel_size[2] = self.voxel_size
        p.n_params = len(self.params)
        p.params = arr(np.asarray(self.params, dtype=np.float64), np.float64, C.c_double) \
            if len(self.params) else None
        return p, keep
