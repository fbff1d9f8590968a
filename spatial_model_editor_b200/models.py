"""Programmatic model builders: the kinetics of the reference's example models
on synthetic geometries, as flat ``Problem`` objects.

Kinetic constants are those of the shipped SBML files
(core/resources/models/gray-scott.xml, gray-scott-3d.xml,
brusselator-model.xml, ABtoC.xml).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import geometry as G
from .pde import ReactionDef, build_rhs
from .problem import DBL_MAX, RK212, Compartment, Membrane, Problem


def dimensionless_diffusion(D: Sequence[float], voxel_size=(1.0, 1.0, 1.0)) -> np.ndarray:
    """[S, 3] D/dx^2, D/dy^2, D/dz^2 (pixelsim_impl.cpp:210-212,236-237)"""
    w, h, d = voxel_size
    return np.array([[v / (w * w), v / (h * h), v / (d * d)] for v in D], dtype=np.float64)


GRAY_SCOTT = dict(
    species=["U", "V"],
    D=[0.2097, 0.105],
    constants={"f": 0.037, "k": 0.06},
    reactions=[
        ReactionDef("r_uv", "U * V^2", {"U": -1.0, "V": 1.0}),
        ReactionDef("r_feed", "f * (1 - U)", {"U": 1.0}),
        ReactionDef("r_kill", "(f + k) * V", {"V": -1.0}),
    ],
)

BRUSSELATOR = dict(
    species=["X", "Y"],
    D=[0.2, 0.5],
    constants={"A": 0.5, "B": 3.0, "k1": 1.0, "k2": 2.0, "k3": 1.0, "k4": 1.0},
    reactions=[
        ReactionDef("R1", "k1 * A", {"X": 1.0}),
        ReactionDef("R2", "k2 * X^2 * Y", {"X": 1.0, "Y": -1.0}),
        ReactionDef("R3", "k3 * B * X", {"X": -1.0, "Y": 1.0}),
        ReactionDef("R4", "k4 * X", {"X": -1.0}),
    ],
)

ABTOC = dict(
    species=["A", "B", "C"],
    D=[0.4, 0.4, 25.0],
    constants={},
    reactions=[ReactionDef("r1", "A * B * k1", {"A": -1.0, "B": -1.0, "C": 1.0},
                           {"k1": 0.1})],
)


def kinetics_expressions(kin: dict, substitutions: Optional[Dict[str, float]] = None
                         ) -> List[str]:
    return build_rhs(kin["species"], kin["reactions"], kin["constants"],
                     substitutions=substitutions)


def gray_scott_initial(nx: int, ny: int, nz: int, seed: int = 0) -> np.ndarray:
    """[N, 2] initial state in voxel-list order (z, x, y): U = 1, V = a field of
    Gaussian spots of period 100 voxels like the fixture's two-Gaussian initial
    condition, tiled over the larger domain."""
    z, x, y = np.meshgrid(np.arange(nz), np.arange(nx), np.arange(ny), indexing="ij")
    px, py, pz = (x % 100).astype(np.float64), (y % 100).astype(np.float64), \
        (z % 100).astype(np.float64)
    r1 = (px - 33.0) ** 2 + (py - 40.0) ** 2 + ((pz - 25.0) ** 2 if nz > 1 else 0.0)
    r2 = (px - 70.0) ** 2 + (py - 66.0) ** 2 + ((pz - 60.0) ** 2 if nz > 1 else 0.0)
    v = np.exp(-r1 / 50.0) + np.exp(-r2 / 50.0)
    u = np.ones_like(v)
    return np.stack([u.reshape(-1), v.reshape(-1)], axis=1)


def box_problem(kin: dict, nx: int, ny: int, nz: int, conc0: Optional[np.ndarray] = None,
                voxel_size=(1.0, 1.0, 1.0), integrator: int = RK212,
                max_err_rel: float = 0.005, max_err_abs: float = DBL_MAX,
                max_timestep: float = DBL_MAX, seed: int = 1000,
                as_voxel_list: bool = False) -> Problem:
    """single compartment filling an nx*ny*nz image"""
    n, S = nx * ny * nz, len(kin["species"])
    if conc0 is None:
        conc0 = random_state(n, S, seed)
    comp = Compartment(
        id="comp", species=list(kin["species"]), expressions=kinetics_expressions(kin),
        diffusion=dimensionless_diffusion(kin["D"], voxel_size),
        conc0=np.ascontiguousarray(conc0, dtype=np.float64).reshape(n, S),
        box=(nx, ny, nz))
    if as_voxel_list:
        labels = np.ones((nz, ny, nx), dtype=np.int32)
        comp.voxels = G.compartment_voxels(labels, 1)
        comp.nn = G.neighbour_table(labels.shape, comp.voxels)
        comp.box = None
    return Problem([comp], [], integrator=integrator, max_err_rel=max_err_rel,
                   max_err_abs=max_err_abs, max_timestep=max_timestep,
                   image_size=(nx, ny, nz))


def random_state(n: int, S: int, seed: int) -> np.ndarray:
    """U(0.05, 1.25) random state, the reference's GPU-vs-CPU parity protocol
    (cudapixelsim_t.cpp:121-132; distribution bounds only — the stream is
    numpy's, not std::mt19937's)"""
    rng = np.random.default_rng(seed)
    return rng.uniform(0.05, 1.25, size=(n, S))


def labelled_problem(labels: np.ndarray, comp_defs: List[dict],
                     membrane_defs: Sequence[dict] = (), voxel_size=(1.0, 1.0, 1.0),
                     integrator: int = RK212, max_err_rel: float = 0.005,
                     max_err_abs: float = DBL_MAX, max_timestep: float = DBL_MAX,
                     seed: int = 1000, vol_over_l3: float = 1.0) -> Problem:
    """General geometry: ``labels[z, y, x]`` image, one compartment per entry of
    comp_defs ({"label", "id", "kin", optional "conc0"}); the colour index of a
    compartment is its position in comp_defs.  membrane_defs entries:
    {"a": idx, "b": idx, "species_a", "species_b", "reactions": [ReactionDef],
    "constants"} with a < b (A = lower colour index, model_membranes_util.cpp)."""
    nz, ny, nx = labels.shape
    comps: List[Compartment] = []
    colour_index = {d["label"]: i for i, d in enumerate(comp_defs)}
    for i, d in enumerate(comp_defs):
        kin = d["kin"]
        vox = G.compartment_voxels(labels, d["label"])
        n, S = len(vox), len(kin["species"])
        conc0 = d.get("conc0")
        if conc0 is None:
            conc0 = random_state(n, S, seed + i)
        comps.append(Compartment(
            id=d.get("id", f"comp{i}"), species=list(kin["species"]),
            expressions=kinetics_expressions(kin),
            diffusion=dimensionless_diffusion(kin["D"], voxel_size),
            conc0=np.ascontiguousarray(conc0, dtype=np.float64).reshape(n, S),
            nn=G.neighbour_table(labels.shape, vox), voxels=vox))
    membs: List[Membrane] = []
    if membrane_defs:
        pairs = G.membrane_voxel_pairs(labels, colour_index)
        for k, md in enumerate(membrane_defs):
            a, b = md["a"], md["b"]
            assert a < b
            key = (comp_defs[a]["label"], comp_defs[b]["label"])
            vp = pairs.get(key, [])
            face_pairs = G.membrane_face_pairs(labels.shape, comps[a].voxels,
                                               comps[b].voxels, vp)
            # variables: [species of A | species of B] (pixelsim_impl.cpp:960-973)
            names_a = md.get("species_a", [f"{s}_a" for s in comps[a].species])
            names_b = md.get("species_b", [f"{s}_b" for s in comps[b].species])
            exprs = build_rhs(list(names_a) + list(names_b), md["reactions"],
                              md.get("constants", {}), scale=vol_over_l3)
            membs.append(Membrane(id=md.get("id", f"membrane{k}"), comp_a=a, comp_b=b,
                                  pairs=face_pairs,
                                  flux_length=G.flux_lengths(voxel_size),
                                  expressions=exprs))
    return Problem(comps, membs, integrator=integrator, max_err_rel=max_err_rel,
                   max_err_abs=max_err_abs, max_timestep=max_timestep,
                   image_size=(nx, ny, nz))


def disc_labels(nx: int, ny: int, nz: int = 1, inner_fraction: float = 0.0) -> np.ndarray:
    """label image: 0 outside, 1 = disc (cylinder) of diameter ~0.63*n like the
    3149-voxel disc of the 100x100 example, optional 2 = concentric inner disc"""
    y, x = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    cx, cy = (nx - 1) / 2.0, (ny - 1) / 2.0
    r2 = ((x - cx) / nx) ** 2 + ((y - cy) / ny) ** 2
    lab = np.zeros((ny, nx), dtype=np.int32)
    lab[r2 <= 0.3166 ** 2] = 1
    if inner_fraction > 0:
        lab[r2 <= (0.3166 * inner_fraction) ** 2] = 2
    return np.repeat(lab[None], nz, axis=0)


def two_compartment_brusselator(nx: int, ny: int, nz: int = 1, km: float = 0.1,
                                **kw) -> Problem:
    """BASELINE config 3 (synthetic: the shipped brusselator is single
    compartment): outer frame / inner square, brusselator kinetics in both,
    coupled by the linear membrane flux km*(X_a - X_b)."""
    lab = np.ones((nz, ny, nx), dtype=np.int32)
    lab[:, ny // 4: ny - ny // 4, nx // 4: nx - nx // 4] = 2
    mem = dict(a=0, b=1, species_a=["X_a", "Y_a"], species_b=["X_b", "Y_b"],
               constants={"km": km},
               reactions=[ReactionDef("flux", "km * (X_a - X_b)",
                                      {"X_a": -1.0, "X_b": 1.0})])
    return labelled_problem(
        lab, [dict(label=1, id="outer", kin=BRUSSELATOR),
              dict(label=2, id="inner", kin=BRUSSELATOR)], [mem], **kw)
