"""Geometry tables of the Pixel solver, restated from the reference.

Input is a label image stack ``labels[z, y, x]`` in QImage orientation (y = 0 is
the top row).  Output are the arrays the solver consumes: per compartment the
voxel list and the 6-neighbour table, per membrane the per-face index pairs.

Reference: geometry::Compartment ctor (core/model/src/geometry.cpp:40-114),
ImageMembranePixels::setImages (core/model/src/model_membranes_util.cpp:65-129),
geometry::Membrane ctor + getFaceDirection (geometry.cpp:18-36, 147-178).
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np


def compartment_voxels(labels: np.ndarray, label: int) -> np.ndarray:
    """[N, 3] int32 (x, y, z) of the voxels of one compartment in the
    reference's enumeration order: z outermost, then x, then y innermost
    (geometry.cpp:55-72)."""
    nz, ny, nx = labels.shape
    mask = labels == label
    # order (z, x, y): transpose to [z, x, y] and take C-order nonzero
    z, x, y = np.nonzero(np.transpose(mask, (0, 2, 1)))
    return np.stack([x, y, z], axis=1).astype(np.int32)


def voxel_index_map(shape: Tuple[int, int, int], voxels: np.ndarray) -> np.ndarray:
    """dense [nz, ny, nx] int64 map voxel -> index in the voxel list, -1 outside"""
    idx = np.full(shape, -1, dtype=np.int64)
    idx[voxels[:, 2], voxels[:, 1], voxels[:, 0]] = np.arange(len(voxels))
    return idx


def neighbour_table(shape: Tuple[int, int, int], voxels: np.ndarray) -> np.ndarray:
    """[N, 6] uint32 neighbour indices in the order +x, -x, +y, -y, +z, -z
    (QImage coordinates); a neighbour outside the compartment is the voxel's own
    index = zero-flux Neumann boundary (geometry.cpp:98-110)."""
    nz, ny, nx = shape
    n = len(voxels)
    idx = voxel_index_map(shape, voxels)
    own = np.arange(n, dtype=np.int64)
    x, y, z = (voxels[:, 0].astype(np.int64), voxels[:, 1].astype(np.int64),
               voxels[:, 2].astype(np.int64))
    nn = np.empty((n, 6), dtype=np.int64)
    for k, (dx, dy, dz) in enumerate(
            [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]):
        xx, yy, zz = x + dx, y + dy, z + dz
        inside = (xx >= 0) & (xx < nx) & (yy >= 0) & (yy < ny) & (zz >= 0) & (zz < nz)
        cand = np.full(n, -1, dtype=np.int64)
        cand[inside] = idx[zz[inside], yy[inside], xx[inside]]
        nn[:, k] = np.where(cand >= 0, cand, own)
    return nn.astype(np.uint32)


def membrane_voxel_pairs(labels: np.ndarray, colour_index: Dict[int, int]
                         ) -> Dict[Tuple[int, int], List[Tuple[tuple, tuple]]]:
    """Pairs of face-adjacent voxels of different colour, keyed by
    (label of lower colour index, label of higher colour index), in the scan
    order of ImageMembranePixels::setImages (model_membranes_util.cpp:77-128):
    x-neighbours (per z, per y, x ascending), then y-neighbours (per z, per x,
    y ascending), then z-neighbours (per x, per y, z ascending).  Each pair is
    (voxel in the lower-index colour, voxel in the higher-index colour)."""
    nz, ny, nx = labels.shape
    # label -> colour index through a lookup table (labels are small ints)
    lo, hi = (int(labels.min()), int(labels.max())) if labels.size else (0, 0)
    lut = np.full(hi - lo + 1, -1, dtype=np.int64)
    for lab, c in colour_index.items():
        if lo <= int(lab) <= hi:
            lut[int(lab) - lo] = c
    ci = lut[labels.astype(np.int64) - lo] if labels.size else labels
    out: Dict[Tuple[int, int], List[Tuple[tuple, tuple]]] = {}

    def add(prev_vox, cur_vox, prev_c, cur_c, prev_l, cur_l):
        if cur_c < prev_c:
            out.setdefault((cur_l, prev_l), []).append((cur_vox, prev_vox))
        elif cur_c > prev_c:
            out.setdefault((prev_l, cur_l), []).append((prev_vox, cur_vox))

    # x-neighbours
    diff = ci[:, :, 1:] != ci[:, :, :-1]
    for z, y, xm in zip(*np.nonzero(diff)):
        x = xm + 1
        add((x - 1, y, z), (x, y, z), ci[z, y, x - 1], ci[z, y, x],
            int(labels[z, y, x - 1]), int(labels[z, y, x]))
    # y-neighbours: per z, per x, y ascending
    diff = ci[:, 1:, :] != ci[:, :-1, :]
    zz, ym, xx = np.nonzero(diff)
    order = np.lexsort((ym, xx, zz))
    for z, y0, x in zip(zz[order], ym[order], xx[order]):
        y = y0 + 1
        add((x, y - 1, z), (x, y, z), ci[z, y - 1, x], ci[z, y, x],
            int(labels[z, y - 1, x]), int(labels[z, y, x]))
    # z-neighbours: per x, per y, z ascending
    if nz > 1:
        diff = ci[1:, :, :] != ci[:-1, :, :]
        zm, yy, xx = np.nonzero(diff)
        order = np.lexsort((zm, yy, xx))
        for z0, y, x in zip(zm[order], yy[order], xx[order]):
            z = z0 + 1
            add((x, y, z - 1), (x, y, z), ci[z - 1, y, x], ci[z, y, x],
                int(labels[z - 1, y, x]), int(labels[z, y, x]))
    return out


def face_direction(a: tuple, b: tuple) -> int:
    """geometry.cpp:18-36: XP, XM, YP, YM, ZP, ZM = 0..5 from sign of (b - a)"""
    if b[0] > a[0]:
        return 0
    if b[0] < a[0]:
        return 1
    if b[1] > a[1]:
        return 2
    if b[1] < a[1]:
        return 3
    if b[2] > a[2]:
        return 4
    return 5


def membrane_face_pairs(shape, voxels_a: np.ndarray, voxels_b: np.ndarray,
                        voxel_pairs: List[Tuple[tuple, tuple]]) -> List[np.ndarray]:
    """Membrane ctor (geometry.cpp:147-172): voxel pairs -> per-face-direction
    (ixA, ixB) index pairs, keeping the pair order within each direction."""
    ia = voxel_index_map(shape, voxels_a)
    ib = voxel_index_map(shape, voxels_b)
    faces: List[list] = [[] for _ in range(6)]
    for a, b in voxel_pairs:
        d = face_direction(a, b)
        faces[d].append((ia[a[2], a[1], a[0]], ib[b[2], b[1], b[0]]))
    return [np.asarray(f, dtype=np.uint32).reshape(-1, 2) for f in faces]


def flux_lengths(voxel_size: Tuple[float, float, float]) -> List[float]:
    """pixelsim_common.hpp:39-54: voxel width / height / depth by direction"""
    w, h, d = voxel_size
    return [w, w, h, h, d, d]


def upscale_labels(labels: np.ndarray, nx: int, ny: int, nz: int) -> np.ndarray:
    """nearest-neighbour upscaling src = floor(i * n_src / n_dst) used for the
    synthetic large geometries."""
    sz, sy, sx = labels.shape
    iz = (np.arange(nz) * sz) // nz
    iy = (np.arange(ny) * sy) // ny
    ix = (np.arange(nx) * sx) // nx
    return labels[np.ix_(iz, iy, ix)]
