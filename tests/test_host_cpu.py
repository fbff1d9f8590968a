"""Host-side logic: geometry tables, expression printing, RHS assembly."""
import numpy as np
import pytest

import spatial_model_editor_b200 as smeb

from spatial_model_editor_b200 import expr as E
from spatial_model_editor_b200 import geometry as G
from spatial_model_editor_b200 import models as M
from spatial_model_editor_b200.pde import ReactionDef, build_rhs


def test_voxel_order_is_z_x_y():
    lab = np.ones((2, 3, 4), dtype=np.int32)  # nz=2, ny=3, nx=4
    v = G.compartment_voxels(lab, 1)
    assert v[0].tolist() == [0, 0, 0] and v[1].tolist() == [0, 1, 0]
    assert v[3].tolist() == [1, 0, 0] and v[12].tolist() == [0, 0, 1]


def test_neighbour_table_box_matches_implicit_rule():
    nx, ny, nz = 4, 3, 2
    lab = np.ones((nz, ny, nx), dtype=np.int32)
    v = G.compartment_voxels(lab, 1)
    nn = G.neighbour_table(lab.shape, v)
    for i, (x, y, z) in enumerate(v):
        exp = [i + ny if x + 1 < nx else i, i - ny if x > 0 else i,
               i + 1 if y + 1 < ny else i, i - 1 if y > 0 else i,
               i + nx * ny if z + 1 < nz else i, i - nx * ny if z > 0 else i]
        assert nn[i].tolist() == exp


def test_neumann_boundary_inside_image():
    lab = np.zeros((1, 5, 5), dtype=np.int32)
    lab[0, 1:4, 1:4] = 1
    v = G.compartment_voxels(lab, 1)
    nn = G.neighbour_table(lab.shape, v)
    # corner voxel (1,1): -x and -y neighbours are itself; z neighbours itself
    assert v[0].tolist() == [1, 1, 0]
    assert nn[0, 1] == 0 and nn[0, 3] == 0 and nn[0, 4] == 0 and nn[0, 5] == 0


def test_membrane_pairs_and_directions():
    lab = np.ones((1, 4, 4), dtype=np.int32)
    lab[0, 1:3, 1:3] = 2
    pairs = G.membrane_voxel_pairs(lab, {1: 0, 2: 1})
    vp = pairs[(1, 2)]
    assert len(vp) == 8
    va, vb = G.compartment_voxels(lab, 1), G.compartment_voxels(lab, 2)
    faces = G.membrane_face_pairs(lab.shape, va, vb, vp)
    assert [len(f) for f in faces] == [2, 2, 2, 2, 0, 0]
    # direction = sign of (B - A): XP pairs have the inner voxel at larger x
    for ia, ib in faces[0]:
        assert vb[ib][0] == va[ia][0] + 1 and vb[ib][1] == va[ia][1]


def test_expression_printing():
    c = E.compile_expressions(["k*A*B^2 - exp(-A)/2"], ["A", "B"], {"k": 0.5})
    assert c == ["(((0.5 * sme_var_0) * (sme_var_1 * sme_var_1)) - (exp((-sme_var_0)) / 2.0))"]
    v = E.evaluate(E.parse("2^3^2 + -1"), {})
    assert v == 511.0


def test_rhs_assembly_drops_zero_stoichiometry():
    rhs = build_rhs(["A", "B", "C"],
                    [ReactionDef("r1", "A*B*k1", {"A": -1, "B": -1, "C": 1}, {"k1": 0.1})])
    assert rhs[0] == rhs[1] == "(-((sme_var_0 * sme_var_1) * 0.1))"
    assert rhs[2] == "((sme_var_0 * sme_var_1) * 0.1)"
    assert build_rhs(["A", "Z"], [ReactionDef("r", "A", {"A": -1})])[1] == "0.0"


def test_two_compartment_problem_tables():
    p = M.two_compartment_brusselator(16, 12, 1)
    a, b = p.compartments
    assert a.n_voxels + b.n_voxels == 16 * 12
    m = p.membranes[0]
    n_pairs = sum(len(x) for x in m.pairs)
    assert n_pairs == 2 * (8 + 6)
    for d in range(6):
        for ia, ib in m.pairs[d]:
            assert ia < a.n_voxels and ib < b.n_voxels


# ---- run-compressed neighbour codes (device layout of voxel-list compartments) ----
def _compress(nn):
    import ctypes as C
    lib = smeb.load_library()
    lib.smeb200_compress_neighbours.restype = C.c_int
    lib.smeb200_compress_neighbours.argtypes = [
        C.POINTER(C.c_uint32), C.c_uint64, C.POINTER(C.c_uint32), C.POINTER(C.c_int32),
        C.POINTER(C.c_uint64)]
    nn = np.ascontiguousarray(nn, dtype=np.uint32)
    n = nn.shape[0]
    codes = np.zeros(n, dtype=np.uint32)
    runs = np.zeros((max(n, 1), 4), dtype=np.int32)
    n_runs = C.c_uint64(runs.shape[0])
    rc = lib.smeb200_compress_neighbours(
        nn.ctypes.data_as(C.POINTER(C.c_uint32)), n,
        codes.ctypes.data_as(C.POINTER(C.c_uint32)),
        runs.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(n_runs))
    return rc, codes, runs[:n_runs.value]


def _decode(codes, runs):
    """what smeb_neighbours (device_kernels.cuh) computes"""
    i = np.arange(len(codes), dtype=np.int64)
    run = (codes & ((1 << 25) - 1)).astype(np.int64)
    fl = (codes >> 25) & 63
    d = runs[run].astype(np.int64)
    nb = np.stack([np.where(fl & 1, i + d[:, 0], i), np.where(fl & 2, i + d[:, 1], i),
                   np.where(fl & 4, i + 1, i), np.where(fl & 8, i - 1, i),
                   np.where(fl & 16, i + d[:, 2], i), np.where(fl & 32, i + d[:, 3], i)], axis=1)
    return nb.astype(np.uint32), (codes >> 31).astype(bool)


@pytest.mark.parametrize("case", ["disc", "cell3d", "blob", "speckle"])
def test_neighbour_codes_decode_to_the_reference_table(case):
    from spatial_model_editor_b200 import geometry as G
    rng = np.random.default_rng(3)
    if case == "disc":
        lab = M.disc_labels(100, 100, inner_fraction=0.5)
    elif case == "cell3d":
        lab = np.ones((9, 14, 11), dtype=np.int32)
        lab[2:7, 3:10, 4:9] = 2
    elif case == "blob":
        z, y, x = np.meshgrid(np.arange(20), np.arange(24), np.arange(28), indexing="ij")
        lab = (((x - 13) ** 2 + (y - 11) ** 2 + 2 * (z - 9) ** 2 < 90)
               | ((x - 20) ** 2 + (y - 6) ** 2 + (z - 14) ** 2 < 30)).astype(np.int32)
    else:  # random speckle: many short runs, several runs per column
        lab = (rng.uniform(size=(4, 30, 30)) < 0.6).astype(np.int32)
    for label in [v for v in np.unique(lab) if v != 0]:
        vox = G.compartment_voxels(lab, int(label))
        nn = G.neighbour_table(lab.shape, vox)
        flagged = nn.copy()
        memb = rng.uniform(size=len(vox)) < 0.1
        flagged[memb, 0] |= np.uint32(0x80000000)
        rc, codes, runs = _compress(flagged)
        assert rc == 0
        dec, mflag = _decode(codes, runs)
        assert np.array_equal(dec, nn)
        assert np.array_equal(mflag, memb)
        if case == "disc":  # one run per column, give or take the rim
            assert len(runs) < 4 * 100


def test_neighbour_codes_refuse_foreign_orderings():
    """a table whose +y neighbour is not i + 1 (not the reference's voxel order) keeps
    the plain layout"""
    nn = np.array([[1, 0, 2, 0, 0, 0], [1, 0, 1, 1, 1, 1], [2, 2, 2, 0, 2, 2]], dtype=np.uint32)
    rc, _, _ = _compress(nn)
    assert rc == 1


def test_symengine_style_expression_text_compiles():
    """What Symbolic::cudaCode prints (SymEngine's CUDA code printer: integer powers as
    pow(x, 2), plain decimal literals, nested parentheses, exp / sqrt / fabs calls)
    goes through smeb200_create's NVRTC path; integer-literal powers become
    multiplications (the CPU solver's LLVM visitor uses llvm.powi)"""
    import spatial_model_editor_b200 as smeb
    from spatial_model_editor_b200 import models as M
    p = M.box_problem(M.GRAY_SCOTT, 8, 8, 4)
    c = p.compartments[0]
    c.expressions = [
        "0.0 + -1.0*sme_var_0*pow(sme_var_1, 2) + 0.037000000000000005*(1.0 - sme_var_0)",
        "0.0 + sme_var_0*pow(sme_var_1, 2) - 0.097*sme_var_1 + 0.0*exp(-pow(sme_var_0, 3))"
        " + 0.0*sqrt(fabs(sme_var_1)) + 0.0*pow(sme_var_0, 0.5)",
    ]
    sim = smeb.PixelSim(p, compile_only=True)
    assert sim.errorMessage() == ""
    src = sim.kernelSource()
    assert "smeb_pow" in src and "pow(sme_var_1, 2)" in src
