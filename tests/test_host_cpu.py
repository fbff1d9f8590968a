"""Host-side logic: geometry tables, expression printing, RHS assembly."""
import numpy as np

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
