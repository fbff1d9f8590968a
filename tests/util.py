import numpy as np

# fixed-step parity bar of BASELINE.json: <= 1e-12 relative in fp64 (stricter
# than the reference's own GPU-vs-CPU bar of rel 1e-11 / abs 1e-13,
# cudapixelsim_t.cpp:46-47)
REL_TOL = 1e-12
ABS_TOL = 1e-14


def max_rel_diff(a, b, abs_floor=ABS_TOL):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.size == 0:
        return 0.0
    d = np.abs(a - b)
    d = np.where(d <= abs_floor, 0.0, d)
    return float(np.max(d / np.maximum(np.abs(b), 1e-300)))


def assert_parity(a, b, rel=REL_TOL, what=""):
    r = max_rel_diff(a, b)
    assert r <= rel, f"{what}: max relative difference {r:.3e} > {rel:.1e}"
    return r


# ---- golden fixtures (tests/golden/*.npz, written by tests/golden/make_golden.py) ----
import json
import os

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name, integrator=None):
    """fixture -> (Problem, meta, labels).  Geometry tables (voxel lists,
    neighbour tables, membrane face pairs) are rebuilt from the label image with
    spatial_model_editor_b200.geometry, so the reference's counts pin them."""
    from spatial_model_editor_b200 import geometry as G
    from spatial_model_editor_b200 import problem as P
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    labels = z["labels"].astype(np.int64)
    nz, ny, nx = labels.shape
    voxel_size = tuple(meta["voxel_size"])
    dx, dy, dz = voxel_size
    comps = []
    for i, c in enumerate(meta["compartments"]):
        vox = G.compartment_voxels(labels, meta["label_of"][c["id"]])
        conc = z[f"conc0_{i}"]
        if conc.shape[0] == 1 and len(vox) != 1:
            conc = np.repeat(conc, len(vox), axis=0)
        d = np.array([[v / (dx * dx), v / (dy * dy), v / (dz * dz)]
                      for v in c["diffusion_constants"]])
        comps.append(P.Compartment(
            id=c["id"], species=c["species"], expressions=c["expressions"], diffusion=d,
            conc0=np.ascontiguousarray(conc, dtype=np.float64),
            nn=G.neighbour_table(labels.shape, vox), voxels=vox,
            n_primary=c.get("n_primary"), non_spatial=c.get("non_spatial", ()),
            inv_storage=None if c.get("inv_storage") is None
            else np.array(c["inv_storage"], dtype=np.float64)))
    membs = []
    if meta["membranes"]:
        order = {}
        for mb in meta["membranes"]:
            for cid in (mb["a"], mb["b"]):
                order.setdefault(meta["label_of"][cid], len(order))
        for lab in meta["label_of"].values():
            order.setdefault(lab, len(order))
        pairs = G.membrane_voxel_pairs(labels, order)
        for mb in meta["membranes"]:
            la, lb = meta["label_of"][mb["a"]], meta["label_of"][mb["b"]]
            key = (la, lb) if order[la] < order[lb] else (lb, la)
            vp = pairs.get(key, [])
            if key != (la, lb):
                vp = [(q, p) for p, q in vp]
            fp = G.membrane_face_pairs(labels.shape, G.compartment_voxels(labels, la),
                                       G.compartment_voxels(labels, lb), vp)
            assert [len(p) for p in fp] == mb["n_pairs"]
            membs.append(P.Membrane(id=mb["id"], comp_a=mb["comp_a"], comp_b=mb["comp_b"],
                                    pairs=fp, flux_length=G.flux_lengths(voxel_size),
                                    expressions=mb["expressions"]))
    problem = P.Problem(
        comps, membs, integrator=meta["integrator"] if integrator is None else integrator,
        max_err_abs=meta["max_err_abs"], max_err_rel=meta["max_err_rel"],
        max_timestep=meta["max_timestep"], image_size=(nx, ny, nz),
        unsupported=meta["unsupported"], params=list(meta.get("params", [])))
    return problem, meta, labels
