"""Generates the golden fixtures under tests/golden/ from the reference's own
SBML example models (run HERE, where /root/reference exists; the fixtures are
what travels to the GPU box).

    python tests/golden/make_golden.py

Each fixture ``<name>.npz`` holds the inputs the reference's Pixel backends
read from ``sme::model::Model`` for that model — label image, compartment ids
and labels, simulated species, diffusion constants, initial concentrations,
reaction expressions (C strings in sme_var_k), membrane ids/orientation and
expressions, voxel size, pixel options — extracted by
``spatial_model_editor_b200.sbml``.  The known answers the tests compare
against are the reference's own (simulate_t.cpp; cited in
tests/test_golden_cpu.py), not outputs of this repository.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from spatial_model_editor_b200 import sbml  # noqa: E402

REF = "/root/reference"
MODELS = {
    "very-simple-model": "core/resources/models/very-simple-model.xml",
    "ABtoC": "core/resources/models/ABtoC.xml",
    "gray-scott": "core/resources/models/gray-scott.xml",
    "gray-scott-3d": "core/resources/models/gray-scott-3d.xml",
    "brusselator-model": "core/resources/models/brusselator-model.xml",
    "liver-simplified": "core/resources/models/liver-simplified.xml",
    "single-compartment-diffusion": "core/resources/models/single-compartment-diffusion.xml",
    "single-compartment-diffusion-3d":
        "core/resources/models/single-compartment-diffusion-3d.xml",
    # 30-voxel disc of simulate_t.cpp:1158-1303 (steady states: uniform, zero-storage,
    # spatially varying diffusion)
    "small-single-compartment-diffusion":
        "test/resources/models/small-single-compartment-diffusion.xml",
    # the rest of the reference's example models (cudapixelsim_t.cpp:542-575 sweeps
    # all of them; NFkappaB.xml, a non-spatial L2 file without geometry, is not one)
    "CalciumWavePropagation3D": "core/resources/models/CalciumWavePropagation3D.xml",
    "FitzhughNagumo3D": "core/resources/models/FitzhughNagumo3D.xml",
    "SelKov3D": "core/resources/models/SelKov3D.xml",
    "circadian-clock": "core/resources/models/circadian-clock.xml",
    "liver-cells": "core/resources/models/liver-cells.xml",
    "very-simple-model-3d": "core/resources/models/very-simple-model-3d.xml",
}


def event_stages(m, **kw):
    """Parameter events in time order with the expressions the re-created simulator
    uses from each event time on (Simulation::applyNextEvent -> initSimulator with
    eventSubstitutions, simulate.cpp:227-270; events at one time are applied
    together, :196-208)."""
    out, subs = [], {}
    params = set(m.parameters) | {v for _, _, v, _ in m.events}
    for t in sorted({t for _, t, _, _ in m.events}):
        ids = []
        for eid, et, var, val in m.events:
            if et == t and var in params:
                subs[var] = val
                ids.append(eid)
        if not ids:
            continue
        p = sbml.to_problem(m, substitutions=dict(subs), **kw)
        out.append(dict(time=t, ids=ids, substitutions=dict(subs),
                        expressions=[c.expressions for c in p.compartments],
                        membrane_expressions=[mb.expressions for mb in p.membranes]))
    return out


def dump(name, m, problem):
    if int(m.labels.max()) > 255:  # raw ARGB sample values: store small labels
        values = sorted(set(m.labels.ravel().tolist()))
        remap = {v: i for i, v in enumerate(values)}
        m.labels = np.vectorize(remap.get)(m.labels).astype(np.int64)
        m.label_of = {k: remap[v] for k, v in m.label_of.items()}
    meta = dict(
        name=name, source=os.path.relpath(m.path, REF), voxel_size=list(m.voxel_size),
        origin=list(m.origin), vol_over_l3=m.vol_over_l3, label_of=m.label_of,
        integrator=problem.integrator, max_err_abs=problem.max_err_abs,
        max_err_rel=problem.max_err_rel, max_timestep=problem.max_timestep,
        unsupported=problem.unsupported, params=list(problem.params),
        events=event_stages(m) if not problem.params else [],
        compartments=[dict(id=c.id, species=c.species, expressions=c.expressions,
                           diffusion_constants=[s.diffusion for s in m.species
                                                if s.id in c.species]
                           + [0.0] * (c.n_species - c.n_primary_species),
                           n_primary=c.n_primary_species,
                           non_spatial=[int(k) for k in c.non_spatial],
                           inv_storage=None if c.inv_storage is None
                           else [float(v) for v in c.inv_storage])
                      for c in problem.compartments],
        membranes=[dict(id=mb.id, comp_a=mb.comp_a, comp_b=mb.comp_b,
                        a=[a for i, a, b in m.membranes if i == mb.id][0],
                        b=[b for i, a, b in m.membranes if i == mb.id][0],
                        expressions=mb.expressions,
                        n_pairs=[int(len(p)) for p in mb.pairs])
                   for mb in problem.membranes])
    arrays = {"labels": m.labels.astype(np.uint8)}
    for i, c in enumerate(problem.compartments):
        conc = c.conc0
        # uniform initial states are stored as one row
        if np.all(conc == conc[0]):
            conc = conc[:1]
        arrays[f"conc0_{i}"] = conc
    np.savez_compressed(os.path.join(HERE, name + ".npz"),
                        meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8), **arrays)
    print(f"{name}: {[c.conc0.shape for c in problem.compartments]}, "
          f"{os.path.getsize(os.path.join(HERE, name + '.npz'))} bytes")


def main():
    for name, rel in MODELS.items():
        m = sbml.load_sbml(os.path.join(REF, rel))
        dump(name, m, sbml.to_problem(m))
    # BASELINE config 5: liver-simplified with the scanned rate constant (IkBa mRNA degradation;
    # TNFa is zero in the stored state, so its degradation rate has no effect) kept as
    # a runtime parameter (sme_param_0), so that all ensemble members share one
    # compiled module
    m = sbml.load_sbml(os.path.join(REF, MODELS["liver-simplified"]))
    dump("liver-simplified-scan", m,
         sbml.to_problem(m, runtime_params=("IkBa_mRNA_degradation_c_k1",)))
    # K1: the 1 x 3 voxel membrane chain of simulate_t.cpp:18-88 — the non-spatial
    # model with a 3-pixel image imported and four reactions moved to membranes
    lab = np.array([[[1], [2], [3]]])
    m = sbml.load_sbml(
        os.path.join(REF, "test/resources/models/very-simple-model-non-spatial.xml"),
        geometry=dict(labels=lab, label_of={"c1": 1, "c2": 2, "c3": 3}),
        reaction_locations={"A_uptake": "c1_c2_membrane", "A_transport": "c2_c3_membrane",
                            "B_transport": "c2_c3_membrane", "B_excretion": "c1_c2_membrane"})
    dump("single-pixel-chain", m, sbml.to_problem(m, integrator=0))
    # reactions that depend on t, x, y (simulate_t.cpp:2009-2207): dA/dt = x,
    # dB/dt = y, dC/dt = time -> state padded with time, x, y, z
    m = sbml.load_sbml(os.path.join(REF, "test/resources/models/txy.xml"))
    dump("txy", m, sbml.to_problem(m, integrator=0))
    # the same with non-unit storage (simulate_t.cpp:2054-2087 sets C to 2)
    for sp in m.species:
        sp.storage = {"A": 0.5, "C": 2.0}.get(sp.id, 1.0)
    dump("txy-storage", m, sbml.to_problem(m, integrator=0))
    # non-unit storage on the 3-d diffusion example (simulate_t.cpp:1153-1155)
    m = sbml.load_sbml(os.path.join(REF, MODELS["single-compartment-diffusion-3d"]))
    for sp in m.species:
        sp.storage = {"slow": 2.0, "fast": 3.0}[sp.id]
    dump("single-compartment-diffusion-3d-storage", m, sbml.to_problem(m))
    # non-spatial species (SimCompartment::spatiallyAverageDcdt): very-simple-model
    # with A_c2 and B_c3 made non-spatial (Species::setIsSpatial(id, false))
    m = sbml.load_sbml(os.path.join(REF, MODELS["very-simple-model"]))
    for sp in m.species:
        if sp.id in ("A_c2", "B_c3"):
            sp.spatial = False
    dump("very-simple-model-non-spatial-species", m, sbml.to_problem(m))


if __name__ == "__main__":
    main()
