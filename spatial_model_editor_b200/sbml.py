"""SBML (level 3 + spatial) model loader: ``.xml`` file -> flat ``Problem``.

Restates, with the standard library's XML parser, what the reference's Pixel
backends pull out of ``sme::model::Model`` in their constructors
(pixelsim.cpp:272-406, cudapixelsim.cpp:857-1150), following SURVEY.md
appendix A:

* sampled-field geometry -> label image (geometry_sampled_field.cpp:99-129:
  samples run z, then SBML row y with y = 0 at the BOTTOM, x fastest);
* voxel size = physical size / image size (model_geometry.cpp:339-350);
* compartments = sampled volumes; simulated species = all but the constant
  non-spatial ones, which become numeric constants together with constant
  parameters and compartment sizes (model_species.cpp:1058-1064,
  model_parameters.cpp:426-463, simulate.cpp:131-156);
* membranes = 2-d compartments with their two adjacent domains (A, B);
* reaction terms through ``pde.build_rhs`` (pde.cpp:44-107), membrane reactions
  scaled by volOverL3 (model_units.cpp:290-314);
* initial concentrations: uniform, sampled field, or analytic expression in
  the voxel-centre coordinates (model_geometry.cpp:1038-1051).

Used to extract the golden fixtures under tests/golden/ from the reference's
own example models, and usable on its own to feed SME models to the B200
backend from Python.
"""
from __future__ import annotations

import math
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

from . import expr as E
from . import geometry as G
from . import problem as P
from .pde import ReactionDef, build_rhs

NS_MATH = "{http://www.w3.org/1998/Math/MathML}"


def _local(tag: str) -> str:
    return tag.rsplit("}", 1)[-1]


def _attr(el, name: str, default=None):
    """attribute by local name, whatever its namespace prefix"""
    for k, v in el.attrib.items():
        if _local(k) == name:
            return v
    return default


def _children(el, name: str):
    return [c for c in el if _local(c.tag) == name]


def _child(el, name: str):
    for c in el:
        if _local(c.tag) == name:
            return c
    return None


def _find_all(el, name: str):
    return [c for c in el.iter() if _local(c.tag) == name]


# ---- MathML -> expression tree (the node format of expr.py) -------------------
_MATHML_BINARY = {"plus": "add", "minus": "sub", "times": "mul", "divide": "div",
                  "power": "pow"}
_MATHML_RELATIONAL = {"lt", "gt", "leq", "geq", "eq", "neq", "and", "or"}


def mathml_to_node(el) -> E.Node:
    tag = _local(el.tag)
    if tag == "math":
        return mathml_to_node(list(el)[0])
    if tag == "cn":
        kind = el.attrib.get("type", "real")
        parts = [el.text or ""] + [(c.tail or "") for c in el]
        if kind == "rational":
            return float(parts[0]) / float(parts[1])
        if kind == "e-notation":
            return float(parts[0]) * 10.0 ** float(parts[1])
        return float(parts[0])
    if tag == "ci":
        return (el.text or "").strip()
    if tag == "csymbol":
        return "time"
    if tag == "exponentiale":
        return math.e
    if tag == "pi":
        return math.pi
    if tag == "true":
        return 1.0
    if tag == "false":
        return 0.0
    if tag == "piecewise":
        pieces = []
        otherwise: E.Node = 0.0
        for c in el:
            if _local(c.tag) == "piece":
                v, cond = list(c)
                pieces.append((mathml_to_node(v), mathml_to_node(cond)))
            elif _local(c.tag) == "otherwise":
                otherwise = mathml_to_node(list(c)[0])
        node = otherwise
        for v, cond in reversed(pieces):
            node = ("call", "piecewise", (cond, v, node))
        return node
    if tag == "apply":
        kids = list(el)
        op = _local(kids[0].tag)
        args = [mathml_to_node(k) for k in kids[1:]]
        if op == "minus" and len(args) == 1:
            return ("neg", args[0])
        if op in ("plus", "times"):
            if not args:
                return 0.0 if op == "plus" else 1.0
            node = args[0]
            for a in args[1:]:
                node = (_MATHML_BINARY[op], node, a)
            return node
        if op in _MATHML_BINARY:
            return (_MATHML_BINARY[op], args[0], args[1])
        if op == "root":
            return ("call", "sqrt", (args[-1],))
        if op == "ci":  # user function call
            return ("call", (kids[0].text or "").strip(), tuple(args))
        if op in _MATHML_RELATIONAL:
            return ("call", op, tuple(args))
        return ("call", op, tuple(args))
    raise E.ExprError(f"unsupported MathML element <{tag}>")


def node_to_infix(node: E.Node) -> str:
    """print a tree as an infix string that expr.parse reads back"""
    if isinstance(node, float):
        return repr(node) if node >= 0 else f"({node!r})"
    if isinstance(node, str):
        return node
    op = node[0]
    if op == "neg":
        return f"(-{node_to_infix(node[1])})"
    if op == "call":
        return f"{node[1]}({', '.join(node_to_infix(a) for a in node[2])})"
    sym = {"add": "+", "sub": "-", "mul": "*", "div": "/", "pow": "^"}[op]
    return f"({node_to_infix(node[1])} {sym} {node_to_infix(node[2])})"


def evaluate_np(node: E.Node, env: Dict[str, np.ndarray]):
    """vectorised evaluation (initial assignments over all voxels)"""
    if isinstance(node, float):
        return node
    if isinstance(node, str):
        if node not in env:
            raise E.ExprError(f"unknown symbol '{node}'")
        return env[node]
    op = node[0]
    if op == "neg":
        return -evaluate_np(node[1], env)
    if op in ("add", "sub", "mul", "div", "pow"):
        a, b = evaluate_np(node[1], env), evaluate_np(node[2], env)
        if op == "add":
            return a + b
        if op == "sub":
            return a - b
        if op == "mul":
            return a * b
        if op == "div":
            return a / b
        return np.power(a, b)
    name, args = node[1], [evaluate_np(a, env) for a in node[2]]
    if name == "piecewise":
        return np.where(np.asarray(args[0]) != 0, args[1], args[2])
    rel = {"lt": np.less, "gt": np.greater, "leq": np.less_equal,
           "geq": np.greater_equal, "eq": np.equal, "neq": np.not_equal,
           "and": np.logical_and, "or": np.logical_or}
    if name in rel:
        return rel[name](args[0], args[1]).astype(np.float64)
    fn = {"ln": np.log, "log": np.log, "abs": np.abs, "fabs": np.abs,
          "power": np.power, "pow": np.power, "min": np.minimum,
          "max": np.maximum}.get(name) or getattr(np, name)
    return fn(*args)


# ---- the parsed model -----------------------------------------------------------
@dataclass
class SbmlSpecies:
    id: str
    compartment: str
    initial: float
    constant: bool
    boundary: bool
    spatial: bool
    diffusion: float = 0.0
    storage: float = 1.0


@dataclass
class SbmlReaction:
    id: str
    location: str
    stoich: Dict[str, float]          # species id -> products minus reactants
    rate: E.Node
    local: Dict[str, float]


@dataclass
class SbmlModel:
    path: str
    species: List[SbmlSpecies]
    compartments: List[Tuple[str, float, int]]      # (id, size, spatial dims)
    parameters: Dict[str, float]                    # constant parameters
    functions: Dict[str, Tuple[List[str], E.Node]]
    rules: Dict[str, E.Node]                        # assignment rules
    initial_assignments: Dict[str, E.Node]
    reactions: List[SbmlReaction]
    labels: np.ndarray                              # [nz, ny, nx], y = 0 at the top
    label_of: Dict[str, int]                        # 3-d compartment id -> label
    membranes: List[Tuple[str, str, str]]           # (id, comp A, comp B)
    origin: Tuple[float, float, float]
    voxel_size: Tuple[float, float, float]
    coord_params: Dict[str, str]                    # parameter id -> "x"|"y"|"z"
    field_params: Dict[str, np.ndarray]             # parameter id -> [nz, ny, nx] (y top)
    vol_over_l3: float
    pixel_options: Dict[str, float] = field(default_factory=dict)
    # SBML events of the form "time >= t: variable = value" (model_events.cpp:197-221,
    # 290-299): (id, time, variable, value).  Simulation applies a parameter event
    # by re-creating the simulator with the new value substituted
    # (simulate.cpp:183-270, 404-445).
    events: List[Tuple[str, float, str, float]] = field(default_factory=list)

    def species_of(self, comp: str) -> List[SbmlSpecies]:
        return [s for s in self.species if s.compartment == comp]

    def is_simulated(self, s: SbmlSpecies) -> bool:
        """model_species.cpp:1058-1064: everything but constant non-spatial"""
        return not (s.constant and not s.spatial)


def _unit(model_el, unit_id: Optional[str]):
    for ud in _find_all(model_el, "unitDefinition"):
        if ud.attrib.get("id") == unit_id:
            u = _find_all(ud, "unit")[0]
            return dict(kind=u.attrib.get("kind"), exponent=float(u.attrib.get("exponent", 1)),
                        scale=float(u.attrib.get("scale", 0)),
                        multiplier=float(u.attrib.get("multiplier", 1)))
    return None


def _vol_over_l3(length, volume) -> float:
    """model_units.cpp:290-314"""
    v = l3 = 1.0
    if volume is not None:
        if volume["kind"] == "metre":
            v = (volume["multiplier"] * 10.0 ** volume["scale"]) ** volume["exponent"]
        elif volume["kind"] == "litre":
            v = volume["multiplier"] * 10.0 ** (volume["scale"] - 3)
    if length is not None and length["kind"] == "metre":
        l3 = (length["multiplier"] * 10.0 ** length["scale"]) ** 3
    return v / l3


def _samples(sf, dtype) -> np.ndarray:
    if _attr(sf, "compression", "uncompressed") != "uncompressed":
        raise ValueError("compressed sampled fields are not supported by this loader")
    return np.array((sf.text or "").split(), dtype=dtype)


def _to_image(values: np.ndarray, nx: int, ny: int, nz: int) -> np.ndarray:
    """SBML sample order (z, y bottom-up, x) -> [nz, ny, nx] with y = 0 at the
    top (geometry_sampled_field.cpp:99-129)"""
    return values.reshape(nz, ny, nx)[:, ::-1, :].copy()


def load_sbml(path: str, geometry: Optional[dict] = None,
              reaction_locations: Optional[Dict[str, str]] = None) -> SbmlModel:
    """``geometry`` (for models without one, like the reference's tests that
    import an image into a non-spatial model, simulate_t.cpp:18-41):
    {"labels": [nz, ny, nx] image with y = 0 at the top, "label_of": {compartment
    id: label}, "voxel_size": (w, h, d)}; membranes are then named
    ``<A>_<B>_membrane`` in label order.  ``reaction_locations`` moves reactions
    (ModelReactions::setLocation, simulate_t.cpp:85-88)."""
    root = ET.parse(path).getroot()
    model = _child(root, "model")
    if model is None:
        model = _find_all(root, "model")[0]
    length = _unit(model, model.attrib.get("lengthUnits"))
    volume = _unit(model, model.attrib.get("volumeUnits"))

    # geometry
    geom = _child(model, "geometry")
    bounds = {}
    coord_ids = {}
    fields = {}
    if geom is None:
        if geometry is None:
            raise ValueError("model has no geometry: pass geometry=")
        return _load_rest(path, model, length, volume, np.asarray(geometry["labels"]),
                          dict(geometry["label_of"]), (0.0, 0.0, 0.0),
                          tuple(geometry.get("voxel_size", (1.0, 1.0, 1.0))), {}, {},
                          reaction_locations, synthesize_membranes=True)
    for cc in _find_all(geom, "coordinateComponent"):
        axis = {"cartesianX": "x", "cartesianY": "y", "cartesianZ": "z"}[_attr(cc, "type")]
        bounds[axis] = (float(_attr(_child(cc, "boundaryMin"), "value")),
                        float(_attr(_child(cc, "boundaryMax"), "value")))
        coord_ids[_attr(cc, "id")] = axis
    sfg = _find_all(geom, "sampledFieldGeometry")[0]
    fields = {_attr(sf, "id"): sf for sf in _find_all(geom, "sampledField")}
    gsf = fields[_attr(sfg, "sampledField")]
    nx, ny = int(_attr(gsf, "numSamples1")), int(_attr(gsf, "numSamples2"))
    nz = int(_attr(gsf, "numSamples3", "1"))
    labels = _to_image(_samples(gsf, np.float64).astype(np.int64), nx, ny, nz)
    value_of_domain = {_attr(sv, "domainType"): int(float(_attr(sv, "sampledValue")))
                       for sv in _find_all(sfg, "sampledVolume")}
    if "z" not in bounds:
        bounds["z"] = (0.0, 1.0)
    origin = (bounds["x"][0], bounds["y"][0], bounds["z"][0])
    voxel_size = ((bounds["x"][1] - bounds["x"][0]) / nx,
                  (bounds["y"][1] - bounds["y"][0]) / ny,
                  (bounds["z"][1] - bounds["z"][0]) / nz)

    label_of = {}
    for c in _find_all(_child(model, "listOfCompartments"), "compartment"):
        cm = _find_all(c, "compartmentMapping")
        if cm and _attr(cm[0], "domainType") in value_of_domain:
            label_of[c.attrib["id"]] = value_of_domain[_attr(cm[0], "domainType")]
    return _load_rest(path, model, length, volume, labels, label_of, origin, voxel_size,
                      coord_ids, fields, reaction_locations, synthesize_membranes=False)


def _divide_by_symbol(node: E.Node, sym: str) -> E.Node:
    """rate / compartment with the factor cancelled symbolically when the rate
    is a product containing it (SimpleSymbolic::divide, model_reactions.cpp:
    117-141): ODE-model kinetic laws are amount/time, spatial ones
    concentration/time"""
    factors = []

    def flatten(n):
        if isinstance(n, tuple) and n[0] == "mul":
            flatten(n[1])
            flatten(n[2])
        else:
            factors.append(n)

    flatten(node)
    if sym in factors:
        factors.remove(sym)
        if not factors:
            return 1.0
        out = factors[0]
        for f in factors[1:]:
            out = ("mul", out, f)
        return out
    return ("div", node, sym)


def _load_rest(path, model, length, volume, labels, label_of, origin, voxel_size, coord_ids,
               fields, reaction_locations, synthesize_membranes) -> SbmlModel:
    compartments = []
    for c in _find_all(_child(model, "listOfCompartments"), "compartment"):
        dims = int(float(c.attrib.get("spatialDimensions", 3)))
        compartments.append((c.attrib["id"], float(c.attrib.get("size", 1.0)), dims))

    # membranes in the reference's order (model_membranes.cpp:105-142): pairs
    # (i < j) of 3-d compartments, j outer; orientation from the membrane id
    ids2d = {cid for cid, _, d in compartments if d == 2}
    comps3d = [cid for cid, _, d in compartments if d == 3 and cid in label_of]
    membranes = []
    for j in range(1, len(comps3d)):
        for i in range(j):
            a, b = comps3d[i], comps3d[j]
            if synthesize_membranes:
                if label_of[b] < label_of[a]:
                    a, b = b, a
                membranes.append((f"{a}_{b}_membrane", a, b))
            elif f"{a}_{b}_membrane" in ids2d:
                membranes.append((f"{a}_{b}_membrane", a, b))
            elif f"{b}_{a}_membrane" in ids2d:
                membranes.append((f"{b}_{a}_membrane", b, a))

    # parameters
    parameters, coord_params, field_params, diffusion = {}, {}, {}, {}
    lop = _child(model, "listOfParameters")
    for p in (_find_all(lop, "parameter") if lop is not None else []):
        pid = p.attrib["id"]
        ref = _find_all(p, "spatialSymbolReference")
        dc = _find_all(p, "diffusionCoefficient")
        if ref:
            r = _attr(ref[0], "spatialRef")
            if r in coord_ids:
                coord_params[pid] = coord_ids[r]
            elif r in fields:
                sf = fields[r]
                fnz = int(_attr(sf, "numSamples3", "1"))
                field_params[pid] = _to_image(
                    _samples(sf, np.float64), int(_attr(sf, "numSamples1")),
                    int(_attr(sf, "numSamples2")), fnz)
            continue
        if dc:
            diffusion[_attr(dc[0], "variable")] = float(p.attrib.get("value", 0.0))
            continue
        if "value" in p.attrib:
            parameters[pid] = float(p.attrib["value"])

    # rules, functions, initial assignments
    rules = {}
    lor = _child(model, "listOfRules")
    for r in (list(lor) if lor is not None else []):
        if _local(r.tag) == "assignmentRule":
            rules[r.attrib["variable"]] = mathml_to_node(_child(r, "math"))
    for v in rules:  # replaced by an AssignmentRule: not a constant
        parameters.pop(v, None)
    functions = {}
    lof = _child(model, "listOfFunctionDefinitions")
    for f in (list(lof) if lof is not None else []):
        lam = list(_child(f, "math"))[0]
        params = [(_find_all(bv, "ci")[0].text or "").strip() for bv in _children(lam, "bvar")]
        body = [k for k in lam if _local(k.tag) != "bvar"][0]
        functions[f.attrib["id"]] = (params, mathml_to_node(body))
    initial_assignments = {}
    loi = _child(model, "listOfInitialAssignments")
    for ia in (list(loi) if loi is not None else []):
        initial_assignments[ia.attrib["symbol"]] = mathml_to_node(_child(ia, "math"))

    # events: trigger "time >= t" (or "t <= time"), one assignment of a number
    # (ModelEvents::getTime / getVariable / getValue, model_events.cpp:168-221,290-299)
    events = []
    loe = _child(model, "listOfEvents")
    for ev in (list(loe) if loe is not None else []):
        trig, asg = _child(ev, "trigger"), _find_all(ev, "eventAssignment")
        if trig is None or not asg:
            continue
        cns = _find_all(trig, "cn")
        maths = _find_all(asg[0], "cn")
        if not cns or not maths:
            continue
        events.append((ev.attrib.get("id", ""), float((cns[0].text or "0").strip()),
                       asg[0].attrib["variable"], float((maths[0].text or "0").strip())))

    # annotation: storage values and pixel options
    storage, pixel_options = {}, {}
    ann = _child(model, "annotation")
    if ann is not None:
        for sv in _find_all(ann, "speciesStorageValues"):
            for item in sv:
                k, v = _child(item, "key"), _child(item, "value")
                if k is not None and v is not None:
                    storage[k.text] = float(v.text)
        for px in _find_all(ann, "pixel"):
            for name in ("integrator", "maxTimestep"):
                el = _child(px, name)
                if el is not None:
                    pixel_options[name] = float(el.text)
            me = _child(px, "maxErr")
            if me is not None:
                pixel_options["maxErrAbs"] = float(_child(me, "abs").text)
                pixel_options["maxErrRel"] = float(_child(me, "rel").text)

    species = []
    for s in _find_all(_child(model, "listOfSpecies"), "species"):
        sid = s.attrib["id"]
        init = s.attrib.get("initialConcentration", s.attrib.get("initialAmount", "0"))
        species.append(SbmlSpecies(
            id=sid, compartment=s.attrib["compartment"], initial=float(init),
            constant=s.attrib.get("constant", "false") == "true",
            boundary=s.attrib.get("boundaryCondition", "false") == "true",
            # a species without the attribute comes from a non-spatial model:
            # SME makes the non-constant ones spatial on import
            spatial=(_attr(s, "isSpatial") == "true") if _attr(s, "isSpatial") is not None
            else s.attrib.get("constant", "false") != "true",
            diffusion=diffusion.get(sid, 0.0), storage=storage.get(sid, 1.0)))

    reactions = []
    lor = _child(model, "listOfReactions")
    by_id = {s.id: s for s in species}
    for r in (_find_all(lor, "reaction") if lor is not None else []):
        stoich: Dict[str, float] = {}
        for lst, sign in (("listOfReactants", -1.0), ("listOfProducts", 1.0)):
            el = _child(r, lst)
            for sr in (_find_all(el, "speciesReference") if el is not None else []):
                sp = by_id[sr.attrib["species"]]
                # model_species.cpp:1066-1075: constant / boundary species do not change
                if sp.constant or sp.boundary:
                    continue
                stoich[sp.id] = stoich.get(sp.id, 0.0) + sign * float(
                    sr.attrib.get("stoichiometry", 1.0))
        kl = _child(r, "kineticLaw")
        if kl is None:
            continue
        local = {lp.attrib["id"]: float(lp.attrib.get("value", 0.0))
                 for lp in _find_all(kl, "localParameter")}
        location = (reaction_locations or {}).get(r.attrib["id"],
                                                  r.attrib.get("compartment", ""))
        if not location:  # non-spatial model: where all its species live
            where = {by_id[sr.attrib["species"]].compartment
                     for sr in _find_all(r, "speciesReference")}
            location = where.pop() if len(where) == 1 else ""
        rate = mathml_to_node(_child(kl, "math"))
        if synthesize_membranes and not r.attrib.get("compartment") \
                and location in {c[0] for c in compartments if c[2] == 3}:
            # ODE model imported into a geometry: compartment reactions are
            # rescaled to concentration/time (expected values of
            # simulate_t.cpp:183-192 pin this: conversion rate 0.3 * A_c3)
            rate = _divide_by_symbol(rate, location)
        reactions.append(SbmlReaction(
            id=r.attrib["id"], location=location,
            stoich={k: v for k, v in stoich.items() if v != 0.0},
            rate=rate, local=local))

    return SbmlModel(path=path, species=species, compartments=compartments,
                     parameters=parameters, functions=functions, rules=rules,
                     initial_assignments=initial_assignments, reactions=reactions,
                     labels=labels, label_of=label_of, membranes=membranes, origin=origin,
                     voxel_size=voxel_size, coord_params=coord_params,
                     field_params=field_params,
                     vol_over_l3=_vol_over_l3(length, volume), pixel_options=pixel_options,
                     events=events)


# ---- SbmlModel -> Problem --------------------------------------------------------
def global_constants(m: SbmlModel) -> Dict[str, float]:
    """model_parameters.cpp:426-463"""
    c: Dict[str, float] = {}
    for s in m.species:
        if s.constant and not s.spatial:
            c[s.id] = s.initial
    c.update(m.parameters)
    for cid, size, _ in m.compartments:
        c[cid] = size
    return c


def _resolved_rules(m: SbmlModel) -> Dict[str, E.Node]:
    rules = dict(m.rules)
    for _ in range(8):  # rules may refer to each other
        rules = {k: E.substitute(v, {r: n for r, n in rules.items() if r != k}, m.functions)
                 for k, v in rules.items()}
    return rules


def voxel_centres(m: SbmlModel, voxels: np.ndarray) -> Dict[str, np.ndarray]:
    """model_geometry.cpp:1038-1051: y is flipped (image row 0 is the top)"""
    nz, ny, nx = m.labels.shape
    ox, oy, oz = m.origin
    dx, dy, dz = m.voxel_size
    return {"x": ox + dx * (voxels[:, 0] + 0.5),
            "y": oy + dy * ((ny - 1 - voxels[:, 1]) + 0.5),
            "z": oz + dz * (voxels[:, 2] + 0.5)}


def initial_concentration(m: SbmlModel, s: SbmlSpecies, voxels: np.ndarray,
                          constants: Dict[str, float]) -> np.ndarray:
    n = len(voxels)
    node = m.initial_assignments.get(s.id)
    if node is None:
        return np.full(n, s.initial, dtype=np.float64)
    centres = voxel_centres(m, voxels)
    env: Dict[str, object] = dict(constants)
    for pid, axis in m.coord_params.items():
        env[pid] = centres[axis]
    for pid, img in m.field_params.items():
        z = np.minimum(voxels[:, 2], img.shape[0] - 1)
        env[pid] = img[z, voxels[:, 1], voxels[:, 0]]
    node = E.substitute(node, {}, m.functions)
    out = evaluate_np(node, env)
    return np.broadcast_to(np.asarray(out, dtype=np.float64), (n,)).copy()


def parameter_value(m: SbmlModel, pid: str) -> float:
    """current value of a global parameter or of ``"<reaction>.<local parameter>"``"""
    if "." in pid:
        rid, lid = pid.split(".", 1)
        for r in m.reactions:
            if r.id == rid and lid in r.local:
                return r.local[lid]
    return global_constants(m)[pid]


def to_problem(m: SbmlModel, integrator: Optional[int] = None,
               substitutions: Optional[Dict[str, float]] = None,
               runtime_params: Tuple[str, ...] = ()) -> P.Problem:
    """what PixelSim / CudaPixelSim build from the Model, as a flat Problem.
    ``runtime_params`` stay symbolic (see pde.build_rhs); their current values
    become ``Problem.params``."""
    constants = global_constants(m)
    rules = _resolved_rules(m)
    env_rules = {k: v for k, v in rules.items()}
    nz, ny, nx = m.labels.shape
    unsupported = 0
    # reactions that depend on time / space: every compartment's state is padded
    # with "time" and / or "x, y, z" components (pixelsim.cpp:282-322,
    # pixelsim_impl.cpp:124-147,312-349)
    axis_id = {axis: pid for pid, axis in m.coord_params.items()}
    coord_ids = [axis_id.get(a, f"sme_coord_{a}") for a in ("x", "y", "z")]
    time_dependent = space_dependent = False
    for r in m.reactions:
        if not r.stoich:
            continue
        rate = E.substitute(r.rate, {k: v for k, v in env_rules.items()
                                     if k not in r.local}, m.functions)
        syms = E.symbols(rate)
        time_dependent = time_dependent or "time" in syms
        space_dependent = space_dependent or bool(syms & set(m.coord_params))
    extra_vars = (["time"] if time_dependent else []) + (coord_ids if space_dependent else [])
    extra_exprs = (["1.0"] if time_dependent else []) + (["0.0"] * 3 if space_dependent else [])

    def reaction_defs(location: str) -> List[ReactionDef]:
        out = []
        for r in m.reactions:
            if r.location != location or not r.stoich:
                continue
            rate = E.substitute(r.rate, {k: v for k, v in env_rules.items()
                                         if k not in r.local}, m.functions)
            out.append(ReactionDef(r.id, node_to_infix(rate), dict(r.stoich), dict(r.local)))
        return out

    comps: List[P.Compartment] = []
    comp_index: Dict[str, int] = {}
    for cid, _, dims in m.compartments:
        if cid not in m.label_of:  # membranes / unsampled compartments
            continue
        sim_species = [s for s in m.species_of(cid) if m.is_simulated(s)]
        if not sim_species:
            continue
        voxels = G.compartment_voxels(m.labels, m.label_of[cid])
        ids = [s.id for s in sim_species]
        non_spatial = [k for k, s in enumerate(sim_species) if not s.spatial]
        storage = [s.storage for s in sim_species] + [1.0] * len(extra_vars)
        dx, dy, dz = m.voxel_size
        # constant species do not diffuse (pixelsim_impl.cpp:233-235)
        dcs = [0.0 if s.constant else s.diffusion for s in sim_species] + [0.0] * len(extra_vars)
        diffusion = np.array([[d / (dx * dx), d / (dy * dy), d / (dz * dz)] for d in dcs])
        columns = [initial_concentration(m, s, voxels, constants) for s in sim_species]
        if time_dependent:
            columns.append(np.zeros(len(voxels)))
        if space_dependent:
            centres = voxel_centres(m, voxels)
            columns += [centres["x"], centres["y"], centres["z"]]
        conc0 = np.stack(columns, axis=1)
        exprs = build_rhs(ids, reaction_defs(cid), constants, functions=m.functions,
                          substitutions=substitutions, runtime_params=runtime_params,
                          extra_vars=extra_vars) + extra_exprs
        comp_index[cid] = len(comps)
        comps.append(P.Compartment(
            id=cid, species=ids + extra_vars, expressions=exprs, diffusion=diffusion,
            conc0=np.ascontiguousarray(conc0), nn=G.neighbour_table(m.labels.shape, voxels),
            voxels=voxels, n_primary=len(ids), non_spatial=non_spatial,
            inv_storage=(None if all(v == 1.0 for v in storage)
                         else np.array([1.0 / v if v != 0.0 else 0.0 for v in storage]))))

    membranes: List[P.Membrane] = []
    colour_index = {label: i for i, label in enumerate(sorted(set(m.label_of.values())))}
    pairs = None
    for mid, a, b in m.membranes:
        defs = reaction_defs(mid)
        if not defs:
            continue
        ia, ib = comp_index.get(a, -1), comp_index.get(b, -1)
        names = ((comps[ia].species[:comps[ia].n_primary_species] if ia >= 0 else [])
                 + (comps[ib].species[:comps[ib].n_primary_species] if ib >= 0 else []))
        if not names:
            continue
        if pairs is None:
            # orientation: the membrane's compartment A must take the lower
            # colour index (model_membranes.cpp:114-120)
            order = {}
            k = 0
            for _, ca, cb in m.membranes:
                for c in (ca, cb):
                    if m.label_of[c] not in order:
                        order[m.label_of[c]] = k
                        k += 1
            for label in m.label_of.values():
                if label not in order:
                    order[label] = k
                    k += 1
            colour_index = order
            pairs = G.membrane_voxel_pairs(m.labels, colour_index)
        la, lb = m.label_of[a], m.label_of[b]
        key = (la, lb) if colour_index[la] < colour_index[lb] else (lb, la)
        vp = pairs.get(key, [])
        if key != (la, lb):
            vp = [(q, p_) for p_, q in vp]
        va = G.compartment_voxels(m.labels, la)
        vb = G.compartment_voxels(m.labels, lb)
        face_pairs = G.membrane_face_pairs(m.labels.shape, va, vb, vp)
        exprs = build_rhs(names, defs, constants, scale=m.vol_over_l3,
                          functions=m.functions, substitutions=substitutions,
                          runtime_params=runtime_params, extra_vars=extra_vars)
        membranes.append(P.Membrane(id=mid, comp_a=ia, comp_b=ib, pairs=face_pairs,
                                    flux_length=G.flux_lengths(m.voxel_size),
                                    expressions=exprs))
    opts = m.pixel_options
    return P.Problem(
        comps, membranes,
        integrator=int(opts.get("integrator", P.RK212)) if integrator is None else integrator,
        max_err_abs=opts.get("maxErrAbs", P.DBL_MAX), max_err_rel=opts.get("maxErrRel", 0.005),
        max_timestep=opts.get("maxTimestep", P.DBL_MAX), image_size=(nx, ny, nz),
        unsupported=unsupported,
        params=[parameter_value(m, pid) for pid in runtime_params])
