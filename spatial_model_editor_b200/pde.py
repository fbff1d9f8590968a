"""Per-species reaction right-hand sides from reactions x stoichiometry.

Restates ``sme::simulate::Pde`` / ``Reaction`` / ``ReacExpr``
(core/simulate/src/pde.cpp:19-108,171-207; pixelsim_impl.cpp:115-148):

    rhs_i = 0.0 + sum_j ((M_ji * (rate_j)) * (scale))

with global constants and reaction-local parameters inlined as numbers.  The
reference lets SymEngine simplify the sum; here zero-stoichiometry terms and
multiplications by one are dropped, which is the part of that simplification
that matters for cost.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

from . import expr as E


@dataclass
class ReactionDef:
    id: str
    rate: str                              # kinetic law, infix
    stoich: Dict[str, float]               # species id -> net stoichiometry (products +, reactants -)
    parameters: Dict[str, float] = field(default_factory=dict)  # local parameters


def _scaled(node, factor: float):
    if factor == 1.0:
        return node
    if factor == -1.0:
        return ("neg", node)
    return ("mul", float(factor), node)


def build_rhs(species: Sequence[str], reactions: Sequence[ReactionDef],
              constants: Optional[Dict[str, float]] = None, scale: float = 1.0,
              functions=None, substitutions: Optional[Dict[str, float]] = None,
              runtime_params: Sequence[str] = (),
              extra_vars: Sequence[str] = ()) -> List[str]:
    """C expressions (in sme_var_k, k = index into ``species``) of the reaction
    term of every species: what ``ReacExpr.expressions`` holds after the
    reference's Pde step, printed for the C-ABI.

    ``runtime_params``: ids of global or reaction-local parameters that stay
    symbolic (printed as ``sme_param_<k>``, k = position in this list) instead
    of being inlined as numbers; a local parameter of one reaction can be named
    ``"<reaction id>.<parameter id>"``.

    ``extra_vars``: the time / space variables ("time", x, y, z ids) the
    expressions may use; they follow the species in the variable numbering
    (ReacExpr, pixelsim_impl.cpp:124-138) and get no expression here."""
    constants = dict(constants or {})
    substitutions = substitutions or {}
    names = {v: f"sme_var_{k}" for k, v in enumerate(list(species) + list(extra_vars))}
    for k, pid in enumerate(runtime_params):
        names[pid] = f"sme_param_{k}"
    out = []
    for sp in species:
        total = None
        for r in reactions:
            m = float(r.stoich.get(sp, 0.0))
            if m == 0.0:
                continue
            env = dict(constants)
            env.update(r.parameters)
            for k, v in substitutions.items():  # pde.cpp:65-77
                if k in env:
                    env[k] = v
            for v in list(species) + list(extra_vars):
                env.pop(v, None)
            for pid in runtime_params:
                if pid.startswith(r.id + ".") and pid[len(r.id) + 1:] in env:
                    env[pid[len(r.id) + 1:]] = pid  # local parameter -> symbol
                else:
                    env.pop(pid, None)
            rate = E.substitute(E.parse(r.rate), env, functions)
            term = _scaled(_scaled(rate, m), scale)
            total = term if total is None else ("add", total, term)
        out.append(E.to_c(total if total is not None else 0.0, names))
    return out
