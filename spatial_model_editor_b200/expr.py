"""Reaction expressions -> C / CUDA-C source.

Stands in, on this path, for the reference's ``Symbolic`` wrapper around
SymEngine (core/common/src/symbolic.cpp:134-251): parse an infix expression,
inline constants and user functions, relabel the variables to
``sme_var_<k>`` (cudapixelsim.cpp:108-136) and print C.  The printed string is
what crosses the C-ABI (``smeb200_compartment.expressions``); the same string is
compiled by NVRTC for the GPU and by gcc for the CPU oracle, so both evaluate
the same operation tree (the text format of the reference's ``cudacode`` output
is not pinned by any reference test; numerical results are).
"""
from __future__ import annotations

import math
import re
from typing import Dict, List, Optional, Sequence, Tuple, Union

Node = Union[tuple, float, str]

_TOKEN = re.compile(
    r"\s*(?:(\d+\.?\d*(?:[eE][-+]?\d+)?|\.\d+(?:[eE][-+]?\d+)?)|([A-Za-z_][A-Za-z0-9_]*)|(\*\*|[-+*/^(),]))")

_FUNCS_1 = {"exp", "log", "ln", "sqrt", "sin", "cos", "tan", "abs", "fabs",
            "tanh", "sinh", "cosh", "asin", "acos", "atan", "floor", "ceil"}
_FUNCS_2 = {"pow", "power", "min", "max"}


class ExprError(ValueError):
    pass


def tokenize(s: str) -> List[Tuple[str, str]]:
    out, pos = [], 0
    s = s.strip()
    while pos < len(s):
        m = _TOKEN.match(s, pos)
        if not m:
            raise ExprError(f"cannot parse '{s}' at position {pos}")
        if m.group(1) is not None:
            out.append(("num", m.group(1)))
        elif m.group(2) is not None:
            out.append(("id", m.group(2)))
        else:
            out.append(("op", m.group(3)))
        pos = m.end()
    return out


class _Parser:
    def __init__(self, tokens):
        self.t = tokens
        self.i = 0

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else ("end", "")

    def next(self):
        tok = self.peek()
        self.i += 1
        return tok

    def expect(self, val):
        tok = self.next()
        if tok[1] != val:
            raise ExprError(f"expected '{val}', got '{tok[1]}'")

    def expr(self) -> Node:
        node = self.term()
        while self.peek() in (("op", "+"), ("op", "-")):
            op = self.next()[1]
            rhs = self.term()
            node = ("add" if op == "+" else "sub", node, rhs)
        return node

    def term(self) -> Node:
        node = self.unary()
        while self.peek() in (("op", "*"), ("op", "/")):
            op = self.next()[1]
            rhs = self.unary()
            node = ("mul" if op == "*" else "div", node, rhs)
        return node

    def unary(self) -> Node:
        if self.peek() == ("op", "-"):
            self.next()
            return ("neg", self.unary())
        if self.peek() == ("op", "+"):
            self.next()
            return self.unary()
        return self.power()

    def power(self) -> Node:
        base = self.atom()
        if self.peek() in (("op", "^"), ("op", "**")):
            self.next()
            exponent = self.unary()  # right associative, binds tighter than unary minus on the left
            return ("pow", base, exponent)
        return base

    def atom(self) -> Node:
        kind, val = self.next()
        if kind == "num":
            return float(val)
        if kind == "id":
            if self.peek() == ("op", "("):
                self.next()
                args = []
                if self.peek() != ("op", ")"):
                    args.append(self.expr())
                    while self.peek() == ("op", ","):
                        self.next()
                        args.append(self.expr())
                self.expect(")")
                return ("call", val, tuple(args))
            return val
        if (kind, val) == ("op", "("):
            node = self.expr()
            self.expect(")")
            return node
        raise ExprError(f"unexpected token '{val}'")


def parse(s: str) -> Node:
    p = _Parser(tokenize(s))
    node = p.expr()
    if p.peek()[0] != "end":
        raise ExprError(f"trailing input in '{s}'")
    return node


def substitute(node: Node, env: Dict[str, Node],
               functions: Optional[Dict[str, Tuple[Sequence[str], Node]]] = None) -> Node:
    """Replace identifiers by nodes from env (constants become floats) and
    inline user function calls (SBML functionDefinitions)."""
    functions = functions or {}
    if isinstance(node, float):
        return node
    if isinstance(node, str):
        if node in env:
            v = env[node]
            return float(v) if isinstance(v, (int, float)) else v
        return node
    if node[0] == "call":
        name, args = node[1], tuple(substitute(a, env, functions) for a in node[2])
        if name in functions:
            params, body = functions[name]
            if len(params) != len(args):
                raise ExprError(f"function '{name}' expects {len(params)} arguments")
            return substitute(body, dict(zip(params, args)), functions)
        return ("call", name, args)
    return (node[0],) + tuple(substitute(a, env, functions) for a in node[1:])


def symbols(node: Node, out=None) -> set:
    out = set() if out is None else out
    if isinstance(node, str):
        out.add(node)
    elif isinstance(node, tuple):
        if node[0] == "call":
            for a in node[2]:
                symbols(a, out)
        else:
            for a in node[1:]:
                symbols(a, out)
    return out


def evaluate(node: Node, env: Dict[str, float]) -> float:
    """numeric evaluation (model loader: initial assignments, constants)"""
    if isinstance(node, float):
        return node
    if isinstance(node, str):
        if node not in env:
            raise ExprError(f"unknown symbol '{node}'")
        return float(env[node])
    op = node[0]
    if op == "neg":
        return -evaluate(node[1], env)
    if op in ("add", "sub", "mul", "div", "pow"):
        a, b = evaluate(node[1], env), evaluate(node[2], env)
        return {"add": a + b, "sub": a - b, "mul": a * b,
                "div": a / b if op == "div" else 0.0,
                "pow": math.pow(a, b) if op == "pow" else 0.0}[op]
    if op == "call":
        name, args = node[1], [evaluate(a, env) for a in node[2]]
        if name in ("ln", "log"):
            return math.log(args[0])
        if name in ("pow", "power"):
            return math.pow(args[0], args[1])
        if name in ("abs", "fabs"):
            return abs(args[0])
        if name in ("min", "max"):
            return min(args) if name == "min" else max(args)
        return getattr(math, name)(*args)
    raise ExprError(f"cannot evaluate node {node!r}")


def _num(v: float) -> str:
    if v == int(v) and abs(v) < 1e15:
        return f"{int(v)}.0"
    return repr(float(v))


def to_c(node: Node, var_names: Dict[str, str]) -> str:
    """print fully parenthesised C; integer powers 2..4 become products so no
    libm/libdevice pow sits on the hot path (gray-scott has V^2)"""
    if isinstance(node, float):
        return _num(node) if node >= 0 else f"({_num(node)})"
    if isinstance(node, str):
        if node not in var_names:
            raise ExprError(f"unknown symbol '{node}' in reaction expression")
        return var_names[node]
    op = node[0]
    if op == "neg":
        return f"(-{to_c(node[1], var_names)})"
    if op in ("add", "sub", "mul", "div"):
        sym = {"add": "+", "sub": "-", "mul": "*", "div": "/"}[op]
        return f"({to_c(node[1], var_names)} {sym} {to_c(node[2], var_names)})"
    if op == "pow":
        return _pow_c(node[1], node[2], var_names)
    if op == "call":
        name, args = node[1], node[2]
        if name in ("pow", "power") and len(args) == 2:
            return _pow_c(args[0], args[1], var_names)
        cname = {"ln": "log", "abs": "fabs", "min": "fmin", "max": "fmax"}.get(name, name)
        if name in _FUNCS_1 and len(args) == 1 or name in _FUNCS_2 and len(args) == 2:
            return f"{cname}({', '.join(to_c(a, var_names) for a in args)})"
        raise ExprError(f"unknown function '{name}'")
    raise ExprError(f"cannot print node {node!r}")


def _pow_c(base: Node, exponent: Node, var_names) -> str:
    b = to_c(base, var_names)
    if isinstance(exponent, float) and exponent == int(exponent) and 1 <= exponent <= 4:
        return "(" + " * ".join([b] * int(exponent)) + ")"
    if isinstance(exponent, float) and exponent == 0.5:
        return f"sqrt({b})"
    return f"pow({b}, {to_c(exponent, var_names)})"


def compile_expressions(expressions: Sequence[str], variables: Sequence[str],
                        constants: Optional[Dict[str, float]] = None,
                        functions: Optional[Dict[str, Tuple[Sequence[str], Node]]] = None
                        ) -> List[str]:
    """``Symbolic(expressions, variables, constants)`` + ``relabel`` +
    ``cudaCode(i)``: returns one C expression per input in sme_var_<k>."""
    names = {v: f"sme_var_{k}" for k, v in enumerate(variables)}
    env: Dict[str, Node] = dict(constants or {})
    for v in variables:
        env.pop(v, None)
    return [to_c(substitute(parse(e), env, functions), names) for e in expressions]
