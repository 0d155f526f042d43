"""Minimal symbolic integer/boolean expressions for tracing tile programs.

The reference builds TVM TIR PrimExprs while it executes a `@T.prim_func` body
(tilelang/language/v2/builder.py).  This backend only needs to recognise the GEMM program
family, so it records the handful of index forms those programs use — `by * block_M`,
`k * block_K`, `k == 0`, `k % 2` — as small expression trees that can be pattern-matched
(`as_affine`) or evaluated (`evaluate`).
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple, Union

Number = Union[int, float, bool]


class Expr:
    __slots__ = ()
    # Python-visible operators -------------------------------------------------------------
    def __add__(self, o): return BinOp("+", self, wrap(o))
    def __radd__(self, o): return BinOp("+", wrap(o), self)
    def __sub__(self, o): return BinOp("-", self, wrap(o))
    def __rsub__(self, o): return BinOp("-", wrap(o), self)
    def __mul__(self, o): return BinOp("*", self, wrap(o))
    def __rmul__(self, o): return BinOp("*", wrap(o), self)
    def __floordiv__(self, o): return BinOp("//", self, wrap(o))
    def __rfloordiv__(self, o): return BinOp("//", wrap(o), self)
    def __mod__(self, o): return BinOp("%", self, wrap(o))
    def __rmod__(self, o): return BinOp("%", wrap(o), self)
    def __neg__(self): return BinOp("-", Const(0), self)
    def __eq__(self, o): return BinOp("==", self, wrap(o))  # noqa: PLW1641 - symbolic, like tir.PrimExpr
    def __ne__(self, o): return BinOp("!=", self, wrap(o))
    def __lt__(self, o): return BinOp("<", self, wrap(o))
    def __le__(self, o): return BinOp("<=", self, wrap(o))
    def __gt__(self, o): return BinOp(">", self, wrap(o))
    def __ge__(self, o): return BinOp(">=", self, wrap(o))
    __hash__ = object.__hash__

    def __bool__(self):
        raise TypeError("symbolic tile-program expression used as a Python bool; "
                        "data-dependent Python control flow is outside the GEMM program family")

    def __index__(self):
        raise TypeError("symbolic tile-program expression used where a concrete integer is required")


class Const(Expr):
    __slots__ = ("value",)

    def __init__(self, value: Number):
        self.value = value

    def __repr__(self):
        return repr(self.value)


class Var(Expr):
    """A block index (`bx`, `by`), a loop variable or a dynamic shape symbol."""
    __slots__ = ("name", "kind", "extent")

    def __init__(self, name: str, kind: str, extent=None):
        self.name, self.kind, self.extent = name, kind, extent

    def __repr__(self):
        return self.name


class BinOp(Expr):
    __slots__ = ("op", "a", "b")

    def __init__(self, op: str, a: Expr, b: Expr):
        self.op, self.a, self.b = op, a, b

    def __repr__(self):
        return f"({self.a!r} {self.op} {self.b!r})"


class Call(Expr):
    """Intrinsic call such as T.max(x, 0)."""
    __slots__ = ("fn", "args")

    def __init__(self, fn: str, args):
        self.fn, self.args = fn, tuple(wrap(a) for a in args)

    def __repr__(self):
        return f"{self.fn}({', '.join(map(repr, self.args))})"


def wrap(x) -> Expr:
    if isinstance(x, Expr):
        return x
    if isinstance(x, (bool, int, float)):
        return Const(x)
    raise TypeError(f"cannot use {type(x).__name__} in a tile-program index expression")


def is_const(e, value=None) -> bool:
    if isinstance(e, (bool, int, float)):
        return value is None or e == value
    return isinstance(e, Const) and (value is None or e.value == value)


def evaluate(e, env: Dict[Var, Number]) -> Number:
    if isinstance(e, (bool, int, float)):
        return e
    if isinstance(e, Const):
        return e.value
    if isinstance(e, Var):
        return env[e]
    if isinstance(e, BinOp):
        a, b = evaluate(e.a, env), evaluate(e.b, env)
        return {"+": lambda: a + b, "-": lambda: a - b, "*": lambda: a * b, "//": lambda: a // b,
                "%": lambda: a % b, "==": lambda: a == b, "!=": lambda: a != b, "<": lambda: a < b,
                "<=": lambda: a <= b, ">": lambda: a > b, ">=": lambda: a >= b}[e.op]()
    if isinstance(e, Call):
        args = [evaluate(a, env) for a in e.args]
        return {"max": max, "min": min}[e.fn](*args)
    raise TypeError(f"cannot evaluate {e!r}")


def as_affine(e) -> Optional[Tuple[Optional[Var], int, int]]:
    """Match `var * scale + offset` (any association).  Returns (var | None, scale, offset)."""
    if isinstance(e, (bool, int)):
        return (None, 0, int(e))
    if isinstance(e, Const):
        return (None, 0, int(e.value)) if isinstance(e.value, (bool, int)) else None
    if isinstance(e, Var):
        return (e, 1, 0)
    if isinstance(e, BinOp):
        l, r = as_affine(e.a), as_affine(e.b)
        if l is None or r is None:
            return None
        (lv, ls, lo), (rv, rs, ro) = l, r
        if e.op in "+-":
            sign = 1 if e.op == "+" else -1
            if lv is not None and rv is not None and lv is not rv:
                return None
            v = lv if lv is not None else rv
            return (v, ls + sign * rs, lo + sign * ro)
        if e.op == "*":
            if lv is not None and rv is not None:
                return None
            if lv is None:
                return (rv, rs * lo, ro * lo)
            return (lv, ls * ro, lo * ro)
    return None


def as_linear(e) -> Optional[Tuple[Dict[Var, int], int]]:
    """Match `sum_i var_i * c_i + offset` over any number of variables.  Returns ({var: c}, offset)."""
    if isinstance(e, (bool, int)):
        return ({}, int(e))
    if isinstance(e, Const):
        return ({}, int(e.value)) if isinstance(e.value, (bool, int)) else None
    if isinstance(e, Var):
        return ({e: 1}, 0)
    if isinstance(e, BinOp):
        l, r = as_linear(e.a), as_linear(e.b)
        if l is None or r is None:
            return None
        (lt, lo), (rt, ro) = l, r
        if e.op in "+-":
            sign = 1 if e.op == "+" else -1
            out = dict(lt)
            for v, c in rt.items():
                out[v] = out.get(v, 0) + sign * c
            return ({v: c for v, c in out.items() if c != 0}, lo + sign * ro)
        if e.op == "*":
            if lt and rt:
                return None
            terms, scale, off = (rt, lo, ro) if not lt else (lt, ro, lo)
            return ({v: c * scale for v, c in terms.items() if c * scale != 0}, off * scale)
    return None
