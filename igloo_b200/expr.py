"""Logical expressions and their lowering to the igloo-cuda expression bytecode.

Host-side counterpart of what the reference's planner does with DataFusion
`Expr`s: `create_physical_expr(expr, &input_schema, ..)` followed by
`expr.display_name()` / `data_type()` / `nullable()` for the output field
(crates/engine/src/physical_planner.rs:60-71, :86-90, :107-116).  The lowered
form is the postfix byte string of include/igloo_expr.h instead of an
`Arc<dyn PhysicalExpr>`.

Type coercion follows what DataFusion leaves behind after
`into_optimized_plan()` for the subset the configs use:
  * both sides of a binary operator end up with one type (I32 < I64 < F64);
  * an integer literal compared with / added to an I32 column stays I32 when it
    fits (DataFusion's unwrap-cast-in-comparison), so `age > 25` is
    `age > Int32(25)`;
  * nothing is coerced silently inside the CUDA library: igl_expr_compile
    rejects mixed types.
"""
from __future__ import annotations

import struct
from typing import List, Optional, Sequence, Tuple

T_BOOL, T_I32, T_I64, T_F64, T_UTF8 = 1, 2, 3, 4, 5
TYPE_NAMES = {T_BOOL: "Boolean", T_I32: "Int32", T_I64: "Int64", T_F64: "Float64", T_UTF8: "Utf8"}

OP_COL, OP_LIT_BOOL, OP_LIT_I32, OP_LIT_I64, OP_LIT_F64, OP_LIT_NULL = 0x01, 0x02, 0x03, 0x04, 0x05, 0x06
OP_EQ, OP_NE, OP_LT, OP_LE, OP_GT, OP_GE = 0x10, 0x11, 0x12, 0x13, 0x14, 0x15
OP_ADD, OP_SUB, OP_MUL, OP_DIV, OP_MOD, OP_NEG = 0x20, 0x21, 0x22, 0x23, 0x24, 0x25
OP_AND, OP_OR, OP_NOT, OP_IS_NULL, OP_IS_NOT_NULL = 0x30, 0x31, 0x32, 0x33, 0x34
OP_CAST = 0x40

_CMP = {OP_EQ: "=", OP_NE: "!=", OP_LT: "<", OP_LE: "<=", OP_GT: ">", OP_GE: ">="}
_ARITH = {OP_ADD: "+", OP_SUB: "-", OP_MUL: "*", OP_DIV: "/", OP_MOD: "%"}
_LOGIC = {OP_AND: "AND", OP_OR: "OR"}


class Field:
    """(name, type, nullable) — arrow::datatypes::Field for the supported subset."""

    def __init__(self, name: str, type_: int, nullable: bool = False):
        self.name, self.type, self.nullable = name, type_, nullable

    def __repr__(self):
        return f"Field({self.name!r}, {TYPE_NAMES[self.type]}, nullable={self.nullable})"

    def __eq__(self, o):
        return (self.name, self.type, self.nullable) == (o.name, o.type, o.nullable)


class PlanError(ValueError):
    """DataFusionError::Plan / ::NotImplemented analogue raised while lowering."""


class Expr:
    def __init__(self, kind: str, op: int = 0, args: Sequence["Expr"] = (), name: str = "", value=None,
                 type_: int = 0, alias: Optional[str] = None):
        self.kind, self.op, self.args, self.name, self.value, self.type, self._alias = (
            kind, op, tuple(args), name, value, type_, alias)

    # -- construction helpers ------------------------------------------------
    def _bin(self, op, other, swap=False):
        o = other if isinstance(other, Expr) else lit(other)
        return Expr("bin", op, (o, self) if swap else (self, o))

    def __add__(self, o): return self._bin(OP_ADD, o)
    def __radd__(self, o): return self._bin(OP_ADD, o, True)
    def __sub__(self, o): return self._bin(OP_SUB, o)
    def __rsub__(self, o): return self._bin(OP_SUB, o, True)
    def __mul__(self, o): return self._bin(OP_MUL, o)
    def __rmul__(self, o): return self._bin(OP_MUL, o, True)
    def __truediv__(self, o): return self._bin(OP_DIV, o)
    def __rtruediv__(self, o): return self._bin(OP_DIV, o, True)
    def __mod__(self, o): return self._bin(OP_MOD, o)
    def __lt__(self, o): return self._bin(OP_LT, o)
    def __le__(self, o): return self._bin(OP_LE, o)
    def __gt__(self, o): return self._bin(OP_GT, o)
    def __ge__(self, o): return self._bin(OP_GE, o)
    def __eq__(self, o): return self._bin(OP_EQ, o)  # type: ignore[override]
    def __ne__(self, o): return self._bin(OP_NE, o)  # type: ignore[override]
    def __and__(self, o): return self._bin(OP_AND, o)
    def __or__(self, o): return self._bin(OP_OR, o)
    def __invert__(self): return Expr("un", OP_NOT, (self,))
    def __neg__(self): return Expr("un", OP_NEG, (self,))
    __hash__ = object.__hash__

    def is_null(self): return Expr("un", OP_IS_NULL, (self,))
    def is_not_null(self): return Expr("un", OP_IS_NOT_NULL, (self,))
    def cast(self, type_: int): return Expr("cast", OP_CAST, (self,), type_=type_)

    def alias(self, name: str) -> "Expr":
        e = Expr(self.kind, self.op, self.args, self.name, self.value, self.type)
        e._alias = name
        return e

    # -- display_name (physical_planner.rs:66) -------------------------------
    def display_name(self) -> str:
        if self._alias is not None:
            return self._alias
        return self._display()

    def _display(self) -> str:
        if self.kind == "col":
            return self.name
        if self.kind == "lit":
            if self.value is None:
                return "NULL"
            return f"{TYPE_NAMES[self.type]}({self.value})"
        if self.kind == "bin":
            sym = _CMP.get(self.op) or _ARITH.get(self.op) or _LOGIC[self.op]
            return f"{self.args[0]._display()} {sym} {self.args[1]._display()}"
        if self.kind == "un":
            a = self.args[0]._display()
            return {OP_NOT: f"NOT {a}", OP_NEG: f"(- {a})", OP_IS_NULL: f"{a} IS NULL",
                    OP_IS_NOT_NULL: f"{a} IS NOT NULL"}[self.op]
        if self.kind == "cast":
            return f"CAST({self.args[0]._display()} AS {TYPE_NAMES[self.type]})"
        raise PlanError(f"bad expression kind {self.kind}")

    def __repr__(self):
        return f"Expr<{self.display_name()}>"


def col(name: str) -> Expr:
    return Expr("col", name=name)


def lit(value, type_: Optional[int] = None) -> Expr:
    if type_ is None:
        if value is None:
            raise PlanError("a NULL literal needs an explicit type")
        if isinstance(value, bool):
            type_ = T_BOOL
        elif isinstance(value, int):
            type_ = T_I64  # DataFusion parses integer literals as Int64
        elif isinstance(value, float):
            type_ = T_F64
        else:
            raise PlanError(f"unsupported literal {value!r}")
    return Expr("lit", value=value, type_=type_)


# -----------------------------------------------------------------------------
# Lowering
# -----------------------------------------------------------------------------
def _emit_lit(value, t: int) -> bytes:
    if value is None:
        return bytes([OP_LIT_NULL, t])
    if t == T_BOOL:
        return bytes([OP_LIT_BOOL, 1 if value else 0])
    if t == T_I32:
        return bytes([OP_LIT_I32]) + struct.pack("<i", int(value))
    if t == T_I64:
        return bytes([OP_LIT_I64]) + struct.pack("<q", int(value))
    if t == T_F64:
        return bytes([OP_LIT_F64]) + struct.pack("<d", float(value))
    raise PlanError(f"unsupported literal type {t}")


_RANK = {T_I32: 1, T_I64: 2, T_F64: 3}


def _coerce_pair(a: Expr, ta: int, b: Expr, tb: int) -> Tuple[Expr, Expr, int]:
    """Numeric binary-operator coercion; literals adapt to the column side."""
    if ta == tb:
        return a, b, ta
    if ta not in _RANK or tb not in _RANK:
        raise PlanError(f"cannot coerce {TYPE_NAMES.get(ta)} and {TYPE_NAMES.get(tb)}")

    def retype_lit(e: Expr, to: int) -> Optional[Expr]:
        if e.kind != "lit" or e.value is None:
            return None
        if to == T_F64:
            return lit(float(e.value), T_F64)
        if to == T_I32 and isinstance(e.value, int) and -2**31 <= e.value < 2**31:
            return lit(e.value, T_I32)
        if to == T_I64 and isinstance(e.value, int):
            return lit(e.value, T_I64)
        return None

    # a literal follows the other side when it can be represented there
    if b.kind == "lit" and a.kind != "lit":
        r = retype_lit(b, ta)
        if r is not None and not (ta != T_F64 and tb == T_F64):
            return a, r, ta
    if a.kind == "lit" and b.kind != "lit":
        r = retype_lit(a, tb)
        if r is not None and not (tb != T_F64 and ta == T_F64):
            return r, b, tb
    t = ta if _RANK[ta] > _RANK[tb] else tb
    ca = a if ta == t else (retype_lit(a, t) or a.cast(t))
    cb = b if tb == t else (retype_lit(b, t) or b.cast(t))
    return ca, cb, t


def lower(e: Expr, schema: Sequence[Field]) -> Tuple[bytes, int, bool]:
    """Expr -> (bytecode, data_type, nullable) against `schema`.

    Mirrors create_physical_expr + data_type() + nullable()
    (physical_planner.rs:60-71)."""
    if e.kind == "col":
        for i, f in enumerate(schema):
            if f.name == e.name:
                return bytes([OP_COL]) + struct.pack("<H", i), f.type, f.nullable
        raise PlanError(f"No field named {e.name!r}")
    if e.kind == "lit":
        return _emit_lit(e.value, e.type), e.type, e.value is None
    if e.kind == "cast":
        bc, t, nl = lower(e.args[0], schema)
        if t == e.type:
            return bc, t, nl
        if t not in _RANK or e.type not in _RANK:
            raise PlanError(f"unsupported CAST {TYPE_NAMES.get(t)} -> {TYPE_NAMES.get(e.type)}")
        return bc + bytes([OP_CAST, e.type]), e.type, nl
    if e.kind == "un":
        bc, t, nl = lower(e.args[0], schema)
        if e.op == OP_NOT:
            if t != T_BOOL:
                raise PlanError("NOT needs a boolean operand")
            return bc + bytes([OP_NOT]), T_BOOL, nl
        if e.op == OP_NEG:
            if t not in _RANK:
                raise PlanError("unary minus needs a numeric operand")
            return bc + bytes([OP_NEG]), t, nl
        return bc + bytes([e.op]), T_BOOL, False
    if e.kind == "bin":
        a, b = e.args
        _, ta, _ = lower(a, schema)
        _, tb, _ = lower(b, schema)
        if e.op in _LOGIC:
            if ta != T_BOOL or tb != T_BOOL:
                raise PlanError("AND/OR need boolean operands")
            rt = T_BOOL
        else:
            if ta == T_BOOL and tb == T_BOOL and e.op in _CMP:
                pass
            else:
                a, b, _t = _coerce_pair(a, ta, b, tb)
            rt = T_BOOL if e.op in _CMP else None
        ba, ta, na = lower(a, schema)
        bb, tb, nb = lower(b, schema)
        if ta != tb:
            raise PlanError(f"type mismatch {TYPE_NAMES.get(ta)} vs {TYPE_NAMES.get(tb)}")
        if ta == T_UTF8:
            raise PlanError("Utf8 columns are pass-through only")
        if rt is None:
            if ta not in _RANK:
                raise PlanError("arithmetic needs numeric operands")
            rt = ta
        return ba + bb + bytes([e.op]), rt, (na or nb)
    raise PlanError(f"bad expression kind {e.kind}")


def projection_field(e: Expr, schema: Sequence[Field]) -> Field:
    """Output field of one projection expression (physical_planner.rs:66-71)."""
    _, t, nl = lower(e, schema)
    return Field(e.display_name(), t, nl)
