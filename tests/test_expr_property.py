"""Property tests (CPU only) of the expression front end: for random well-typed trees the Python
lowering (igloo_b200/expr.py) and the library's igl_expr_compile must agree on result type and
nullability; random byte strings must be rejected or accepted cleanly, never crash."""
import ctypes

import pytest
from hypothesis import given, settings, strategies as st

from igloo_b200._lib import lib
from igloo_b200.expr import Field, col, lit, lower, T_BOOL, T_F64, T_I32, T_I64

SCHEMA = [Field("a", T_I32), Field("b", T_I32, True), Field("c", T_I64, True), Field("d", T_I64), Field("x", T_F64),
          Field("y", T_F64, True), Field("p", T_BOOL), Field("q", T_BOOL, True)]
TYPES = (ctypes.c_int32 * len(SCHEMA))(*[f.type for f in SCHEMA])
NULLS = bytes(int(f.nullable) for f in SCHEMA)

lib.igl_expr_compile.restype = ctypes.c_int32
lib.igl_expr_compile.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_int32),
                                 ctypes.c_char_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_void_p)]
lib.igl_expr_type.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]
lib.igl_expr_destroy.argtypes = [ctypes.c_void_p]


def compile_(bc):
    h = ctypes.c_void_p()
    rc = lib.igl_expr_compile(None, bc, len(bc), TYPES, NULLS, len(SCHEMA), ctypes.byref(h))
    return rc, h


def leaves(t):
    cols = [col(f.name) for f in SCHEMA if f.type == t]
    lits = {T_I32: st.integers(-2**31, 2**31 - 1).map(lambda v: lit(v, T_I32)),
            T_I64: st.integers(-2**63, 2**63 - 1).map(lambda v: lit(v, T_I64)),
            T_F64: st.floats(allow_nan=False, allow_infinity=False).map(lambda v: lit(v, T_F64)),
            T_BOOL: st.booleans().map(lambda v: lit(v, T_BOOL))}[t]
    return st.one_of(st.sampled_from(cols), lits, st.just(lit(None, t)))


def numeric(t, depth):
    if depth == 0:
        return leaves(t)
    sub = numeric(t, depth - 1)
    others = [u for u in (T_I32, T_I64, T_F64) if u != t]
    return st.one_of(
        leaves(t),
        st.tuples(sub, sub, st.sampled_from("+-*/%")).map(lambda a: {"+": a[0] + a[1], "-": a[0] - a[1], "*": a[0] * a[1],
                                                                    "/": a[0] / a[1], "%": a[0] % a[1]}[a[2]]),
        sub.map(lambda e: -e),
        st.sampled_from(others).flatmap(lambda u: numeric(u, depth - 1)).map(lambda e: e.cast(t)),
    )


def boolean(depth):
    if depth == 0:
        return leaves(T_BOOL)
    sub = boolean(depth - 1)
    cmp_ = st.sampled_from([T_I32, T_I64, T_F64]).flatmap(
        lambda t: st.tuples(numeric(t, depth - 1), numeric(t, depth - 1), st.sampled_from(["<", "<=", ">", ">=", "==", "!="])))
    return st.one_of(
        leaves(T_BOOL),
        cmp_.map(lambda a: {"<": a[0] < a[1], "<=": a[0] <= a[1], ">": a[0] > a[1], ">=": a[0] >= a[1],
                            "==": a[0] == a[1], "!=": a[0] != a[1]}[a[2]]),
        st.tuples(sub, sub, st.booleans()).map(lambda a: (a[0] & a[1]) if a[2] else (a[0] | a[1])),
        sub.map(lambda e: ~e),
        st.sampled_from([T_I32, T_I64, T_F64, T_BOOL]).flatmap(lambda t: numeric(t, depth - 1) if t != T_BOOL else sub)
          .flatmap(lambda e: st.sampled_from([e.is_null(), e.is_not_null()])),
    )


@settings(max_examples=300, deadline=None)
@given(st.one_of(boolean(3), numeric(T_I32, 3), numeric(T_I64, 3), numeric(T_F64, 3)))
def test_lowering_and_compile_agree_on_type_and_nullability(e):
    bc, t, nullable = lower(e, SCHEMA)
    rc, h = compile_(bc)
    assert rc == 0, (lib.igl_last_error(), bc.hex())
    tt, nn = ctypes.c_int32(), ctypes.c_int32()
    lib.igl_expr_type(h, ctypes.byref(tt), ctypes.byref(nn))
    lib.igl_expr_destroy(h)
    assert (tt.value, bool(nn.value)) == (t, bool(nullable))


@settings(max_examples=500, deadline=None)
@given(st.binary(max_size=64))
def test_arbitrary_bytes_never_crash_the_compiler(bc):
    rc, h = compile_(bc)
    assert rc in (0, -1, -2)
    if rc == 0:
        lib.igl_expr_destroy(h)
