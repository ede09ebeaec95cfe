"""GPU parity: fused Filter/Projection kernels (through the C ABI) vs the CPU oracle."""
import json
import os

import numpy as np
import pyarrow as pa
import pytest

from igloo_b200.expr import Field, col, lit, lower, T_BOOL, T_I32, T_I64, T_F64
from oracle import oracle as orc
from tests.util import arrow_to_columns, assert_same_multiset, columns_to_arrow, tight_fields

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from igloo_b200.device import Context
    c = Context(0)
    yield c
    c.close()


def _fields(tbl):
    return tight_fields(tbl)


def gpu_filter_project(ctx, tbl, pred, proj, two_pass=None, expect_specialised=None):
    """two_pass: None = the library's own choice by input size; False / True force the one-pass
    (in-kernel ordered scan) or two-pass (survivor bitmap + tile offsets) form of the filter."""
    from igloo_b200.device import DeviceColumn, FilterProject
    from igloo_b200.arrow import schema_of
    fields = _fields(tbl)
    cols = [DeviceColumn.from_arrow(tbl.column(i), ctx.device) for i in range(tbl.num_columns)]
    old = os.environ.pop("IGL_FP_TWO_PASS", None)
    try:
        if two_pass is not None:
            os.environ["IGL_FP_TWO_PASS"] = "1" if two_pass else "0"
        fp = FilterProject(ctx, fields, pred, proj)   # the switch is read when the plan is created
    finally:
        os.environ.pop("IGL_FP_TWO_PASS", None)
        if old is not None:
            os.environ["IGL_FP_TWO_PASS"] = old
    out, n = fp.run(cols, tbl.num_rows)
    if expect_specialised is not None and tbl.num_rows > 0:
        assert fp.is_specialised == expect_specialised
    arrays = [c.to_arrow(n) for c in out]
    return pa.Table.from_arrays(arrays, schema=schema_of(fp.out_fields))


def oracle_filter_project(tbl, pred, proj):
    fields = _fields(tbl)
    t = orc.Table.scan(arrow_to_columns(tbl))
    if pred is not None:
        t = t.filter(lower(pred, fields)[0])
    if proj:
        t = t.project([lower(e, fields)[0] for e in proj], [e.display_name() for e in proj])
    return columns_to_arrow(t.columns()), t


def check(ctx, tbl, pred, proj, exact_order=True):
    want, _ = oracle_filter_project(tbl, pred, proj)
    for two_pass in ([False, True] if pred is not None else [None]):
        got = _check_one(gpu_filter_project(ctx, tbl, pred, proj, two_pass), want, exact_order)
    return got


def _check_one(got, want, exact_order):
    assert got.num_rows == want.num_rows
    assert [f.type for f in got.schema] == [f.type for f in want.schema]
    assert got.schema.names == want.schema.names
    if exact_order:  # the kernel preserves row order, like filter_record_batch
        for i in range(got.num_columns):
            a, b = got.column(i).combine_chunks(), want.column(i).combine_chunks()
            assert a.null_count == b.null_count
            if a.type == pa.float64():  # bit-exact, NaN-safe
                av = a.fill_null(0.0).to_numpy(zero_copy_only=False).view(np.int64)
                bv = b.fill_null(0.0).to_numpy(zero_copy_only=False).view(np.int64)
                assert np.array_equal(av, bv)
                assert a.is_null().equals(b.is_null())
            else:
                assert a.equals(b), (i, a[:10], b[:10])
    else:
        assert_same_multiset(got, want)
    return got


def c2_table(n, seed=0x16100002):
    return pa.table({"a": orc.gen_i32_uniform(seed, 0, 0, n, 1000), "b": orc.gen_i32_uniform(seed, 1, 0, n, 1 << 31),
                     "x": orc.gen_f64_unit(seed, 2, 0, n)})


@pytest.mark.parametrize("n", [0, 1, 31, 127, 895, 896, 1792, 1793, 5000, 200_003])
def test_c2_shapes_and_ragged_sizes(ctx, n):
    t = c2_table(n)
    check(ctx, t, col("a") < 250, [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")])


@pytest.mark.parametrize("thr", [0, 10, 500, 900, 1000])
def test_c2_selectivity_sweep(ctx, thr):
    t = c2_table(300_000)
    check(ctx, t, col("a") < thr, [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")])


def test_filter_only_passes_all_columns_and_projection_only(ctx):
    t = c2_table(100_000)
    got = check(ctx, t, col("a") < 100, [])
    assert got.schema.names == ["a", "b", "x"]
    check(ctx, t, None, [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")])


def test_golden_users_numeric_part(ctx):
    """integration_test.rs:21-29: WHERE age > 30 -> ages {32, 35} (Utf8 name is
    exercised through the operator API)."""
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "users.json")))
    t = pa.table({"id": pa.array(g["rows"]["id"], pa.int32()), "age": pa.array(g["rows"]["age"], pa.int32())})
    got = check(ctx, t, col("age") > 30, [col("id"), col("age")])
    assert sorted(got.column("age").to_pylist()) == [32, 35]
    got = check(ctx, t, col("age") > 25, [col("age")])
    assert sorted(got.column("age").to_pylist()) == [28, 30, 32, 35]


def test_literal_projection(ctx):
    t = pa.table({"d": pa.array([0], pa.int32())})
    got = check(ctx, t, None, [lit(42).alias("answer")])
    assert got.column("answer").to_pylist() == [42] and got.schema.field("answer").type == pa.int64()


def test_nulls_kleene_and_bool_outputs(ctx):
    rng = np.random.default_rng(5)
    n = 50_000
    a = pa.array(rng.integers(-50, 50, n), pa.int32(), mask=rng.random(n) < 0.2)
    b = pa.array(rng.random(n) < 0.5, pa.bool_(), mask=rng.random(n) < 0.3)
    c = pa.array(rng.integers(-2**40, 2**40, n), pa.int64(), mask=rng.random(n) < 0.1)
    t = pa.table({"a": a, "b": b, "c": c})
    check(ctx, t, (col("a") > 0) | col("b"), [])
    check(ctx, t, (col("a") > 0) & col("b"), [col("a"), col("c")])
    check(ctx, t, ~col("b"), [col("a").is_null().alias("n"), col("c").is_not_null().alias("m"), (col("a") > 3).alias("g")])
    check(ctx, t, col("a").is_null(), [])
    check(ctx, t, None, [(col("a").cast(T_I64) + col("c")).alias("s"), (-col("a")).alias("neg"), (col("a") == col("a")).alias("eq")])
    check(ctx, t, lit(None, T_BOOL), [])       # NULL predicate drops everything
    check(ctx, t, None, [lit(None, T_I64).alias("nul"), col("b")])


def test_wrapping_division_cast(ctx):
    from igloo_b200._lib import IglooCudaError
    a = np.array([2**31 - 1, -2**31, 7, -7, 0, 123456], np.int32)
    t = pa.table({"a": a, "b": np.array([2, 3, -2, 4, 5, -1], np.int32), "x": np.array([1.9, -1.9, 0.0, 1e9, -3.5, 2.5])})
    check(ctx, t, None, [(col("a") + lit(1, T_I32)).alias("p"), (col("a") * lit(3, T_I32)).alias("m"), (-col("a")).alias("n"),
                         (col("a") - col("b")).alias("d")])
    check(ctx, t.slice(2), None, [(col("a") / col("b")).alias("q"), (col("a") % col("b")).alias("r")])
    check(ctx, t, None, [(col("x") / lit(0.0)).alias("q"), (col("x") % lit(0.7)).alias("r"), col("x").cast(T_I32).alias("i"),
                         col("a").cast(T_F64).alias("f"), col("a").cast(T_I64).alias("l")])
    with pytest.raises(IglooCudaError) as e:
        gpu_filter_project(ctx, t, None, [(col("a") / lit(0, T_I32)).alias("q")])
    assert e.value.code == -6
    with pytest.raises(IglooCudaError) as e:
        gpu_filter_project(ctx, t, None, [(col("a") % lit(0, T_I32)).alias("q")])
    assert e.value.code == -6
    with pytest.raises(IglooCudaError) as e:
        gpu_filter_project(ctx, t, None, [(col("a") / lit(-1, T_I32)).alias("q")])
    assert e.value.code == -8
    big = pa.table({"c": np.array([1, 2**40], np.int64)})
    with pytest.raises(IglooCudaError) as e:
        gpu_filter_project(ctx, big, None, [col("c").cast(T_I32).alias("i")])
    assert e.value.code == -9
    # a null divisor row is not evaluated
    tn = pa.table({"a": pa.array([1, 2], pa.int32()), "b": pa.array([None, 1], pa.int32())})
    check(ctx, tn, None, [(col("a") / col("b")).alias("q")])


def test_projection_sees_only_rows_the_filter_kept(ctx):
    """Projection(Filter(x)): projection.rs evaluates its expressions on the batches filter.rs
    already compacted, so `a / b WHERE b <> 0` succeeds although b holds zeros."""
    from igloo_b200._lib import IglooCudaError
    rng = np.random.default_rng(11)
    n = 70_001
    t = pa.table({"a": rng.integers(-1000, 1000, n).astype(np.int32), "b": rng.integers(-3, 4, n).astype(np.int32),
                  "c": rng.integers(-2**40, 2**40, n).astype(np.int64)})
    check(ctx, t, col("b") != lit(0, T_I32), [(col("a") / col("b")).alias("q"), (col("a") % col("b")).alias("r")])
    check(ctx, t, (col("c") < lit(2**31, T_I64)) & (col("c") >= lit(-2**31, T_I64)), [col("c").cast(T_I32).alias("i")])
    for two_pass in (False, True):   # ... and still fails when a kept row divides by zero
        with pytest.raises(IglooCudaError) as e:
            gpu_filter_project(ctx, t, col("b") >= lit(0, T_I32), [(col("a") / col("b")).alias("q")], two_pass)
        assert e.value.code == -6


def test_run_time_specialised_kernels_match_the_interpreter_and_the_oracle(ctx, monkeypatch):
    """IGL_FP_JIT=1: every pass is compiled for the plan with NVRTC (fp_jit.cpp).  Same cases as the
    interpreter gets — fast handlers, generic null-aware steps, temporaries, division errors."""
    from igloo_b200._lib import IglooCudaError
    monkeypatch.setenv("IGL_FP_JIT", "1")
    t = c2_table(100_003)
    want, _ = oracle_filter_project(t, col("a") < 250, [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")])
    for two_pass in (False, True):
        got = gpu_filter_project(ctx, t, col("a") < 250, [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")],
                                 two_pass, expect_specialised=True)
        _check_one(got, want, True)
    rng = np.random.default_rng(17)
    n = 40_000
    tn = pa.table({"a": pa.array(rng.integers(-50, 50, n), pa.int32(), mask=rng.random(n) < 0.2),
                   "b": pa.array(rng.random(n) < 0.5, pa.bool_(), mask=rng.random(n) < 0.3),
                   "c": pa.array(rng.integers(-2**40, 2**40, n), pa.int64(), mask=rng.random(n) < 0.1),
                   "x": pa.array(rng.random(n) * 100 - 50)})
    check(ctx, tn, (col("a") > 0) | col("b"), [col("a").is_null().alias("n"), (col("a").cast(T_I64) + col("c")).alias("s"),
                                               ((col("x") + 1.0) * (col("x") - 1.0)).alias("p"), (col("c") % lit(7, T_I64)).alias("m")])
    check(ctx, tn, None, [(-col("a")).alias("neg"), (col("x") / lit(0.0)).alias("q"), col("x").cast(T_I32).alias("i")])
    with pytest.raises(IglooCudaError) as e:
        gpu_filter_project(ctx, tn, None, [(col("a") / lit(0, T_I32)).alias("q")])
    assert e.value.code == -6
    monkeypatch.setenv("IGL_FP_JIT", "0")
    gpu_filter_project(ctx, t, col("a") < 250, [col("b")], True, expect_specialised=False)
    # by default a plan specialises itself once it has been given enough rows in total: a stream of chunks counts
    monkeypatch.delenv("IGL_FP_JIT")
    monkeypatch.setenv("IGL_FP_JIT_MIN_ROWS", "250000")
    from igloo_b200.device import DeviceColumn, FilterProject
    cols = [DeviceColumn.from_arrow(t.column(i), ctx.device) for i in range(t.num_columns)]
    fp = FilterProject(ctx, _fields(t), col("a") < 250, [(col("b") + col("a")).alias("s")])
    seen = []
    for _ in range(4):
        out, n = fp.run(cols, t.num_rows)
        seen.append((fp.is_specialised, n, out[0].data[:n].cpu().numpy().copy()))
    assert [x[0] for x in seen] == [False, False, True, True]       # 100 003 rows per call, threshold 250 000
    assert all(x[1] == seen[0][1] and np.array_equal(x[2], seen[0][2]) for x in seen)
    # igl_fp_prepare: the same specialisation ahead of time, before the plan has seen a single row
    fp2 = FilterProject(ctx, _fields(t), col("a") < 250, [(col("b") + col("a")).alias("s")])
    assert not fp2.is_specialised
    assert fp2.prepare() and fp2.is_specialised
    out, n = fp2.run(cols, t.num_rows)
    assert n == seen[0][1] and np.array_equal(out[0].data[:n].cpu().numpy(), seen[0][2])


def test_predicate_must_be_boolean(ctx):
    from igloo_b200._lib import IglooCudaError
    from igloo_b200.device import FilterProject
    with pytest.raises(IglooCudaError) as e:
        FilterProject(ctx, [Field("a", T_I32)], col("a") + lit(1, T_I32), [])
    assert e.value.code == -7 and "boolean" in e.value.message


def test_float_total_order(ctx):
    nan = float("nan")
    t = pa.table({"x": np.array([0.0, -0.0, nan, 1.0, float("inf"), -nan, -1.0])})
    for e in (col("x") == lit(nan), col("x") < lit(0.0), col("x") > lit(float("inf")), col("x") >= lit(-0.0), col("x") != lit(0.0)):
        check(ctx, t, e, [])


def test_deep_expression_and_many_columns(ctx):
    rng = np.random.default_rng(11)
    n = 20_000
    cols = {f"c{i}": rng.integers(-1000, 1000, n).astype(np.int64) for i in range(10)}
    t = pa.table(cols)
    e = ((col("c0") + col("c1")) * (col("c2") - col("c3"))) + ((col("c4") * col("c5")) - (col("c6") + col("c7")))
    check(ctx, t, (col("c8") > col("c9")) & ((col("c0") % lit(7)) != lit(0)), [e.alias("e")] + [col(f"c{i}") for i in range(10)])


def test_full_scale_properties(ctx):
    """Size-independent checks at 64 Mi rows (device-generated C2 data): the
    surviving count equals the exact count of a<T, sum(s) - sum(b|a<T) = sum(a|a<T)
    via the digest of recomputed columns, and idempotence of a second run."""
    from igloo_b200.device import DeviceColumn, FilterProject
    n = 1 << 26
    seed = 0x16100002
    a = ctx.gen(1, seed, 0, 0, n, 1000)
    b = ctx.gen(1, seed, 1, 0, n, 1 << 31)
    x = ctx.gen(3, seed, 2, 0, n)
    fields = [Field("a", T_I32), Field("b", T_I32), Field("x", T_F64)]
    fused = FilterProject(ctx, fields, col("a") < 250, [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")])
    out1, n1 = fused.run([a, b, x], n)
    # unfused: Filter then Projection must give the identical stream
    filt = FilterProject(ctx, fields, col("a") < 250, [])
    mid, nm = filt.run([a, b, x], n)
    proj = FilterProject(ctx, fields, None, [(col("b") + col("a")).alias("s"), (col("x") * 1.5 + 0.25).alias("y")])
    out2, n2 = proj.run(mid, nm)
    assert n1 == nm == n2
    assert abs(n1 / n - 0.25) < 1e-3
    assert ctx.digest(out1, n1) == ctx.digest(out2, n2)
    import torch
    assert torch.equal(out1[0].data[:n1], out2[0].data[:n2]) and torch.equal(out1[1].data[:n1], out2[1].data[:n2])
    # oracle on the first 1 Mi rows: the prefix of the output must match exactly (order preserved)
    m = 1 << 20
    ha = orc.gen_i32_uniform(seed, 0, 0, m, 1000); hb = orc.gen_i32_uniform(seed, 1, 0, m, 1 << 31); hx = orc.gen_f64_unit(seed, 2, 0, m)
    assert np.array_equal(a.data[:m].cpu().numpy(), ha) and np.array_equal(x.data[:m].cpu().numpy(), hx)
    keep = ha < 250
    k = int(keep.sum())
    assert np.array_equal(out1[0].data[:k].cpu().numpy(), (hb[keep].astype(np.int64) + ha[keep]).astype(np.int32))
    assert np.array_equal(out1[1].data[:k].cpu().numpy(), hx[keep] * 1.5 + 0.25)
