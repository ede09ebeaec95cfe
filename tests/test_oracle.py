"""CPU tests: the oracle against the reference's golden vectors and against
known-answer cases authored from the cited reference lines."""
import json
import os

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc
import pytest

from igloo_b200.expr import Field, col, lit, lower, T_BOOL, T_I32, T_I64, T_F64, T_UTF8
from oracle import oracle as orc
from tests.util import arrow_to_columns, assert_same_multiset, columns_to_arrow

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "users.json")))


def users_table():
    r = GOLD["rows"]
    return pa.table({"id": pa.array(r["id"], pa.int32()), "name": pa.array(r["name"], pa.string()),
                     "age": pa.array(r["age"], pa.int32())},
                    schema=pa.schema([pa.field("id", pa.int32(), False), pa.field("name", pa.string(), False),
                                      pa.field("age", pa.int32(), False)]))


USERS_SCHEMA = [Field("id", T_I32), Field("name", T_UTF8), Field("age", T_I32)]


@pytest.mark.parametrize("case", GOLD["cases"], ids=lambda c: c["sql"])
def test_golden_users_filter_project(case):
    """integration_test.rs:21-29,59,72-75 — the reference's only known answer."""
    t = orc.Table.scan(arrow_to_columns(users_table()))
    pred, ty, _ = lower(col("age") > case["threshold"], USERS_SCHEMA)
    assert ty == T_BOOL
    f = t.filter(pred)
    fs = [Field(c.name, c.type) for c in f.columns()]
    exprs = [lower(col("name"), fs)[0], lower(col("age"), fs)[0]]
    p = f.project(exprs, ["name", "age"])
    got = columns_to_arrow(p.columns())
    rows = sorted(zip(got.column("name").to_pylist(), got.column("age").to_pylist()), key=lambda x: x[1])
    assert [list(r) for r in rows] == case["expect_sorted_by_age"]


def test_golden_literal_projection():
    """SELECT 42 AS answer => one row (lib.rs:160-183): literal broadcast."""
    one = orc.Table.scan([orc.Column("dummy", orc.T_I32, np.zeros(1, np.int32))])
    p = one.project([lower(lit(42), [])[0]], ["answer"])
    c = p.columns()[0]
    assert c.type == orc.T_I64 and c.to_pylist() == [GOLD["literal"]["value"]] and p.nrows == GOLD["literal"]["rows"]


def test_age_gt_25_lowering_is_int32():
    """`age > 25` reaches the operator as `age > Int32(25)`."""
    bc, _, _ = lower(col("age") > 25, USERS_SCHEMA)
    assert bc == bytes([0x01, 2, 0, 0x03, 25, 0, 0, 0, 0x14])


def test_filter_drops_empty_batches_projection_never():
    """filter.rs:59-63 drops 0-row batches; projection.rs:67 always emits."""
    n = 4096
    a = np.arange(n, dtype=np.int32)
    t = orc.Table.scan([orc.Column("a", orc.T_I32, a)], batch_rows=1024)
    sch = [Field("a", T_I32)]
    f = t.filter(lower(col("a") < 1500, sch)[0])
    assert f.nrows == 1500 and f.batch_sizes() == [1024, 476]
    p = f.project([lower(col("a") + 1, sch)[0]], ["b"])
    assert p.batch_sizes() == [1024, 476]
    none = t.filter(lower(col("a") < 0, sch)[0])
    assert none.nrows == 0 and none.nbatches == 0
    assert none.project([lower(col("a"), sch)[0]], ["a"]).nbatches == 0


def test_filter_null_mask_rows_dropped_and_kleene():
    a = pa.array([1, None, 3, None, 5, 6], pa.int32())
    b = pa.array([True, True, None, None, False, True], pa.bool_())
    t = orc.Table.scan(arrow_to_columns(pa.table({"a": a, "b": b})))
    sch = [Field("a", T_I32, True), Field("b", T_BOOL, True)]
    # a > 2: null rows dropped
    assert columns_to_arrow(t.filter(lower(col("a") > 2, sch)[0]).columns()).column("a").to_pylist() == [3, 5, 6]
    # (a > 2) OR b  — Kleene: null OR true = true, null OR null = null (dropped)
    got = columns_to_arrow(t.filter(lower((col("a") > 2) | col("b"), sch)[0]).columns()).column("a").to_pylist()
    assert got == [1, None, 3, 5, 6]
    # (a > 2) AND b — null AND false = false, true AND null = null
    got = columns_to_arrow(t.filter(lower((col("a") > 2) & col("b"), sch)[0]).columns()).column("a").to_pylist()
    assert got == [6]
    # IS NULL never yields null
    got = columns_to_arrow(t.filter(lower(col("a").is_null(), sch)[0]).columns()).column("b").to_pylist()
    assert got == [True, None]


def test_wrapping_and_errors():
    a = np.array([2**31 - 1, -2**31, 7], np.int32)
    t = orc.Table.scan([orc.Column("a", orc.T_I32, a)])
    sch = [Field("a", T_I32)]
    p = t.project([lower(col("a") + lit(1, T_I32), sch)[0], lower(col("a") * lit(2, T_I32), sch)[0],
                   lower(-col("a"), sch)[0]], ["x", "y", "z"])
    x, y, z = (c.to_pylist() for c in p.columns())
    assert x == [-2**31, -2**31 + 1, 8] and y == [-2, 0, 14] and z == [-(2**31 - 1), -2**31, -7]
    with pytest.raises(orc.OracleError, match="Divide by zero"):
        t.project([lower(col("a") / lit(0, T_I32), sch)[0]], ["q"])
    with pytest.raises(orc.OracleError, match="Divide by zero"):
        t.project([lower(col("a") % lit(0, T_I32), sch)[0]], ["q"])
    with pytest.raises(orc.OracleError, match="overflow"):
        t.project([lower(col("a") / lit(-1, T_I32), sch)[0]], ["q"])
    r = t.project([lower(col("a") % lit(-1, T_I32), sch)[0], lower(col("a") / lit(2, T_I32), sch)[0],
                   lower(col("a") % lit(4, T_I32), sch)[0]], ["m", "d", "r"])
    m, d, rr = (c.to_pylist() for c in r.columns())
    assert m == [0, 0, 0] and d == [1073741823, -1073741824, 3] and rr == [3, 0, 3]
    # a null divisor row is not evaluated (arrow visits valid slots only)
    tn = orc.Table.scan(arrow_to_columns(pa.table({"a": pa.array([1, 2], pa.int32()), "b": pa.array([None, 1], pa.int32())})))
    q = tn.project([lower(col("a") / col("b"), [Field("a", T_I32), Field("b", T_I32, True)])[0]], ["q"])
    assert q.columns()[0].to_pylist() == [None, 2]


def test_float_total_order_and_no_fma():
    nan = float("nan")
    x = np.array([0.0, -0.0, nan, 1.0, float("inf")], np.float64)
    t = orc.Table.scan([orc.Column("x", orc.T_F64, x)])
    sch = [Field("x", T_F64)]
    f = lambda e: columns_to_arrow(t.filter(lower(e, sch)[0]).columns()).column("x").to_pylist()
    assert len(f(col("x") == lit(nan))) == 1                      # NaN == NaN
    assert [repr(v) for v in f(col("x") < lit(0.0))] == ["-0.0"]  # -0.0 < +0.0
    assert len(f(col("x") > lit(float("inf")))) == 1               # NaN sorts above +inf
    # x*1.5+0.25 must be two roundings (no contraction)
    rng = np.random.default_rng(7)
    xs = rng.random(20000)
    p = orc.Table.scan([orc.Column("x", orc.T_F64, xs)]).project([lower(col("x") * 1.5 + 0.25, sch)[0]], ["y"])
    assert np.array_equal(p.columns()[0].values, (xs * 1.5) + 0.25)
    # f64 / 0 is not an error
    d = t.project([lower(col("x") / lit(0.0), sch)[0]], ["d"]).columns()[0].values
    assert np.isnan(d[0]) and np.isinf(d[3])


def test_cast():
    a = np.array([1, -5, 2**40], np.int64)
    t = orc.Table.scan([orc.Column("a", orc.T_I64, a)])
    sch = [Field("a", T_I64)]
    assert t.project([lower(col("a").cast(T_F64), sch)[0]], ["f"]).columns()[0].to_pylist() == [1.0, -5.0, float(2**40)]
    with pytest.raises(orc.OracleError, match="Cast"):
        t.project([lower(col("a").cast(T_I32), sch)[0]], ["i"])
    x = np.array([1.9, -1.9, 3.0], np.float64)
    tx = orc.Table.scan([orc.Column("x", orc.T_F64, x)])
    assert tx.project([lower(col("x").cast(T_I32), [Field("x", T_F64)])[0]], ["i"]).columns()[0].to_pylist() == [1, -1, 3]


def test_pyarrow_second_opinion_filter_project():
    """Arrow C++ agrees on data away from NaN / -0.0 (last row)."""
    n = 50000
    a = orc.gen_i32_uniform(0x16100002, 0, 0, n, 1000)
    b = orc.gen_i32_uniform(0x16100002, 1, 0, n, 1 << 31)
    x = orc.gen_f64_unit(0x16100002, 2, 0, n)
    sch = [Field("a", T_I32), Field("b", T_I32), Field("x", T_F64)]
    t = orc.Table.scan([orc.Column("a", orc.T_I32, a), orc.Column("b", orc.T_I32, b), orc.Column("x", orc.T_F64, x)])
    f = t.filter(lower(col("a") < 250, sch)[0])
    p = f.project([lower(col("b") + col("a"), sch)[0], lower(col("x") * 1.5 + 0.25, sch)[0]], ["s", "y"])
    m = a < 250
    assert np.array_equal(p.columns()[0].values, (b[m].astype(np.int64) + a[m]).astype(np.int32))
    assert np.array_equal(p.columns()[1].values, x[m] * 1.5 + 0.25)
    tb = pa.table({"a": a, "b": b, "x": x})
    ft = tb.filter(pc.less(tb["a"], 250))
    assert np.array_equal(pc.add(ft["b"], ft["a"]).to_numpy(), p.columns()[0].values)


# ---------------------------------------------------------------------------
# Hash join: known answers authored from hash_join.rs (parity unpinned upstream)
# ---------------------------------------------------------------------------
def _join(left, right, lk, rk, jt=orc.JOIN_INNER, mode=0, threads=1):
    L = orc.Table.scan(arrow_to_columns(left))
    R = orc.Table.scan(arrow_to_columns(right))
    ls = [Field(f.name, {pa.int32(): T_I32, pa.int64(): T_I64, pa.float64(): T_F64, pa.string(): T_UTF8, pa.bool_(): T_BOOL}[f.type], f.nullable) for f in left.schema]
    rs = [Field(f.name, {pa.int32(): T_I32, pa.int64(): T_I64, pa.float64(): T_F64, pa.string(): T_UTF8, pa.bool_(): T_BOOL}[f.type], f.nullable) for f in right.schema]
    J = L.hash_join(R, [lower(col(lk), ls)[0]], [lower(col(rk), rs)[0]], jt, mode, threads)
    return J, columns_to_arrow(J.columns())


@pytest.mark.parametrize("mode", [0, 1])
def test_join_duplicates_order_and_rename(mode):
    left = pa.table({"k": pa.array([1, 2, 3, 2], pa.int64()), "w": pa.array([10, 20, 30, 40], pa.int64())})
    right = pa.table({"k": pa.array([2, 9, 2, 1], pa.int64()), "v": pa.array([200, 900, 201, 100], pa.int64())})
    J, out = _join(left, right, "k", "k", mode=mode)
    # schema: left fields then right, clash renamed right_k (hash_join.rs:54-63)
    assert out.schema.names == ["k", "w", "right_k", "v"]
    # probe order, then build arrival order (hash_join.rs:125-126,176-181)
    assert out.column("w").to_pylist() == [10, 20, 20, 40, 40]
    assert out.column("v").to_pylist() == [100, 200, 201, 200, 201]
    assert J.nbatches == J.nrows == 5  # one 1-row batch per emit


@pytest.mark.parametrize("mode", [0, 1])
def test_join_left_null_fill_right_is_inner_full_is_left(mode):
    left = pa.table({"k": pa.array([1, 5], pa.int64()), "w": pa.array([10, 50], pa.int64())})
    right = pa.table({"k": pa.array([1, 7], pa.int64()), "v": pa.array([100, 700], pa.int64())})
    _, inner = _join(left, right, "k", "k", orc.JOIN_INNER, mode)
    _, lft = _join(left, right, "k", "k", orc.JOIN_LEFT, mode)
    _, rgt = _join(left, right, "k", "k", orc.JOIN_RIGHT, mode)
    _, full = _join(left, right, "k", "k", orc.JOIN_FULL, mode)
    assert inner.to_pydict() == {"k": [1], "w": [10], "right_k": [1], "v": [100]}
    assert lft.to_pydict() == {"k": [1, 5], "w": [10, 50], "right_k": [1, None], "v": [100, None]}
    assert rgt.to_pydict() == inner.to_pydict()      # Right never emits unmatched right rows
    assert full.to_pydict() == lft.to_pydict()       # Full == Left (hash_join.rs:189, :263-276 unreachable)


@pytest.mark.parametrize("mode", [0, 1])
def test_join_null_keys_match_and_float_identity(mode):
    left = pa.table({"k": pa.array([None, 1], pa.int64()), "w": pa.array([1, 2], pa.int64())})
    right = pa.table({"k": pa.array([None, None, 1], pa.int64()), "v": pa.array([7, 8, 9], pa.int64())})
    _, out = _join(left, right, "k", "k", mode=mode)
    assert out.column("v").to_pylist() == [7, 8, 9]   # NULL == NULL via the printed form "null"
    nan = float("nan")
    lf = pa.table({"k": pa.array([nan, 0.0, -0.0], pa.float64()), "w": pa.array([1, 2, 3], pa.int64())})
    rf = pa.table({"k": pa.array([nan, -0.0], pa.float64()), "v": pa.array([10, 20], pa.int64())})
    _, of = _join(lf, rf, "k", "k", mode=mode)
    assert of.column("w").to_pylist() == [1, 3] and of.column("v").to_pylist() == [10, 20]


@pytest.mark.parametrize("mode", [0, 1])
def test_join_type_name_is_part_of_key(mode):
    left = pa.table({"k": pa.array([1, 2], pa.int32())})
    right = pa.table({"j": pa.array([1, 2], pa.int64())})
    _, out = _join(left, right, "k", "j", mode=mode)
    assert out.num_rows == 0                            # "PrimitiveArray<Int32>" != "PrimitiveArray<Int64>"
    _, lo = _join(left, right, "k", "j", orc.JOIN_LEFT, mode)
    assert lo.to_pydict() == {"k": [1, 2], "j": [None, None]}


def test_join_utf8_key_and_payload_faithful():
    left = pa.table({"name": pa.array(["Bob", "Eve", "Zed"]), "age": pa.array([30, 32, 1], pa.int32())})
    right = pa.table({"name": pa.array(["Eve", "Bob"]), "city": pa.array(["Oslo", "Rome"])})
    _, out = _join(left, right, "name", "name")
    assert out.to_pydict() == {"name": ["Bob", "Eve"], "age": [30, 32], "right_name": ["Bob", "Eve"], "city": ["Rome", "Oslo"]}


def test_join_utf8_key_null_the_string_null_and_the_empty_string():
    """The printed key (hash_join.rs:120, arrow-array's Debug) writes a NULL as `null` and the string "null" as
    `"null"` with its quotes, so they differ; NULL == NULL like every key; "" is a value of its own; and a Utf8 key
    never meets a key of another type (`StringArray` is in the printed form).  What the GPU string-key tests
    (tests/test_gpu_operators.py) are held against."""
    left = pa.table({"name": pa.array(["null", None, "", "a", "a"], pa.string()), "w": pa.array([1, 2, 3, 4, 5], pa.int32())})
    right = pa.table({"name": pa.array([None, "null", "a", "", None], pa.string()), "v": pa.array([10, 20, 30, 40, 50], pa.int32())})
    _, out = _join(left, right, "name", "name")
    got = sorted(zip(out.column("w").to_pylist(), out.column("v").to_pylist()))
    assert got == [(1, 20), (2, 10), (2, 50), (3, 40), (4, 30), (5, 30)]
    _, lo = _join(pa.table({"name": pa.array(["zz"], pa.string()), "w": pa.array([9], pa.int32())}), right, "name", "name", orc.JOIN_LEFT)
    assert lo.to_pydict() == {"name": ["zz"], "w": [9], "right_name": [None], "v": [None]}
    ints = pa.table({"name": pa.array([1, 2], pa.int64()), "v": pa.array([1, 2], pa.int32())})
    _, none = _join(pa.table({"name": pa.array(["1", "2"], pa.string())}), ints, "name", "name")
    assert none.num_rows == 0


def test_join_fast_equals_faithful_random():
    rng = np.random.default_rng(3)
    nl, nr = 5000, 700
    lk = rng.integers(0, 400, nl)
    rk = rng.integers(0, 400, nr)
    left = pa.table({"k": pa.array(lk, pa.int64(), mask=rng.random(nl) < 0.02), "w": pa.array(rng.integers(0, 1 << 40, nl), pa.int64())})
    right = pa.table({"k": pa.array(rk, pa.int64(), mask=rng.random(nr) < 0.02), "v": pa.array(rng.random(nr), pa.float64())})
    for jt in (orc.JOIN_INNER, orc.JOIN_LEFT):
        J0, a = _join(left, right, "k", "k", jt, 0)
        J1, b = _join(left, right, "k", "k", jt, 1, threads=3)
        assert a.equals(b)                       # same rows in the same order
        assert J0.digest() == J1.digest()
        assert_same_multiset(a, b)


def test_generator_and_permutation():
    bits = 12
    p = orc.gen_perm_pow2(0, 1 << bits, bits)
    assert sorted(p.tolist()) == list(range(1 << bits))
    for x in (0, 1, 77, 4095):
        assert orc.lib().orc_perm_pow2_inv(orc.lib().orc_perm_pow2(x, bits), bits) == x
    for b in (27, 28):
        for x in (0, 12345, (1 << b) - 1):
            assert orc.lib().orc_perm_pow2_inv(orc.lib().orc_perm_pow2(x, b), b) == x
    a = orc.gen_i32_uniform(1, 0, 0, 100000, 1000)
    assert a.min() >= 0 and a.max() < 1000 and abs(a.mean() - 499.5) < 5
    assert np.array_equal(orc.gen_i32_uniform(1, 0, 500, 100, 1000), a[500:600])   # range addressable
    x = orc.gen_f64_unit(1, 2, 0, 1000)
    assert x.min() >= 0 and x.max() < 1
    # splitmix64 known answer (first output of the reference SplitMix64 seeded with 0)
    assert orc.lib().orc_splitmix64(0) == 0xE220A8397B1DCDAF


@pytest.mark.parametrize("mode", [0, 1])
def test_pyarrow_second_opinion_join(mode):
    """Acero's hash join (Arrow C++) agrees with the restated join on data away from the documented
    divergences (no NULL / NaN / -0.0 keys): Inner and Left, duplicates on both sides, as multisets."""
    rng = np.random.default_rng(21)
    nl, nb = 5000, 1200
    left = pa.table({"k": pa.array(rng.integers(0, 700, nl), pa.int64()), "w": pa.array(rng.integers(0, 1 << 40, nl), pa.int64())})
    right = pa.table({"k": pa.array(rng.integers(0, 600, nb), pa.int64()), "v": pa.array(rng.random(nb)),
                      "z": pa.array(rng.integers(0, 99, nb), pa.int32(), mask=rng.random(nb) < 0.2)})
    for jt, acero in ((orc.JOIN_INNER, "inner"), (orc.JOIN_LEFT, "left outer")):
        _, got = _join(left, right, "k", "k", jt, mode)
        want = left.join(right.rename_columns(["rk", "v", "z"]), keys="k", right_keys="rk", join_type=acero, coalesce_keys=False)
        assert got.num_rows == want.num_rows
        g = sorted(zip(*[got.column(c).to_pylist() for c in ("k", "w", "right_k", "v", "z")]), key=repr)
        w = sorted(zip(*[want.column(c).to_pylist() for c in ("k", "w", "rk", "v", "z")]), key=repr)
        assert g == w


def test_golden_scan_fixture_csv():
    """The reference's tiny scan fixture (connectors/filesystem/test_data.csv), read the way the harness
    reads CSV (pyarrow), goes through the restated Filter: col_a > 1 keeps (2, bar)."""
    import io
    import json
    import os
    import pyarrow.csv as pacsv
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "scan_fixture.json")))
    tbl = pacsv.read_csv(io.BytesIO(g["csv"].encode()))
    assert tbl.to_pydict() == g["rows"]
    sch = [Field("col_a", T_I64), Field("col_b", T_UTF8)]
    t = orc.Table.scan(arrow_to_columns(tbl))
    out = columns_to_arrow(t.filter(lower(col("col_a") > lit(1, T_I64), sch)[0]).columns())
    assert out.to_pydict() == {"col_a": [2], "col_b": ["bar"]}
