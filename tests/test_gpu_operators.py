"""GPU parity through the streaming operator API (Arrow C Data Interface push/pull),
written like the reference's own integration test (crates/engine/tests/integration_test.rs)."""
import json
import os

import numpy as np
import pyarrow as pa
import pytest

from igloo_b200.expr import Field, col, lit, lower, T_BOOL, T_F64, T_I32, T_I64
from oracle import oracle as orc
from tests.util import arrow_to_columns, assert_same_multiset, columns_to_arrow

pytestmark = pytest.mark.gpu
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "users.json")))


@pytest.fixture(scope="module")
def ctx():
    from igloo_b200.device import Context
    c = Context(0, coalesce_rows=4096)   # small chunks: every test crosses chunk boundaries
    yield c
    c.close()


def users():
    r = GOLD["rows"]
    return pa.table({"id": pa.array(r["id"], pa.int32()), "name": pa.array(r["name"], pa.string()), "age": pa.array(r["age"], pa.int32())},
                    schema=pa.schema([pa.field("id", pa.int32(), False), pa.field("name", pa.string(), False), pa.field("age", pa.int32(), False)]))


def collect(plan):
    batches = list(plan.execute())
    return pa.Table.from_batches(batches, schema=plan.schema()), batches


def test_c1_select_name_age_where_age_gt(ctx):
    """BASELINE.json configs[0] on the integration_test.rs rows, via the planner:
    SELECT name, age FROM users WHERE age > 25 -> 4 rows; WHERE age > 30 -> (Eve,32),(Charlie,35)."""
    from igloo_b200.physical_planner import Filter, MemoryCatalog, PhysicalPlanner, Projection, TableScan
    cat = MemoryCatalog()
    cat.register_table("users", users())
    for fuse in (True, False):
        planner = PhysicalPlanner(cat, ctx, fuse=fuse)
        for case in GOLD["cases"]:
            plan = planner.create_physical_plan(Projection(Filter(TableScan("users", [1, 2]), col("age") > case["threshold"]),
                                                           [col("name"), col("age")]))
            t, _ = collect(plan)
            assert t.schema.names == ["name", "age"] and t.schema.field("age").type == pa.int32()
            rows = sorted(zip(t.column("name").to_pylist(), t.column("age").to_pylist()), key=lambda x: x[1])
            assert [list(r) for r in rows] == case["expect_sorted_by_age"]


def test_filter_exec_streams_and_drops_empty_chunks(ctx):
    from igloo_b200.operators import FilterExec, MemoryScanExec
    n = 20_000
    rng = np.random.default_rng(0)
    tbl = pa.table({"a": np.arange(n, dtype=np.int32), "s": pa.array([f"row{i % 97}" * (i % 3) for i in range(n)]),
                    "x": pa.array(rng.random(n), pa.float64(), mask=rng.random(n) < 0.1)},
                   schema=pa.schema([pa.field("a", pa.int32(), False), pa.field("s", pa.string(), False), pa.field("x", pa.float64(), True)]))
    f = FilterExec(MemoryScanExec(tbl), (col("a") < 6000) | (col("a") >= 19_000), ctx)
    got, batches = collect(f)
    keep = (np.arange(n) < 6000) | (np.arange(n) >= 19_000)
    assert got.schema == tbl.schema
    assert got.equals(tbl.filter(pa.array(keep)))             # order preserved, strings gathered, nulls kept
    assert all(b.num_rows > 0 for b in batches) and len(batches) == 3   # chunks 0,1 and 4 survive; 2,3 are dropped
    none, nb = collect(FilterExec(MemoryScanExec(tbl), col("a") < 0, ctx))
    assert none.num_rows == 0 and nb == []
    # oracle agreement on the numeric + string columns
    o = orc.Table.scan(arrow_to_columns(tbl))
    fields = [Field("a", T_I32), Field("s", 5), Field("x", T_F64, True)]
    want = columns_to_arrow(o.filter(lower((col("a") < 6000) | (col("a") >= 19_000), fields)[0]).columns())
    assert_same_multiset(got.cast(want.schema), want)


def test_projection_exec_matches_oracle_and_pair_equals_fused(ctx):
    from igloo_b200.operators import FilterExec, FusedFilterProjectExec, MemoryScanExec, ProjectionExec
    n = 50_000
    a = orc.gen_i32_uniform(0x16100002, 0, 0, n, 1000); b = orc.gen_i32_uniform(0x16100002, 1, 0, n, 1 << 31); x = orc.gen_f64_unit(0x16100002, 2, 0, n)
    sch = pa.schema([pa.field("a", pa.int32(), False), pa.field("b", pa.int32(), False), pa.field("x", pa.float64(), False)])
    tbl = pa.table({"a": a, "b": b, "x": x}, schema=sch)
    exprs = [((col("b") + col("a")).alias("s"), "s"), ((col("x") * 1.5 + 0.25).alias("y"), "y")]
    out_schema = pa.schema([pa.field("s", pa.int32(), False), pa.field("y", pa.float64(), False)])
    pair = ProjectionExec(FilterExec(MemoryScanExec(tbl), col("a") < 250, ctx), exprs, out_schema, ctx)
    fused = FusedFilterProjectExec(MemoryScanExec(tbl), col("a") < 250, exprs, out_schema, ctx)
    t1, _ = collect(pair)
    t2, _ = collect(fused)
    assert t1.equals(t2)
    m = a < 250
    assert np.array_equal(t1.column("s").to_numpy(), (b[m].astype(np.int64) + a[m]).astype(np.int32))
    assert np.array_equal(t1.column("y").to_numpy(), x[m] * 1.5 + 0.25)
    # SELECT 42 AS answer (lib.rs:160-183): literal projection over a one-row relation
    one = pa.table({"d": pa.array([0], pa.int32())})
    t, _ = collect(ProjectionExec(MemoryScanExec(one), [(lit(42).alias("answer"), "answer")], pa.schema([pa.field("answer", pa.int64(), False)]), ctx))
    assert t.column("answer").to_pylist() == [42]


def test_stream_error_is_surfaced(ctx):
    from igloo_b200._lib import IglooCudaError
    from igloo_b200.operators import MemoryScanExec, ProjectionExec
    tbl = pa.table({"a": pa.array([1, 2, 0, 4], pa.int32())}, schema=pa.schema([pa.field("a", pa.int32(), False)]))
    p = ProjectionExec(MemoryScanExec(tbl), [((lit(10, T_I32) / col("a")).alias("q"), "q")], pa.schema([pa.field("q", pa.int32(), False)]), ctx)
    with pytest.raises(IglooCudaError) as e:
        collect(p)
    assert e.value.code == -6 and "Divide by zero" in e.value.message


@pytest.mark.parametrize("jt", [0, 1])
def test_hash_join_exec_streaming(ctx, jt):
    from igloo_b200.operators import HashJoinExec, MemoryScanExec
    rng = np.random.default_rng(5)
    nb, nl = 9_000, 30_000
    right = pa.table({"k": rng.permutation(20_000)[:nb].astype(np.int64), "v": rng.integers(0, 1 << 40, nb)},
                     schema=pa.schema([pa.field("k", pa.int64(), False), pa.field("v", pa.int64(), False)]))
    left = pa.table({"k": rng.integers(0, 20_000, nl), "w": rng.random(nl)},
                    schema=pa.schema([pa.field("k", pa.int64(), False), pa.field("w", pa.float64(), False)]))
    j = HashJoinExec(MemoryScanExec(left), MemoryScanExec(right), [col("k")], [col("k")], jt, ctx)
    assert j.schema().names == ["k", "w", "right_k", "v"]
    got, _ = collect(j)
    if jt == 0:   # inner join on non-null int64 keys: the two key columns of a batch share one host buffer
        assert got.column("right_k").equals(got.column("k"))
    L = orc.Table.scan(arrow_to_columns(left)); R = orc.Table.scan(arrow_to_columns(right))
    key = lower(col("k"), [Field("k", T_I64)])[0]
    want = columns_to_arrow(L.hash_join(R, [key], [key], jt, 1, 2).columns())
    assert_same_multiset(got.cast(pa.schema([pa.field(f.name, f.type, True) for f in got.schema])),
                         want.cast(pa.schema([pa.field(f.name, f.type, True) for f in want.schema])))


def test_hash_join_duplicates_and_nullable_columns_streaming(ctx):
    from igloo_b200.operators import HashJoinExec, MemoryScanExec
    rng = np.random.default_rng(6)
    nb, nl = 8_192, 12_000
    right = pa.table({"k": pa.array(rng.integers(0, 500, nb), pa.int64(), mask=rng.random(nb) < 0.01),
                      "v": pa.array(rng.integers(0, 99, nb), pa.int32(), mask=rng.random(nb) < 0.3)})
    left = pa.table({"k": pa.array(rng.integers(0, 600, nl), pa.int64(), mask=rng.random(nl) < 0.01), "w": pa.array(rng.random(nl))})
    for jt in (0, 1):
        got, _ = collect(HashJoinExec(MemoryScanExec(left), MemoryScanExec(right), [col("k")], [col("k")], jt, ctx))
        L = orc.Table.scan(arrow_to_columns(left)); R = orc.Table.scan(arrow_to_columns(right))
        key = lower(col("k"), [Field("k", T_I64, True)])[0]
        want = columns_to_arrow(L.hash_join(R, [key], [key], jt, 1, 2).columns())
        assert_same_multiset(got.cast(pa.schema([pa.field(f.name, f.type, True) for f in got.schema])),
                             want.cast(pa.schema([pa.field(f.name, f.type, True) for f in want.schema])))


def test_planner_join_of_filtered_scans(ctx):
    from igloo_b200.physical_planner import Filter, Join, MemoryCatalog, PhysicalPlanner, Projection, TableScan
    rng = np.random.default_rng(8)
    orders = pa.table({"o_orderkey": np.arange(1, 5001, dtype=np.int64), "o_total": rng.random(5000)})
    items = pa.table({"l_orderkey": rng.integers(1, 6000, 40_000), "l_qty": rng.integers(1, 50, 40_000).astype(np.int32)})
    cat = MemoryCatalog(); cat.register_table("orders", orders); cat.register_table("lineitem", items)
    plan = PhysicalPlanner(cat, ctx).create_physical_plan(
        Join(Filter(TableScan("lineitem"), col("l_qty") > 10), TableScan("orders"), [(col("l_orderkey"), col("o_orderkey"))], "Inner"))
    got, _ = collect(plan)
    m = (items.column("l_qty").to_numpy() > 10) & (items.column("l_orderkey").to_numpy() <= 5000)
    assert got.num_rows == int(m.sum()) and got.schema.names == ["l_orderkey", "l_qty", "o_orderkey", "o_total"]
    assert np.array_equal(np.sort(got.column("l_orderkey").to_numpy()), np.sort(items.column("l_orderkey").to_numpy()[m]))
    assert np.array_equal(got.column("l_orderkey").to_numpy(), got.column("o_orderkey").to_numpy())


def test_golden_scan_fixture_through_filter_exec(ctx):
    """tests/golden/scan_fixture.json (the reference's test_data.csv): CSV -> MemoryScanExec -> FilterExec on
    the GPU, the Utf8 column passing through by row id."""
    import io
    import pyarrow.csv as pacsv
    from igloo_b200.operators import FilterExec, MemoryScanExec
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "scan_fixture.json")))
    tbl = pacsv.read_csv(io.BytesIO(g["csv"].encode()))
    got, _ = collect(FilterExec(MemoryScanExec(tbl), col("col_a") > lit(1, T_I64), ctx))
    assert got.to_pydict() == {"col_a": [2], "col_b": ["bar"]}


def test_hash_join_exec_key_lists_computed_keys_and_boolean_columns(ctx):
    """HashJoinExec::new takes key VECTORS of arbitrary expressions (hash_join.rs:35-41, 109-122) and
    carries every column type of its children: two key pairs (one computed), a Boolean payload column
    on each side, streamed in small batches; Inner and Left against the oracle."""
    from igloo_b200.operators import HashJoinExec, MemoryScanExec
    rng = np.random.default_rng(12)
    nb, nl = 6_000, 25_000
    right = pa.table({"a": rng.integers(0, 80, nb).astype(np.int32), "b": rng.integers(0, 60, nb).astype(np.int32),
                      "ok": pa.array(rng.random(nb) < 0.4, pa.bool_()), "v": rng.integers(0, 1 << 40, nb)},
                     schema=pa.schema([pa.field("a", pa.int32(), False), pa.field("b", pa.int32(), False),
                                       pa.field("ok", pa.bool_(), False), pa.field("v", pa.int64(), False)]))
    left = pa.table({"a": rng.integers(0, 90, nl).astype(np.int32), "c": rng.integers(0, 30, nl).astype(np.int32),
                     "hot": pa.array(rng.random(nl) < 0.5, pa.bool_()), "w": rng.random(nl)},
                    schema=pa.schema([pa.field("a", pa.int32(), False), pa.field("c", pa.int32(), False),
                                      pa.field("hot", pa.bool_(), False), pa.field("w", pa.float64(), False)]))
    lf = [Field("a", T_I32), Field("c", T_I32), Field("hot", 1), Field("w", 4)]
    rf = [Field("a", T_I32), Field("b", T_I32), Field("ok", 1), Field("v", T_I64)]
    lkeys, rkeys = [col("a"), col("c") * 2], [col("a"), col("b")]
    for jt in (0, 1):
        j = HashJoinExec(MemoryScanExec(left, 4096), MemoryScanExec(right, 4096), lkeys, rkeys, jt, ctx)
        assert j.schema().names == ["a", "c", "hot", "w", "right_a", "b", "ok", "v"]
        got, _ = collect(j)
        L = orc.Table.scan(arrow_to_columns(left)); R = orc.Table.scan(arrow_to_columns(right))
        want = columns_to_arrow(L.hash_join(R, [lower(e, lf)[0] for e in lkeys], [lower(e, rf)[0] for e in rkeys], jt, 0, 1).columns())
        assert got.num_rows == want.num_rows > 0
        assert_same_multiset(got.cast(pa.schema([pa.field(f.name, f.type, True) for f in got.schema])),
                             want.cast(pa.schema([pa.field(f.name, f.type, True) for f in want.schema])))


def test_fragment_exchange_wire_format_carries_gpu_operator_output(ctx):
    """a Compute fragment's FilterExec output travels as RecordBatchMessage wire bytes
    (distributed.proto:53-57; the stubs at distributed_executor.rs:203-222) and feeds the probe side
    of a HashJoinExec in the dependent Join fragment."""
    from igloo_b200.exchange import (FragmentType, QueryFragment, RecordBatchMessage, WireScanExec, deserialize_batch,
                                     execute_fragment)
    from igloo_b200.operators import FilterExec, HashJoinExec, MemoryScanExec
    rng = np.random.default_rng(21)
    nb, nl = 3_000, 40_000
    orders = pa.table({"o_key": np.arange(nb, dtype=np.int64), "o_total": rng.random(nb)},
                      schema=pa.schema([pa.field("o_key", pa.int64(), False), pa.field("o_total", pa.float64(), False)]))
    items = pa.table({"l_key": rng.integers(0, 2 * nb, nl), "l_qty": rng.integers(1, 50, nl).astype(np.int32)},
                     schema=pa.schema([pa.field("l_key", pa.int64(), False), pa.field("l_qty", pa.int32(), False)]))
    compute = QueryFragment("f1", FragmentType.Compute, FilterExec(MemoryScanExec(items, 8192), col("l_qty") > 40, ctx), "worker-0")
    wire = execute_fragment(compute)
    assert wire and all(isinstance(w, bytes) for w in wire)
    assert sum(RecordBatchMessage.from_wire(w).num_rows for w in wire) == int((items.column("l_qty").to_numpy() > 40).sum())
    join = QueryFragment("f2", FragmentType.Join,
                         HashJoinExec(WireScanExec(wire, items.schema), MemoryScanExec(orders), [col("l_key")], [col("o_key")], 0, ctx),
                         "coordinator", dependencies=["f1"])
    got = pa.Table.from_batches([deserialize_batch(RecordBatchMessage.from_wire(w)) for w in execute_fragment(join)])
    m = (items.column("l_qty").to_numpy() > 40) & (items.column("l_key").to_numpy() < nb)
    assert got.num_rows == int(m.sum()) and got.schema.names == ["l_key", "l_qty", "o_key", "o_total"]
    assert np.array_equal(np.sort(got.column("l_key").to_numpy()), np.sort(items.column("l_key").to_numpy()[m]))
    assert np.array_equal(got.column("o_total").to_numpy(), orders.column("o_total").to_numpy()[got.column("o_key").to_numpy()])


@pytest.mark.parametrize("jt", [0, 1])
def test_hash_join_exec_carries_utf8_columns_on_both_sides(ctx, jt):
    """The reference's own fixture shape, users(id, name, age) (integration_test.rs:21-29), joined with
    an orders table that carries a nullable string as well: HashJoinExec stores and emits whole rows
    (hash_join.rs:125, 229-236), so string columns ride along on either side.  On the GPU the join sees
    row ids in their place and the strings are gathered afterwards."""
    from igloo_b200.operators import HashJoinExec, MemoryScanExec
    rng = np.random.default_rng(77)
    names = GOLD["rows"]["name"] if "rows" in GOLD and "name" in GOLD["rows"] else ["Alice", "Bob", "Charlie", "Diana", "Eve"]
    nu, no = 3_000, 20_000
    users = pa.table({"id": pa.array(np.arange(1, nu + 1), pa.int32()),
                      "name": pa.array([f"{names[i % len(names)]}-{i}" for i in range(nu)], pa.string()),
                      "age": pa.array(rng.integers(18, 80, nu), pa.int32())},
                     schema=pa.schema([pa.field("id", pa.int32(), False), pa.field("name", pa.string(), False), pa.field("age", pa.int32(), False)]))
    notes = [None if rng.random() < 0.2 else ("n" * int(rng.integers(0, 12))) + str(i) for i in range(no)]
    orders = pa.table({"user_id": pa.array(rng.integers(1, nu + 500, no), pa.int32()), "amount": rng.random(no),
                       "note": pa.array(notes, pa.string())})
    # orders probe, users build (strings on both sides); then the other way round
    for left, right, lk, rk in ((orders, users, "user_id", "id"), (users, orders, "id", "user_id")):
        j = HashJoinExec(MemoryScanExec(left, 2048), MemoryScanExec(right, 1000), [col(lk)], [col(rk)], jt, ctx)
        got, _ = collect(j)
        L = orc.Table.scan(arrow_to_columns(left)); R = orc.Table.scan(arrow_to_columns(right))
        from tests.util import tight_fields
        lf, rf = tight_fields(left), tight_fields(right)
        want = columns_to_arrow(L.hash_join(R, [lower(col(lk), lf)[0]], [lower(col(rk), rf)[0]], jt, 0, 1).columns())
        assert got.schema.names == want.schema.names and got.num_rows == want.num_rows > 0
        assert_same_multiset(got.cast(pa.schema([pa.field(f.name, f.type, True) for f in got.schema])),
                             want.cast(pa.schema([pa.field(f.name, f.type, True) for f in want.schema])))


def _string_key_tables(seed, nl, nr, nkeys_distinct):
    rng = np.random.default_rng(seed)
    def names(n):
        out = []
        for _ in range(n):
            u = rng.random()
            k = int(rng.integers(0, nkeys_distinct))
            out.append(None if u < 0.05 else "null" if u < 0.07 else "" if u < 0.09 else f"user-{k}" + "x" * (k % 7))
        return out
    left = pa.table({"name": pa.array(names(nl), pa.string()), "v": pa.array(rng.integers(0, 1000, nl), pa.int32()),
                     "grp": pa.array(rng.integers(0, 3, nl), pa.int64())})
    right = pa.table({"tag": pa.array(rng.integers(0, 3, nr), pa.int64()), "name": pa.array(names(nr), pa.string()), "w": rng.random(nr)})
    return left, right


def _check_join(ctx, left, right, lkeys, rkeys, jt):
    from igloo_b200.operators import HashJoinExec, MemoryScanExec
    from tests.util import tight_fields
    j = HashJoinExec(MemoryScanExec(left, 1500), MemoryScanExec(right, 700), [col(k) for k in lkeys], [col(k) for k in rkeys], jt, ctx)
    got, _ = collect(j)
    L = orc.Table.scan(arrow_to_columns(left)); R = orc.Table.scan(arrow_to_columns(right))
    lf, rf = tight_fields(left), tight_fields(right)
    want = columns_to_arrow(L.hash_join(R, [lower(col(k), lf)[0] for k in lkeys], [lower(col(k), rf)[0] for k in rkeys], jt, 0, 1).columns())
    assert got.schema.names == want.schema.names and got.num_rows == want.num_rows > 0
    assert_same_multiset(got.cast(pa.schema([pa.field(f.name, f.type, True) for f in got.schema])),
                         want.cast(pa.schema([pa.field(f.name, f.type, True) for f in want.schema])))
    return got.num_rows


@pytest.mark.parametrize("jt", [0, 1])
def test_hash_join_exec_on_a_string_key(ctx, jt):
    """hash_join.rs:117-122 joins on the printed key, so a Utf8 key is a key like any other: NULL == NULL,
    the string "null" is not NULL, the empty string is a value.  On the GPU: a 64-bit hash of the bytes as
    the device key, every joined pair compared byte for byte afterwards.  Unique and duplicated build keys."""
    left, right = _string_key_tables(5, 9_000, 2_500, 4_000)      # duplicates on both sides (NULL, "null", "" and repeats)
    _check_join(ctx, left, right, ["name"], ["name"], jt)
    uniq = pa.table({"name": pa.array([f"user-{k}" + "x" * (k % 7) for k in range(3_000)], pa.string()), "w": np.arange(3_000) * 0.5},
                    schema=pa.schema([pa.field("name", pa.string(), False), pa.field("w", pa.float64(), False)]))
    _check_join(ctx, left, uniq, ["name"], ["name"], jt)            # unique build keys, no NULLs on the build side


def test_hash_join_exec_on_a_string_key_and_an_integer_key(ctx):
    left, right = _string_key_tables(6, 8_000, 3_000, 500)
    _check_join(ctx, left, right, ["name", "grp"], ["name", "tag"], 0)


def test_string_key_hash_collisions_are_filtered_out(ctx, monkeypatch):
    """With the hash cut to 6 bits nearly every bucket mixes different strings: the byte-for-byte check
    must remove exactly the false pairs from an Inner join, and a Left join must refuse rather than guess."""
    from igloo_b200._lib import IglooCudaError
    from igloo_b200.operators import HashJoinExec, MemoryScanExec
    monkeypatch.setenv("IGL_UTF8_KEY_HASH_BITS", "6")
    left, right = _string_key_tables(7, 5_000, 1_200, 900)
    _check_join(ctx, left, right, ["name"], ["name"], 0)
    with pytest.raises(IglooCudaError) as e:
        collect(HashJoinExec(MemoryScanExec(left), MemoryScanExec(right), [col("name")], [col("name")], 1, ctx))
    assert e.value.code == -2 and "different strings" in str(e.value)


def test_string_key_paired_with_an_integer_key_is_refused(ctx):
    from igloo_b200._lib import IglooCudaError
    from igloo_b200.operators import HashJoinExec, MemoryScanExec
    left, right = _string_key_tables(8, 100, 100, 10)
    with pytest.raises(IglooCudaError) as e:
        collect(HashJoinExec(MemoryScanExec(left), MemoryScanExec(right), [col("name")], [col("tag")], 0, ctx))
    assert e.value.code == -2
