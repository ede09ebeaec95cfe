"""CPU tests of the host logic above the C ABI: the planner arms mirror
physical_planner.rs:23-140 (no device work happens until execute())."""
import pyarrow as pa
import pytest

from igloo_b200.expr import PlanError, col, lit
from igloo_b200.operators import FilterExec, FusedFilterProjectExec, HashJoinExec, MemoryScanExec, PhysicalPlan, ProjectionExec
from igloo_b200.physical_planner import Aggregate, Filter, Join, MemoryCatalog, PhysicalPlanner, Projection, TableScan


def catalog():
    cat = MemoryCatalog()
    cat.register_table("users", pa.table({"id": pa.array([1, 2], pa.int32()), "name": ["a", "b"], "age": pa.array([25, 30], pa.int32())}))
    cat.register_table("orders", pa.table({"id": pa.array([1], pa.int32()), "total": [1.5]}))
    return cat


def test_projection_schema_names_types_nullability():
    p = PhysicalPlanner(catalog(), fuse=False).create_physical_plan(
        Projection(TableScan("users"), [col("name"), (col("age") + 1).alias("next"), col("age") > 25]))
    assert isinstance(p, PhysicalPlan) and p.kind == PhysicalPlan.Projection and isinstance(p.exec, ProjectionExec)
    s = p.schema()
    assert s.names == ["name", "next", "age > Int64(25)"]   # display_name() of the expression as written (physical_planner.rs:66)
    assert [f.type for f in s] == [pa.string(), pa.int32(), pa.bool_()]


def test_filter_arm_and_fusion_peephole():
    cat = catalog()
    unfused = PhysicalPlanner(cat, fuse=False).create_physical_plan(Projection(Filter(TableScan("users"), col("age") > 25), [col("name")]))
    assert isinstance(unfused.exec, ProjectionExec) and isinstance(unfused.exec.input.exec, FilterExec)
    fused = PhysicalPlanner(cat, fuse=True).create_physical_plan(Projection(Filter(TableScan("users"), col("age") > 25), [col("name")]))
    assert isinstance(fused.exec, FusedFilterProjectExec) and isinstance(fused.exec.input.exec, MemoryScanExec)
    assert fused.schema() == unfused.schema()
    f = PhysicalPlanner(cat).create_physical_plan(Filter(TableScan("users"), col("age") > 25))
    assert f.kind == PhysicalPlan.Filter and f.schema() == cat.get_table("users").schema   # filter.rs:70-72


def test_join_arm_schema_and_type_mapping():
    cat = catalog()
    j = PhysicalPlanner(cat).create_physical_plan(Join(TableScan("users"), TableScan("orders"), [(col("id"), col("id"))], "Left"))
    assert isinstance(j.exec, HashJoinExec) and j.schema().names == ["id", "name", "age", "right_id", "total"]   # hash_join.rs:54-63
    with pytest.raises(NotImplementedError, match="Join type"):
        PhysicalPlanner(cat).create_physical_plan(Join(TableScan("users"), TableScan("orders"), [(col("id"), col("id"))], "LeftSemi"))


def test_unsupported_nodes_and_missing_tables():
    cat = catalog()
    with pytest.raises(NotImplementedError, match="not implemented"):          # physical_planner.rs:136-138
        PhysicalPlanner(cat).create_physical_plan(Aggregate(TableScan("users")))
    with pytest.raises(PlanError, match="not found"):                          # :32-33
        PhysicalPlanner(cat).create_physical_plan(TableScan("nope"))
    with pytest.raises(PlanError, match="No field named"):
        PhysicalPlanner(cat).create_physical_plan(Filter(TableScan("users"), col("zzz") > 1))
    from igloo_b200._lib import IglooCudaError
    with pytest.raises(IglooCudaError, match="boolean"):                       # filter.rs:50-55
        PhysicalPlanner(cat).create_physical_plan(Filter(TableScan("users"), col("age") + 1))
