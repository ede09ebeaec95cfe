"""LogicalPlan -> PhysicalPlan, mirroring crates/engine/src/physical_planner.rs:23-140.

The reference consumes DataFusion's LogicalPlan; the four node kinds it handles
(TableScan :28-50, Projection :52-80, Filter :82-94, Join :96-134) are modelled
here as small dataclasses, everything else raises NotImplementedError like the
fall-through arm (:136-138).  Operator SELECTION happens here: the three hot-path
nodes become GPU operators, and Projection(Filter(x)) is fused into one operator.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple, Union

import pyarrow as pa

from .arrow import fields_of, schema_of
from .expr import Expr, Field, PlanError, col, lower, projection_field
from .operators import (ExecutionPlan, FilterExec, FusedFilterProjectExec, HashJoinExec, JoinType, MemoryScanExec,
                        ParquetScanExec, PhysicalPlan, ProjectionExec)


@dataclass
class TableScan:
    table_name: str
    projection: Optional[List[int]] = None


@dataclass
class Projection:
    input: "LogicalPlan"
    expr: List[Expr]


@dataclass
class Filter:
    input: "LogicalPlan"
    predicate: Expr


@dataclass
class Join:
    left: "LogicalPlan"
    right: "LogicalPlan"
    on: List[Tuple[Expr, Expr]]
    join_type: str = "Inner"


@dataclass
class Aggregate:  # stands for every node the reference does not lower
    input: "LogicalPlan"


LogicalPlan = Union[TableScan, Projection, Filter, Join, Aggregate]


class MemoryCatalog:
    """crates/common/src/catalog.rs:5-27: name -> table provider (a pyarrow Table or a Parquet path)."""

    def __init__(self):
        self._tables: Dict[str, Union[pa.Table, Tuple[str, pa.Schema]]] = {}

    def register_table(self, name: str, table):
        self._tables[name] = table

    def get_table(self, name: str):
        return self._tables.get(name)


class PhysicalPlanner:
    def __init__(self, catalog: MemoryCatalog, ctx=None, fuse: bool = True):
        self.catalog, self.ctx, self.fuse = catalog, ctx, fuse

    def create_physical_plan(self, plan: LogicalPlan) -> ExecutionPlan:
        if isinstance(plan, TableScan):                                         # :28-50
            t = self.catalog.get_table(plan.table_name)
            if t is None:
                raise PlanError(f"Table '{plan.table_name}' not found")
            if isinstance(t, pa.Table):
                tbl = t.select(plan.projection) if plan.projection is not None else t
                return PhysicalPlan(MemoryScanExec(tbl))
            path, schema = t
            projected = pa.schema([schema.field(i) for i in plan.projection]) if plan.projection is not None else schema
            return PhysicalPlan(ParquetScanExec(path, projected, plan.projection))
        if isinstance(plan, Projection):                                        # :52-80
            fused_filter = None
            if self.fuse and isinstance(plan.input, Filter):
                fused_filter = plan.input
                input_ = self.create_physical_plan(fused_filter.input)
            else:
                input_ = self.create_physical_plan(plan.input)
            fields = fields_of(input_.schema())
            exprs, out_fields = [], []
            for e in plan.expr:
                f = projection_field(e, fields)                                 # name / data_type / nullable, :66-71
                exprs.append((e, f.name))
                out_fields.append(f)
            schema = schema_of(out_fields)
            if fused_filter is not None:
                self._check_predicate(fused_filter.predicate, fields)
                return PhysicalPlan(FusedFilterProjectExec(input_, fused_filter.predicate, exprs, schema, self.ctx))
            return PhysicalPlan(ProjectionExec(input_, exprs, schema, self.ctx))
        if isinstance(plan, Filter):                                            # :82-94
            input_ = self.create_physical_plan(plan.input)
            self._check_predicate(plan.predicate, fields_of(input_.schema()))
            return PhysicalPlan(FilterExec(input_, plan.predicate, self.ctx))
        if isinstance(plan, Join):                                              # :96-134
            left = self.create_physical_plan(plan.left)
            right = self.create_physical_plan(plan.right)
            jt = {"Inner": JoinType.Inner, "Left": JoinType.Left, "Right": JoinType.Right, "Full": JoinType.Full}.get(plan.join_type)
            if jt is None:
                raise NotImplementedError(f"Join type {plan.join_type} not implemented")   # :127-129
            lk = [l for l, _ in plan.on]
            rk = [r for _, r in plan.on]
            return PhysicalPlan(HashJoinExec(left, right, lk, rk, jt, self.ctx))
        raise NotImplementedError(f"Logical plan {type(plan).__name__} not implemented")     # :136-138

    @staticmethod
    def _check_predicate(pred: Expr, fields: Sequence[Field]):
        lower(pred, fields)  # raises PlanError on unknown columns / type errors, like create_physical_expr
