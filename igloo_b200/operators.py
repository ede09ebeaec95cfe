"""Host-side mirror of the reference's operator API, running on the GPU library.

Same names and argument meaning as crates/engine/src:
  ExecutionPlan (physical_plan.rs:10-17)  execute() -> stream of RecordBatch, schema()
  FilterExec::new(input, predicate)                         filter.rs:19
  ProjectionExec::new(input, [(expr, name)], schema)        projection.rs:19-29
  HashJoinExec::new(left, right, left_keys, right_keys, jt) hash_join.rs:35-76
  ParquetScanExec::new(file_path, schema, projection)       parquet_scan.rs:21-32
  PhysicalPlan enum + From impls                            physical_plan.rs:20-71
`execute()` is a Python generator of pyarrow.RecordBatch (the async BoxStream of
the reference); batches cross the C ABI through the Arrow C Data Interface
(`_export_to_c` / `_import_from_c`).  Every operator body is a call sequence into
libigloo_cuda.so — there is no pyarrow-compute or CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterator, List, Optional, Sequence, Tuple

import pyarrow as pa

from . import _lib
from ._lib import ArrowArray, ArrowSchema, IGL_AGAIN, IGL_EOS, IGL_OK, check, lib
from .arrow import fields_of, schema_of
from .expr import Expr, Field, PlanError, col, lower, projection_field

SCAN_BATCH_ROWS = 1024  # parquet_scan.rs:54

_default_ctx = None


def default_context():
    """One igl_ctx per process on cuda:0 unless the caller passes its own."""
    global _default_ctx
    if _default_ctx is None:
        from .device import Context
        _default_ctx = Context(0)
    return _default_ctx


class JoinType:  # hash_join.rs:16-21
    Inner, Left, Right, Full = 0, 1, 2, 3


class ExecutionPlan:
    def execute(self) -> Iterator[pa.RecordBatch]:
        raise NotImplementedError

    def schema(self) -> pa.Schema:
        raise NotImplementedError


# ---- C Data Interface plumbing ------------------------------------------------
def _export_schema(schema: pa.Schema) -> ArrowSchema:
    s = ArrowSchema()
    schema._export_to_c(C.addressof(s))
    return s


def _release_schema(s: ArrowSchema):
    if s.release:
        C.CFUNCTYPE(None, C.POINTER(ArrowSchema))(s.release)(C.byref(s))


def _push(fn, op, batch: pa.RecordBatch):
    a = ArrowArray()
    batch._export_to_c(C.addressof(a))
    try:
        check(fn(op, C.byref(a)))
    finally:  # the library copied what it needs; the batch is only borrowed
        if a.release:
            C.CFUNCTYPE(None, C.POINTER(ArrowArray))(a.release)(C.byref(a))


def _drain(op, schema: pa.Schema, until_eos: bool) -> Iterator[pa.RecordBatch]:
    while True:
        a = ArrowArray()
        rc = check(lib.igl_op_pull(op, C.byref(a)))
        if rc == IGL_OK:
            s = _export_schema(schema)
            yield pa.RecordBatch._import_from_c(C.addressof(a), C.addressof(s))
        elif rc == IGL_EOS or (rc == IGL_AGAIN and not until_eos):
            return
        else:
            raise RuntimeError("igl_op_pull returned IGL_AGAIN after finish_input")


def _op_schema(op) -> pa.Schema:
    s = ArrowSchema()
    check(lib.igl_op_schema(op, C.byref(s)))
    return pa.Schema._import_from_c(C.addressof(s))


# ---- scans (upstream neighbours of the hot path; host side) ---------------------
class MemoryScanExec(ExecutionPlan):
    """An in-memory table streamed as 1024-row batches (what ParquetScanExec emits)."""

    def __init__(self, table: pa.Table, batch_rows: int = SCAN_BATCH_ROWS):
        self.table, self.batch_rows = table, batch_rows

    def execute(self):
        yield from self.table.to_batches(max_chunksize=self.batch_rows)

    def schema(self):
        return self.table.schema


class ParquetScanExec(ExecutionPlan):
    """parquet_scan.rs:40-85: host Parquet reader -> 1024-row batches, optional
    column projection.  Decoding stays on the host."""

    def __init__(self, file_path: str, schema: pa.Schema, projection: Optional[Sequence[int]] = None):
        self.file_path, self._schema, self.projection = file_path, schema, projection

    def execute(self):
        import pyarrow.parquet as pq
        f = pq.ParquetFile(self.file_path)
        names = None
        if self.projection is not None:
            names = [f.schema_arrow.names[i] for i in self.projection]
        for b in f.iter_batches(batch_size=SCAN_BATCH_ROWS, columns=names):
            yield b

    def schema(self):
        return self._schema


# ---- the three hot-path operators ---------------------------------------------------
class _FilterProjectBase(ExecutionPlan):
    def __init__(self, input: ExecutionPlan, predicate: Optional[Expr], expressions: Sequence[Tuple[Expr, str]],
                 schema: Optional[pa.Schema], ctx=None):
        self.input, self.predicate, self.expressions, self._schema = input, predicate, list(expressions), schema
        self.ctx = ctx

    def _create(self):
        ctx = self.ctx or default_context()
        in_schema = self.input.schema()
        fields = fields_of(in_schema)
        pred = lower(self.predicate, fields)[0] if self.predicate is not None else None
        projs = [lower(e, fields)[0] for e, _ in self.expressions]
        names = [n.encode() for _, n in self.expressions]
        n = len(projs)
        parr = (C.c_char_p * max(1, n))(*projs)
        plen = (C.c_size_t * max(1, n))(*[len(p) for p in projs])
        pnam = (C.c_char_p * max(1, n))(*names)
        s = _export_schema(in_schema)
        h = C.c_void_p()
        try:
            check(lib.igl_op_filter_project_create(ctx._h, C.byref(s), pred, len(pred) if pred else 0,
                                                   parr, plen, pnam, n, C.byref(h)))
        finally:
            _release_schema(s)
        return h

    def execute(self):
        op = self._create()
        try:
            out_schema = self.schema()
            for batch in self.input.execute():
                _push(lib.igl_op_push, op, batch)
                yield from _drain(op, out_schema, until_eos=False)
            check(lib.igl_op_finish_input(op))
            yield from _drain(op, out_schema, until_eos=True)
        finally:
            lib.igl_op_destroy(op)


class FilterExec(_FilterProjectBase):
    """filter.rs:13-72.  Keeps rows whose predicate is valid and true; the schema is
    the input schema; chunks that keep no row are not emitted."""

    def __init__(self, input: ExecutionPlan, predicate: Expr, ctx=None):
        super().__init__(input, predicate, [], None, ctx)
        _, t, _ = lower(predicate, fields_of(input.schema()))
        if t != 1:
            raise _lib.IglooCudaError(-7, "Filter predicate must return boolean array")  # filter.rs:50-55

    def schema(self):
        return self.input.schema()


class ProjectionExec(_FilterProjectBase):
    """projection.rs:12-76.  `expressions` = [(expr, name)], `schema` from the planner."""

    def __init__(self, input: ExecutionPlan, expressions: Sequence[Tuple[Expr, str]], schema: pa.Schema, ctx=None):
        super().__init__(input, None, expressions, schema, ctx)

    def schema(self):
        return self._schema


class FusedFilterProjectExec(_FilterProjectBase):
    """Projection(Filter(x)) as one kernel launch per chunk — what the planner emits
    when it sees physical_planner.rs:52-94 back to back.  Same stream as the pair."""

    def __init__(self, input, predicate, expressions, schema, ctx=None):
        super().__init__(input, predicate, expressions, schema, ctx)

    def schema(self):
        return self._schema


class HashJoinExec(ExecutionPlan):
    """hash_join.rs:25-216.  Builds from `right`, probes with `left`; output schema =
    left fields then right fields, name clashes renamed right_<name>."""

    def __init__(self, left: ExecutionPlan, right: ExecutionPlan, left_keys: Sequence[Expr], right_keys: Sequence[Expr],
                 join_type: int, ctx=None):
        if len(left_keys) != len(right_keys) or not left_keys:
            raise ValueError("join keys come in pairs (physical_planner.rs:106-120)")
        self.left, self.right, self.left_keys, self.right_keys = left, right, list(left_keys), list(right_keys)
        self.join_type, self.ctx = join_type, ctx
        lf, rf = fields_of(left.schema()), fields_of(right.schema())
        names = {f.name for f in lf}
        left_join = join_type in (JoinType.Left, JoinType.Full)
        out = [pa.field(f.name, left.schema().field(i).type, f.nullable) for i, f in enumerate(lf)]
        for i, f in enumerate(rf):
            out.append(pa.field("right_" + f.name if f.name in names else f.name, right.schema().field(i).type,
                                f.nullable or left_join))
        self._schema = pa.schema(out)

    def schema(self):
        return self._schema

    def execute(self):
        ctx = self.ctx or default_context()
        lf, rf = fields_of(self.left.schema()), fields_of(self.right.schema())
        lks = [lower(e, lf)[0] for e in self.left_keys]
        rks = [lower(e, rf)[0] for e in self.right_keys]
        nk = len(lks)
        ls, rs = _export_schema(self.left.schema()), _export_schema(self.right.schema())
        h = C.c_void_p()
        try:
            check(lib.igl_op_hash_join_create_ex(ctx._h, self.join_type, C.byref(ls), C.byref(rs),
                                                 (C.c_char_p * nk)(*lks), (C.c_size_t * nk)(*[len(b) for b in lks]),
                                                 (C.c_char_p * nk)(*rks), (C.c_size_t * nk)(*[len(b) for b in rks]), nk, C.byref(h)))
        finally:
            _release_schema(ls)
            _release_schema(rs)
        op = h
        try:
            for batch in self.right.execute():          # build phase drains the right child (hash_join.rs:100-128)
                _push(lib.igl_op_push_build, op, batch)
            check(lib.igl_op_finish_build(op))
            for batch in self.left.execute():           # probe phase streams the left child (:131-213)
                _push(lib.igl_op_push, op, batch)
                yield from _drain(op, self._schema, until_eos=False)
            check(lib.igl_op_finish_input(op))
            yield from _drain(op, self._schema, until_eos=True)
        finally:
            lib.igl_op_destroy(op)


class PhysicalPlan(ExecutionPlan):
    """physical_plan.rs:20-71: closed set of plan nodes with delegating execute/schema."""

    ParquetScan, Projection, Filter, HashJoin = "ParquetScan", "Projection", "Filter", "HashJoin"

    def __init__(self, exec_: ExecutionPlan):
        kinds = {ParquetScanExec: self.ParquetScan, MemoryScanExec: self.ParquetScan, ProjectionExec: self.Projection,
                 FusedFilterProjectExec: self.Projection, FilterExec: self.Filter, HashJoinExec: self.HashJoin}
        if type(exec_) not in kinds:
            raise TypeError(f"{type(exec_).__name__} is not a PhysicalPlan variant")
        self.kind, self.exec = kinds[type(exec_)], exec_

    @staticmethod
    def from_(exec_: ExecutionPlan) -> "PhysicalPlan":
        return PhysicalPlan(exec_)

    def execute(self):
        return self.exec.execute()

    def schema(self):
        return self.exec.schema()
