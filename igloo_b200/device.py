"""Device-resident columns and thin Python handles over the C ABI.

PyTorch is used for one thing only: owning device memory (`torch.empty(...,
device="cuda")`) and naming the current CUDA stream.  Every computation is a
call into libigloo_cuda.so through ctypes; there is no torch or CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import DevCol, check, lib
from .expr import Expr, Field, lower, T_BOOL, T_F64, T_I32, T_I64, T_UTF8

_TORCH_DT = None


def _torch():
    import torch
    global _TORCH_DT
    if _TORCH_DT is None:
        _TORCH_DT = {T_I32: torch.int32, T_I64: torch.int64, T_F64: torch.float64}
    return torch


def _pad64(n: int) -> int:
    return (n + 63) // 64 * 64 + 64


class Context:
    """igl_ctx: one per GPU (device ordinal), owns streams and the buffer pool."""

    def __init__(self, device: int = 0, coalesce_rows: int = 0, hash_load_factor: float = 0.0):
        cfg = _lib.Config(device, 0, coalesce_rows, hash_load_factor, 0)
        h = C.c_void_p()
        check(lib.igl_ctx_create(C.byref(cfg), C.byref(h)))
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            lib.igl_ctx_destroy(self._h)
            self._h = None

    __del__ = close

    @property
    def launches(self) -> int:
        return lib.igl_ctx_launch_count(self._h)

    def stream(self) -> int:
        torch = _torch()
        # torch's default stream is the legacy default stream (handle 0); the C ABI
        # reads a NULL stream as "the context's own stream", so name it explicitly
        # (cudaStreamLegacy == 0x1) to keep kernels on the stream torch events see.
        return torch.cuda.current_stream(self.device).cuda_stream or 1

    def compile(self, e: Expr, schema: Sequence[Field]) -> "CompiledExpr":
        return CompiledExpr(self, e, schema)

    # synthetic data -----------------------------------------------------------
    def gen(self, kind: int, seed: int, col_id: int, row0: int, n: int, m: int = 0, bits: int = 0,
            keys: Optional["DeviceColumn"] = None) -> "DeviceColumn":
        torch = _torch()
        t = {1: T_I32, 2: T_I64, 3: T_F64}.get(kind, T_I64)   # igl_gen_kind: 4-8 are Int64 generators
        out = DeviceColumn.empty(t, n, self.device)
        check(lib.igl_gen_column(self._h, kind, seed, col_id, row0, n, m, bits,
                                 keys.data.data_ptr() if keys is not None else None,
                                 out.data.data_ptr(), self.stream()))
        return out

    def digest(self, cols: Sequence["DeviceColumn"], nrows: int) -> Tuple[int, int, int]:
        arr = (DevCol * len(cols))(*[c.c_struct() for c in cols])
        out = (C.c_uint64 * 3)()
        check(lib.igl_digest_rows(self._h, arr, len(cols), nrows, out, self.stream()))
        return int(out[0]), int(out[1]), int(out[2])


class DeviceColumn:
    """One Arrow column resident in HBM: values (+ validity bitmap)."""

    def __init__(self, type_: int, length: int, data, validity=None):
        self.type, self.length, self.data, self.validity = type_, length, data, validity

    @staticmethod
    def empty(type_: int, n: int, device: int = 0, nullable: bool = False) -> "DeviceColumn":
        torch = _torch()
        dev = f"cuda:{device}"
        if type_ == T_BOOL:
            data = torch.empty(_pad64((n + 7) // 8), dtype=torch.uint8, device=dev)
        else:
            w = 4 if type_ == T_I32 else 8
            data = torch.empty(_pad64(n * w) // w, dtype=_TORCH_DT[type_], device=dev)
        validity = torch.empty(_pad64((n + 7) // 8), dtype=torch.uint8, device=dev) if nullable else None
        return DeviceColumn(type_, n, data, validity)

    @staticmethod
    def from_arrow(arr, device: int = 0) -> "DeviceColumn":
        """Pinned-free convenience upload used by the tests (the streaming upload
        path lives behind the operator API)."""
        import pyarrow as pa
        torch = _torch()
        from .arrow import PA_TO_T
        if isinstance(arr, pa.ChunkedArray):
            arr = arr.combine_chunks() if arr.num_chunks != 1 else arr.chunk(0)
        if arr.offset != 0:
            arr = pa.concat_arrays([arr])  # re-materialise at offset 0
        t = PA_TO_T[arr.type]
        n = len(arr)
        bufs = arr.buffers()
        col = DeviceColumn.empty(t, n, device, nullable=arr.null_count > 0)

        def put(dst, buf, nbytes):
            if nbytes == 0 or buf is None:
                return
            host = torch.frombuffer(memoryview(buf)[:nbytes], dtype=torch.uint8)
            dst.view(torch.uint8)[:nbytes].copy_(host)
        if t == T_BOOL:
            put(col.data, bufs[1], (n + 7) // 8)
        else:
            w = 4 if t == T_I32 else 8
            put(col.data, bufs[1], n * w)
        if col.validity is not None:
            put(col.validity, bufs[0], (n + 7) // 8)
        return col

    def to_arrow(self, nrows: Optional[int] = None):
        import pyarrow as pa
        from .arrow import T_TO_PA
        n = self.length if nrows is None else nrows
        if self.type == T_BOOL:
            data = pa.py_buffer(self.data[: (n + 7) // 8 + 1].cpu().numpy().tobytes())
        else:
            data = pa.py_buffer(self.data[:n].cpu().numpy().tobytes())
        valid = None
        if self.validity is not None:
            valid = pa.py_buffer(self.validity[: (n + 7) // 8 + 1].cpu().numpy().tobytes())
        return pa.Array.from_buffers(T_TO_PA[self.type], n, [valid, data])

    def c_struct(self) -> DevCol:
        return DevCol(self.type, 0, self.data.data_ptr(),
                      self.validity.data_ptr() if self.validity is not None else None, None)


class CompiledExpr:
    """igl_expr: `Arc<dyn PhysicalExpr>` lowered against a child schema."""

    def __init__(self, ctx: Context, e: Expr, schema: Sequence[Field]):
        self.bytecode, self.type, self.nullable = lower(e, schema)
        self.name = e.display_name()
        types = (C.c_int32 * max(1, len(schema)))(*[f.type for f in schema])
        nulls = bytes(1 if f.nullable else 0 for f in schema)
        h = C.c_void_p()
        check(lib.igl_expr_compile(ctx._h, self.bytecode, len(self.bytecode), types, nulls, len(schema), C.byref(h)))
        self._h = h
        t, nl = C.c_int32(), C.c_int32()
        check(lib.igl_expr_type(h, C.byref(t), C.byref(nl)))
        assert t.value == self.type, (t.value, self.type)

    def __del__(self):
        if getattr(self, "_h", None):
            lib.igl_expr_destroy(self._h)
            self._h = None


class FilterProject:
    """igl_fp: FilterExec (pred only), ProjectionExec (proj only) or both fused."""

    def __init__(self, ctx: Context, schema: Sequence[Field], pred: Optional[Expr], proj: Sequence[Expr] = ()):
        self.ctx, self.schema = ctx, list(schema)
        self._pred = ctx.compile(pred, schema) if pred is not None else None
        self._proj = [ctx.compile(e, schema) for e in proj]
        types = (C.c_int32 * len(schema))(*[f.type for f in schema])
        nulls = bytes(1 if f.nullable else 0 for f in schema)
        parr = (C.c_void_p * max(1, len(self._proj)))(*[p._h for p in self._proj])
        h = C.c_void_p()
        check(lib.igl_fp_create(ctx._h, types, nulls, len(schema), self._pred._h if self._pred else None,
                                parr, len(self._proj), C.byref(h)))
        self._h = h
        self.out_fields: List[Field] = []
        for j in range(lib.igl_fp_num_outputs(h)):
            t, nl = C.c_int32(), C.c_int32()
            check(lib.igl_fp_output_type(h, j, C.byref(t), C.byref(nl)))
            name = self._proj[j].name if self._proj else self.schema[j].name
            self.out_fields.append(Field(name, t.value, bool(nl.value)))

    def __del__(self):
        # a handle must not outlive its context (igl_fp_destroy returns blocks to the pool)
        if getattr(self, "_h", None) and getattr(self.ctx, "_h", None):
            lib.igl_fp_destroy(self._h)
        self._h = None

    @property
    def is_specialised(self) -> bool:
        """True once the plan runs on kernels compiled for it at run time (igl_fp_is_specialised)."""
        return bool(lib.igl_fp_is_specialised(self._h))

    def prepare(self) -> bool:
        """igl_fp_prepare: compile the plan's specialised kernels now; True when they are in use."""
        check(lib.igl_fp_prepare(self._h))
        return self.is_specialised

    def alloc_outputs(self, nrows: int) -> List[DeviceColumn]:
        return [DeviceColumn.empty(f.type, nrows, self.ctx.device, nullable=f.nullable) for f in self.out_fields]

    def _arrs(self, cols, out):
        a = (DevCol * len(cols))(*[c.c_struct() for c in cols])
        o = (DevCol * len(out))(*[c.c_struct() for c in out])
        return a, o

    def run(self, cols: Sequence[DeviceColumn], nrows: int, out: Optional[List[DeviceColumn]] = None):
        out = out if out is not None else self.alloc_outputs(nrows)
        a, o = self._arrs(cols, out)
        n = C.c_int64(0)
        check(lib.igl_fp_run(self._h, a, nrows, o, C.byref(n), self.ctx.stream()))
        return out, n.value

    def launch(self, cols, nrows, out):
        a, o = self._arrs(cols, out)
        check(lib.igl_fp_launch(self._h, a, nrows, o, self.ctx.stream()))

    def finish(self) -> int:
        n = C.c_int64(0)
        check(lib.igl_fp_finish(self._h, C.byref(n), self.ctx.stream()))
        return n.value


JOIN_INNER, JOIN_LEFT, JOIN_RIGHT, JOIN_FULL = 0, 1, 2, 3


def join_output_fields(left: Sequence[Field], right: Sequence[Field], join_type: int = JOIN_INNER) -> List[Field]:
    """HashJoinExec::new output schema (hash_join.rs:35-76): left fields, then
    right fields, a right field whose name exists on the left renamed
    `right_<name>` keeping type and nullability."""
    names = {f.name for f in left}
    out = [Field(f.name, f.type, f.nullable) for f in left]
    for f in right:
        out.append(Field("right_" + f.name if f.name in names else f.name, f.type, f.nullable))
    return out


class HashJoin:
    """igl_join: build from the right child, probe with the left (hash_join.rs:100-213)."""

    def __init__(self, ctx: Context, join_type: int, left: Sequence[Field], right: Sequence[Field],
                 left_key, right_key):
        """left_key / right_key: one expression or a list of them, pairwise (HashJoinExec::new takes
        vectors, hash_join.rs:38-39); anything beyond one plain column goes through igl_join_create_ex."""
        self.ctx, self.join_type = ctx, join_type
        self.left, self.right = list(left), list(right)
        lks = list(left_key) if isinstance(left_key, (list, tuple)) else [left_key]
        rks = list(right_key) if isinstance(right_key, (list, tuple)) else [right_key]
        assert len(lks) == len(rks) and lks, "join keys come in pairs"
        self._lks = [ctx.compile(e, left) for e in lks]
        self._rks = [ctx.compile(e, right) for e in rks]
        self._lk, self._rk = self._lks[0], self._rks[0]
        lt = (C.c_int32 * len(left))(*[f.type for f in left])
        rt = (C.c_int32 * len(right))(*[f.type for f in right])
        ln = bytes(1 if f.nullable else 0 for f in left)
        rn = bytes(1 if f.nullable else 0 for f in right)
        h = C.c_void_p()
        la = (C.c_void_p * len(lks))(*[k._h for k in self._lks])
        ra = (C.c_void_p * len(rks))(*[k._h for k in self._rks])
        check(lib.igl_join_create_ex(ctx._h, join_type, lt, ln, len(left), rt, rn, len(right), la, ra, len(lks), C.byref(h)))
        self._h = h
        self.out_fields = join_output_fields(left, right, join_type)
        self._right_cols = None
        self.left_join = join_type in (JOIN_LEFT, JOIN_FULL)

    def __del__(self):
        if getattr(self, "_h", None) and getattr(self.ctx, "_h", None):
            lib.igl_join_destroy(self._h)
        self._h = None

    def build(self, cols: Sequence[DeviceColumn], nrows: int):
        self._right_cols = list(cols)  # borrowed by the library until destroy
        a = (DevCol * len(cols))(*[c.c_struct() for c in cols])
        check(lib.igl_join_build(self._h, a, nrows, self.ctx.stream()))

    @property
    def has_duplicates(self) -> bool:
        return bool(lib.igl_join_has_duplicates(self._h))

    def probe_count(self, cols: Sequence[DeviceColumn], nrows: int) -> int:
        a = (DevCol * len(cols))(*[c.c_struct() for c in cols])
        n = C.c_int64(0)
        check(lib.igl_join_probe_count(self._h, a, nrows, C.byref(n), self.ctx.stream()))
        return n.value

    def phase_ms(self):
        """Device time of the last build / probe: build partitioning, build, probe partitioning, probe kernel."""
        out = (C.c_double * 4)()
        check(lib.igl_join_phase_ms(self._h, out))
        return [float(x) for x in out]

    def alloc_outputs(self, capacity: int, left_cols: Sequence[DeviceColumn]) -> List[DeviceColumn]:
        out = []
        nl = len(self.left)
        for i, f in enumerate(self.out_fields):
            if i < nl:
                nullable = left_cols[i].validity is not None
            else:
                nullable = self.left_join or self._right_cols[i - nl].validity is not None
            out.append(DeviceColumn.empty(f.type, capacity, self.ctx.device, nullable=nullable))
        return out

    def probe(self, cols: Sequence[DeviceColumn], nrows: int, out: Optional[List[DeviceColumn]] = None,
              capacity: Optional[int] = None):
        if out is None:
            capacity = self.probe_count(cols, nrows) if (self.has_duplicates or capacity is None) else capacity
            out = self.alloc_outputs(capacity, cols)
        elif capacity is None:
            capacity = out[0].length
        a = (DevCol * len(cols))(*[c.c_struct() for c in cols])
        o = (DevCol * len(out))(*[c.c_struct() for c in out])
        n = C.c_int64(0)
        check(lib.igl_join_probe(self._h, a, nrows, o, capacity, C.byref(n), self.ctx.stream()))
        return out, n.value
