"""Shared helpers for the parity tests: pyarrow <-> oracle columns, multiset compare."""
from __future__ import annotations

import math
import struct
from typing import List, Sequence

import numpy as np
import pyarrow as pa

from oracle import oracle as orc

PA_TO_T = {pa.bool_(): orc.T_BOOL, pa.int32(): orc.T_I32, pa.int64(): orc.T_I64,
           pa.float64(): orc.T_F64, pa.string(): orc.T_UTF8}
T_TO_PA = {v: k for k, v in PA_TO_T.items()}


def arrow_to_columns(tbl) -> List[orc.Column]:
    """pyarrow Table / RecordBatch -> oracle columns."""
    if isinstance(tbl, pa.RecordBatch):
        tbl = pa.Table.from_batches([tbl])
    tbl = tbl.combine_chunks()
    out = []
    for f in tbl.schema:
        arr = tbl.column(f.name)
        arr = arr.chunk(0) if arr.num_chunks else pa.array([], f.type)
        t = PA_TO_T[f.type]
        n = len(arr)
        valid = None
        if arr.null_count:
            valid = np.array([0 if v is None else 1 for v in arr.to_pylist()], np.uint8)
        if t == orc.T_UTF8:
            py = arr.to_pylist()
            bs = [b"" if s is None else s.encode() for s in py]
            off = np.zeros(n + 1, np.int32)
            np.cumsum([len(b) for b in bs], out=off[1:])
            vals = np.frombuffer(b"".join(bs), np.uint8)
            out.append(orc.Column(f.name, t, vals, valid, off, f.nullable))
        elif t == orc.T_BOOL:
            vals = np.array([1 if v else 0 for v in arr.to_pylist()], np.uint8)
            out.append(orc.Column(f.name, t, vals, valid, None, f.nullable))
        else:
            vals = arr.fill_null(0).to_numpy(zero_copy_only=False) if arr.null_count else arr.to_numpy(zero_copy_only=False)
            out.append(orc.Column(f.name, t, vals, valid, None, f.nullable))
    return out


def columns_to_arrow(cols: Sequence[orc.Column]) -> pa.Table:
    arrays, fields = [], []
    for c in cols:
        mask = None if c.valid is None else (c.valid == 0)
        if c.type == orc.T_UTF8:
            arrays.append(pa.array(c.to_pylist(), pa.string()))
        elif c.type == orc.T_BOOL:
            arrays.append(pa.array(c.values.astype(bool), pa.bool_(), mask=mask))
        else:
            arrays.append(pa.array(c.values, T_TO_PA[c.type], mask=mask))
        fields.append(pa.field(c.name, T_TO_PA[c.type], c.nullable))
    return pa.Table.from_arrays(arrays, schema=pa.schema(fields))


def _key(v):
    """Sort/compare key that is exact on bits (NaN == NaN, -0.0 != 0.0)."""
    if v is None:
        return (0, 0)
    if isinstance(v, float):
        return (1, struct.unpack("<q", struct.pack("<d", v))[0])
    if isinstance(v, str):
        return (2, v)
    return (1, int(v))


def rows_multiset(tbl) -> List[tuple]:
    if isinstance(tbl, pa.RecordBatch):
        tbl = pa.Table.from_batches([tbl])
    cols = [tbl.column(i).to_pylist() for i in range(tbl.num_columns)]
    rows = [tuple(_key(c[i]) for c in cols) for i in range(tbl.num_rows)]
    rows.sort()
    return rows


def assert_same_multiset(got, want, float_rel: float = 0.0):
    """Bit-exact sorted-multiset equality; `float_rel` > 0 relaxes f64 columns
    (north_star: float projections within 1e-12 relative)."""
    assert got.num_columns == want.num_columns, (got.schema, want.schema)
    assert got.num_rows == want.num_rows, (got.num_rows, want.num_rows)
    for fg, fw in zip(got.schema, want.schema):
        assert fg.type == fw.type and fg.name == fw.name, (fg, fw)
    if float_rel == 0.0:
        assert rows_multiset(got) == rows_multiset(want)
        return
    # sort on the non-float columns, then compare floats with tolerance
    fcols = [i for i, f in enumerate(got.schema) if f.type == pa.float64()]
    icols = [i for i in range(got.num_columns) if i not in fcols]

    def split(t):
        cols = [t.column(i).to_pylist() for i in range(t.num_columns)]
        rows = [(tuple(_key(cols[i][r]) for i in icols), tuple(cols[i][r] for i in fcols)) for r in range(t.num_rows)]
        rows.sort(key=lambda x: (x[0], tuple(_key(v) for v in x[1])))
        return rows
    for (ig, fg_), (iw, fw_) in zip(split(got), split(want)):
        assert ig == iw
        for a, b in zip(fg_, fw_):
            if a is None or b is None:
                assert a is b
            elif math.isnan(a) or math.isnan(b):
                assert math.isnan(a) and math.isnan(b)
            else:
                assert abs(a - b) <= float_rel * max(abs(a), abs(b)), (a, b)


def tight_fields(tbl):
    """Fields whose nullable flag is true only where nulls actually occur, so
    the no-null kernel variants get exercised (pyarrow marks everything nullable)."""
    from igloo_b200.arrow import PA_TO_T
    from igloo_b200.expr import Field
    return [Field(f.name, PA_TO_T[f.type], tbl.column(f.name).null_count > 0) for f in tbl.schema]
