"""pyarrow <-> igloo type mapping and schema helpers."""
from __future__ import annotations

from typing import List

import pyarrow as pa

from .expr import Field, T_BOOL, T_F64, T_I32, T_I64, T_UTF8

PA_TO_T = {pa.bool_(): T_BOOL, pa.int32(): T_I32, pa.int64(): T_I64, pa.float64(): T_F64, pa.string(): T_UTF8}
T_TO_PA = {v: k for k, v in PA_TO_T.items()}


def fields_of(schema: pa.Schema) -> List[Field]:
    out = []
    for f in schema:
        if f.type not in PA_TO_T:
            raise NotImplementedError(f"Arrow type {f.type} is not supported by igloo-cuda")
        out.append(Field(f.name, PA_TO_T[f.type], f.nullable))
    return out


def schema_of(fields) -> pa.Schema:
    return pa.schema([pa.field(f.name, T_TO_PA[f.type], f.nullable) for f in fields])
