"""ctypes front end of the CPU oracle (oracle/igloo_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(igloo_b200/) never imports this module.

Tables go in and out as lists of (name, type, values, valid) numpy columns so
the tests can compare them with pyarrow RecordBatches from the CUDA path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

T_BOOL, T_I32, T_I64, T_F64, T_UTF8 = 1, 2, 3, 4, 5
_NP = {T_BOOL: np.uint8, T_I32: np.int32, T_I64: np.int64, T_F64: np.float64}
JOIN_INNER, JOIN_LEFT, JOIN_RIGHT, JOIN_FULL = 0, 1, 2, 3


def build(force: bool = False) -> str:
    """Compile liboracle.so with the committed Makefile (g++ only)."""
    if force or not os.path.exists(_LIB_PATH) or (
        os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "igloo_oracle.cpp"))
    ):
        subprocess.check_call(["make", "-C", _HERE, "-s", "liboracle.so"])
    return _LIB_PATH


class _Col(C.Structure):
    _fields_ = [("type", C.c_int32), ("nullable", C.c_int32), ("n", C.c_int64),
                ("data", C.c_void_p), ("valid", C.c_void_p), ("offsets", C.c_void_p),
                ("name", C.c_char_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_last_error.restype = C.c_char_p
        L.orc_scan.restype = C.c_void_p
        L.orc_scan.argtypes = [C.POINTER(_Col), C.c_int, C.c_int64, C.c_int64]
        L.orc_filter.restype = C.c_void_p
        L.orc_filter.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        L.orc_project.restype = C.c_void_p
        L.orc_project.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t),
                                  C.POINTER(C.c_char_p), C.c_int]
        L.orc_hash_join.restype = C.c_void_p
        L.orc_hash_join.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t),
                                    C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.c_int, C.c_int, C.c_int, C.c_int]
        L.orc_table_ncols.argtypes = [C.c_void_p]
        L.orc_table_nrows.restype = C.c_int64
        L.orc_table_nrows.argtypes = [C.c_void_p]
        L.orc_table_nbatches.restype = C.c_int64
        L.orc_table_nbatches.argtypes = [C.c_void_p]
        L.orc_table_batch_size.restype = C.c_int64
        L.orc_table_batch_size.argtypes = [C.c_void_p, C.c_int64]
        L.orc_table_col.argtypes = [C.c_void_p, C.c_int, C.POINTER(_Col)]
        L.orc_table_digest.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        L.orc_table_free.argtypes = [C.c_void_p]
        L.orc_splitmix64.restype = C.c_uint64
        L.orc_splitmix64.argtypes = [C.c_uint64]
        for fn in ("orc_gen_u64",):
            getattr(L, fn).argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int64, C.c_void_p]
        L.orc_gen_i32_uniform.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int64, C.c_uint64, C.c_void_p]
        L.orc_gen_i64_uniform.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int64, C.c_uint64, C.c_void_p]
        L.orc_gen_f64_unit.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int64, C.c_void_p]
        L.orc_gen_perm_pow2.argtypes = [C.c_int64, C.c_int64, C.c_int, C.c_void_p]
        L.orc_perm_pow2.restype = C.c_uint64
        L.orc_perm_pow2.argtypes = [C.c_uint64, C.c_int]
        L.orc_perm_pow2_inv.restype = C.c_uint64
        L.orc_perm_pow2_inv.argtypes = [C.c_uint64, C.c_int]
        L.orc_bench_filter_project.restype = C.c_double
        L.orc_bench_filter_project.argtypes = [C.c_uint64, C.c_int64, C.c_int32, C.POINTER(C.c_int64)]
        L.orc_bench_join.restype = C.c_double
        L.orc_bench_join.argtypes = [C.c_uint64, C.c_int, C.c_int64, C.c_int, C.c_int, C.POINTER(C.c_int64)]
        _lib = L
    return _lib


class OracleError(RuntimeError):
    pass


class Column:
    """One host column: values (numpy), optional byte-per-row validity, UTF8 offsets."""

    def __init__(self, name: str, type_: int, values, valid=None, offsets=None, nullable: Optional[bool] = None):
        self.name, self.type = name, type_
        if type_ == T_UTF8:
            self.values = np.ascontiguousarray(values, dtype=np.uint8)
            self.offsets = np.ascontiguousarray(offsets, dtype=np.int32)
        else:
            self.values = np.ascontiguousarray(values, dtype=_NP[type_])
            self.offsets = None
        self.valid = None if valid is None else np.ascontiguousarray(valid, dtype=np.uint8)
        self.nullable = bool(self.valid is not None) if nullable is None else nullable

    def __len__(self):
        return len(self.offsets) - 1 if self.type == T_UTF8 else len(self.values)

    def to_pylist(self):
        n = len(self)
        out = []
        for i in range(n):
            if self.valid is not None and not self.valid[i]:
                out.append(None)
            elif self.type == T_UTF8:
                out.append(bytes(self.values[self.offsets[i]:self.offsets[i + 1]]).decode())
            elif self.type == T_BOOL:
                out.append(bool(self.values[i]))
            else:
                out.append(self.values[i].item())
        return out


class Table:
    """Owning handle on an oracle stream (a concatenated table + batch sizes)."""

    def __init__(self, handle, keep=None):
        if not handle:
            raise OracleError(lib().orc_last_error().decode())
        self._h = handle
        self._keep = keep

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_table_free(self._h)
            self._h = None

    @staticmethod
    def scan(cols: Sequence[Column], batch_rows: int = 1024) -> "Table":
        n = len(cols[0]) if cols else 0
        arr = (_Col * max(1, len(cols)))()
        keep = []
        for j, c in enumerate(cols):
            assert len(c) == n
            nm = c.name.encode()
            keep.append(nm)
            arr[j] = _Col(c.type, int(c.nullable), n, c.values.ctypes.data,
                          None if c.valid is None else c.valid.ctypes.data,
                          None if c.offsets is None else c.offsets.ctypes.data, nm)
        return Table(lib().orc_scan(arr, len(cols), n, batch_rows), keep=(keep, cols))

    @property
    def nrows(self) -> int:
        return lib().orc_table_nrows(self._h)

    @property
    def nbatches(self) -> int:
        return lib().orc_table_nbatches(self._h)

    def batch_sizes(self) -> List[int]:
        return [lib().orc_table_batch_size(self._h, i) for i in range(self.nbatches)]

    def columns(self) -> List[Column]:
        out = []
        for j in range(lib().orc_table_ncols(self._h)):
            c = _Col()
            lib().orc_table_col(self._h, j, C.byref(c))
            n = c.n
            valid = None
            if c.valid:
                valid = np.ctypeslib.as_array(C.cast(c.valid, C.POINTER(C.c_uint8)), shape=(n,)).copy()
            if c.type == T_UTF8:
                off = np.ctypeslib.as_array(C.cast(c.offsets, C.POINTER(C.c_int32)), shape=(n + 1,)).copy()
                nb = int(off[-1])
                vals = (np.ctypeslib.as_array(C.cast(c.data, C.POINTER(C.c_uint8)), shape=(nb,)).copy()
                        if nb else np.zeros(0, np.uint8))
                out.append(Column(c.name.decode(), c.type, vals, valid, off, bool(c.nullable)))
            else:
                dt = np.dtype(_NP[c.type])
                if n:
                    raw = np.ctypeslib.as_array(C.cast(c.data, C.POINTER(C.c_uint8)), shape=(n * dt.itemsize,)).copy()
                    vals = raw.view(dt)
                else:
                    vals = np.zeros(0, dt)
                out.append(Column(c.name.decode(), c.type, vals, valid, None, bool(c.nullable)))
        return out

    def digest(self):
        d = (C.c_uint64 * 3)()
        lib().orc_table_digest(self._h, d)
        return tuple(int(v) for v in d)

    # operators -------------------------------------------------------------
    def filter(self, pred: bytes) -> "Table":
        """FilterExec::execute (filter.rs:39-68)."""
        return Table(lib().orc_filter(self._h, pred, len(pred)))

    def project(self, exprs: Sequence[bytes], names: Sequence[str]) -> "Table":
        """ProjectionExec::execute (projection.rs:48-72)."""
        n = len(exprs)
        e = (C.c_char_p * n)(*exprs)
        ln = (C.c_size_t * n)(*[len(x) for x in exprs])
        nm = (C.c_char_p * n)(*[s.encode() for s in names])
        return Table(lib().orc_project(self._h, e, ln, nm, n))

    def hash_join(self, right: "Table", left_keys: Sequence[bytes], right_keys: Sequence[bytes],
                  join_type: int = JOIN_INNER, mode: int = 0, threads: int = 1) -> "Table":
        """HashJoinExec::execute with self as the left (probe) child (hash_join.rs:100-213)."""
        n = len(left_keys)
        lk = (C.c_char_p * n)(*left_keys)
        ll = (C.c_size_t * n)(*[len(x) for x in left_keys])
        rk = (C.c_char_p * n)(*right_keys)
        rl = (C.c_size_t * n)(*[len(x) for x in right_keys])
        return Table(lib().orc_hash_join(self._h, right._h, lk, ll, rk, rl, n, join_type, mode, threads))


# generator ------------------------------------------------------------------
def gen_i32_uniform(seed, col, row0, n, m):
    out = np.empty(n, np.int32)
    lib().orc_gen_i32_uniform(seed, col, row0, n, m, out.ctypes.data)
    return out


def gen_i64_uniform(seed, col, row0, n, m=0):
    out = np.empty(n, np.int64)
    lib().orc_gen_i64_uniform(seed, col, row0, n, m, out.ctypes.data)
    return out


def gen_f64_unit(seed, col, row0, n):
    out = np.empty(n, np.float64)
    lib().orc_gen_f64_unit(seed, col, row0, n, out.ctypes.data)
    return out


def gen_perm_pow2(row0, n, bits):
    out = np.empty(n, np.int64)
    lib().orc_gen_perm_pow2(row0, n, bits, out.ctypes.data)
    return out


def bench_filter_project(seed, nrows, threshold):
    rows = C.c_int64(0)
    s = lib().orc_bench_filter_project(seed, nrows, threshold, C.byref(rows))
    return s, rows.value


def bench_join(seed, build_bits, nprobe, mode=0, threads=1):
    rows = C.c_int64(0)
    s = lib().orc_bench_join(seed, build_bits, nprobe, mode, threads, C.byref(rows))
    return s, rows.value
