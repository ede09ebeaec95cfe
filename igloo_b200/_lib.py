"""ctypes binding of libigloo_cuda.so (include/igloo_cuda.h).

The library is the product: if it is missing or does not load, importing this
module raises — there is no Python / CPU fallback path.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# IGLOO_CUDA_LIB: load another build of the same library (kernel A/B experiments)
LIB_PATH = os.environ.get("IGLOO_CUDA_LIB") or os.path.join(_HERE, "libigloo_cuda.so")

IGL_OK, IGL_AGAIN, IGL_EOS = 0, 1, 2
ERR_NAMES = {-1: "INVALID_ARG", -2: "UNSUPPORTED", -3: "OOM", -4: "CUDA", -5: "NCCL", -6: "DIV_BY_ZERO",
             -7: "PREDICATE_NOT_BOOLEAN", -8: "ARITH_OVERFLOW", -9: "CAST", -10: "INTERNAL"}


class IglooCudaError(RuntimeError):
    """DataFusionError analogue: `code` is the igl_status error class."""

    def __init__(self, code: int, msg: str):
        super().__init__(f"[{ERR_NAMES.get(code, code)}] {msg}")
        self.code = code
        self.message = msg


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("flags", C.c_int32), ("coalesce_rows", C.c_int64),
                ("hash_load_factor", C.c_double), ("pinned_ring_bytes", C.c_int64)]


class DevCol(C.Structure):
    _fields_ = [("type", C.c_int32), ("reserved", C.c_int32), ("data", C.c_void_p),
                ("validity", C.c_void_p), ("offsets", C.c_void_p)]


class ArrowSchema(C.Structure):
    pass


class ArrowArray(C.Structure):
    pass


ArrowSchema._fields_ = [("format", C.c_char_p), ("name", C.c_char_p), ("metadata", C.c_char_p),
                        ("flags", C.c_int64), ("n_children", C.c_int64),
                        ("children", C.POINTER(C.POINTER(ArrowSchema))), ("dictionary", C.POINTER(ArrowSchema)),
                        ("release", C.c_void_p), ("private_data", C.c_void_p)]
ArrowArray._fields_ = [("length", C.c_int64), ("null_count", C.c_int64), ("offset", C.c_int64),
                       ("n_buffers", C.c_int64), ("n_children", C.c_int64), ("buffers", C.POINTER(C.c_void_p)),
                       ("children", C.POINTER(C.POINTER(ArrowArray))), ("dictionary", C.POINTER(ArrowArray)),
                       ("release", C.c_void_p), ("private_data", C.c_void_p)]

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(igloo_b200 has no CPU fallback)")

lib = C.CDLL(LIB_PATH)

_vp, _i32, _i64, _u64, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_size_t
_P = C.POINTER


def _sig(name, restype, *argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = list(argtypes)
    return fn


_sig("igl_last_error", C.c_char_p)
_sig("igl_version", _i32)
_sig("igl_ctx_create", _i32, _P(Config), _P(_vp))
_sig("igl_ctx_destroy", None, _vp)
_sig("igl_ctx_device", _i32, _vp)
_sig("igl_ctx_launch_count", _i64, _vp)
_sig("igl_dev_alloc", _i32, _vp, _sz, _P(_vp))
_sig("igl_dev_free", _i32, _vp, _vp)
_sig("igl_host_alloc_pinned", _i32, _vp, _sz, _P(_vp))
_sig("igl_host_free_pinned", _i32, _vp, _vp)
_sig("igl_memcpy_h2d", _i32, _vp, _vp, _vp, _sz, _vp)
_sig("igl_memcpy_d2h", _i32, _vp, _vp, _vp, _sz, _vp)
_sig("igl_stream_sync", _i32, _vp, _vp)
_sig("igl_expr_compile", _i32, _vp, C.c_char_p, _sz, _P(_i32), C.c_char_p, _i32, _P(_vp))
_sig("igl_expr_type", _i32, _vp, _P(_i32), _P(_i32))
_sig("igl_expr_destroy", None, _vp)
_sig("igl_fp_create", _i32, _vp, _P(_i32), C.c_char_p, _i32, _vp, _P(_vp), _i32, _P(_vp))
_sig("igl_fp_num_outputs", _i32, _vp)
_sig("igl_fp_output_type", _i32, _vp, _i32, _P(_i32), _P(_i32))
_sig("igl_fp_run", _i32, _vp, _P(DevCol), _i64, _P(DevCol), _P(_i64), _vp)
_sig("igl_fp_launch", _i32, _vp, _P(DevCol), _i64, _P(DevCol), _vp)
_sig("igl_fp_finish", _i32, _vp, _P(_i64), _vp)
_sig("igl_fp_destroy", None, _vp)
_sig("igl_fp_is_specialised", _i32, _vp)
_sig("igl_fp_prepare", _i32, _vp)
_sig("igl_fp_jit_compile_only", _i32, _vp, _P(_i64))
_sig("igl_jit_selftest", _i32, _P(_i64))
_sig("igl_fp_run_static", _i32, _vp, _P(DevCol), _i64, _P(DevCol), _P(_i64), _vp)
_sig("igl_join_create", _i32, _vp, _i32, _P(_i32), C.c_char_p, _i32, _P(_i32), C.c_char_p, _i32, _vp, _vp, _P(_vp))
_sig("igl_join_create_ex", _i32, _vp, _i32, _P(_i32), C.c_char_p, _i32, _P(_i32), C.c_char_p, _i32, _P(_vp), _P(_vp), _i32, _P(_vp))
_sig("igl_join_build", _i32, _vp, _P(DevCol), _i64, _vp)
_sig("igl_join_probe_count", _i32, _vp, _P(DevCol), _i64, _P(_i64), _vp)
_sig("igl_join_probe", _i32, _vp, _P(DevCol), _i64, _P(DevCol), _i64, _P(_i64), _vp)
_sig("igl_join_has_duplicates", _i32, _vp)
_sig("igl_join_partition_bits", _i32, _vp, _i64)
_sig("igl_join_phase_ms", _i32, _vp, _P(C.c_double))
_sig("igl_join_set_prepartitioned", _i32, _vp, _i32)
_sig("igl_join_destroy", None, _vp)
_sig("igl_partition_histogram", _i32, _vp, _P(DevCol), _i64, _i32, _vp, _vp)
_sig("igl_partition_scatter", _i32, _vp, _P(DevCol), _i32, _i32, _i64, _i32, _P(_vp), _P(_i64), _vp)
_sig("igl_partition_radix_supported", _i32, _P(DevCol), _i32)
_sig("igl_partition_histogram_radix", _i32, _vp, _P(DevCol), _i64, _i32, _i32, _vp, _vp)
_sig("igl_partition_scatter_radix", _i32, _vp, _P(DevCol), _i32, _i32, _i64, _i32, _i32, _P(_vp), _P(_i64), _vp)
_sig("igl_ipc_export", _i32, _vp, _vp, C.c_char_p)
_sig("igl_ipc_open", _i32, _vp, C.c_char_p, _P(_vp))
_sig("igl_ipc_close", _i32, _vp, _vp)
_sig("igl_comm_unique_id", _i32, C.c_char_p)
_sig("igl_comm_init", _i32, _vp, _i32, _i32, C.c_char_p)
_sig("igl_comm_rank", _i32, _vp, _P(_i32), _P(_i32))
_sig("igl_join_shuffle_create", _i32, _vp, _i32, _P(_i32), C.c_char_p, _i32, _P(_i32), C.c_char_p, _i32, _vp, _vp, _i32, _P(_vp))
_sig("igl_join_shuffle_run", _i32, _vp, _P(DevCol), _i64, _P(DevCol), _i64, _P(_i64), _vp)
_sig("igl_join_shuffle_output", _i32, _vp, _P(DevCol), _P(_i64))
_sig("igl_join_shuffle_phase_ms", _i32, _vp, _P(C.c_double))
_sig("igl_join_shuffle_stats", _i32, _vp, _P(_i64))
_sig("igl_join_shuffle_destroy", None, _vp)
_sig("igl_shuffle_plan", _i32, _i32, _i32, _i32, _P(_i64), _P(_i64), _P(_i32), _P(_i64), _P(_i64), _P(_i64))
_sig("igl_shuffle_plan_ex", _i32, _i32, _i32, _i32, _i32, _P(_i64), _P(_i64), _P(_i32), _P(_i64), _P(_i64), _P(_i64))
_sig("igl_gen_column", _i32, _vp, _i32, _u64, _u64, _i64, _i64, _u64, _i32, _vp, _vp, _vp)
_sig("igl_digest_rows", _i32, _vp, _P(DevCol), _i32, _i64, _P(_u64), _vp)
_sig("igl_op_filter_project_create", _i32, _vp, _P(ArrowSchema), C.c_char_p, _sz, _P(C.c_char_p), _P(_sz),
     _P(C.c_char_p), _i32, _P(_vp))
_sig("igl_op_hash_join_create", _i32, _vp, _i32, _P(ArrowSchema), _P(ArrowSchema), C.c_char_p, _sz,
     C.c_char_p, _sz, _P(_vp))
_sig("igl_op_hash_join_create_ex", _i32, _vp, _i32, _P(ArrowSchema), _P(ArrowSchema), _P(C.c_char_p), _P(_sz),
     _P(C.c_char_p), _P(_sz), _i32, _P(_vp))
_sig("igl_op_schema", _i32, _vp, _P(ArrowSchema))
_sig("igl_op_push_build", _i32, _vp, _P(ArrowArray))
_sig("igl_op_finish_build", _i32, _vp)
_sig("igl_op_push", _i32, _vp, _P(ArrowArray))
_sig("igl_op_finish_input", _i32, _vp)
_sig("igl_op_pull", _i32, _vp, _P(ArrowArray))
_sig("igl_op_destroy", None, _vp)


def check(rc: int) -> int:
    if rc < 0:
        raise IglooCudaError(rc, lib.igl_last_error().decode(errors="replace"))
    return rc
