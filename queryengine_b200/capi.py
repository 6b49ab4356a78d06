"""ctypes binding of the C ABI in include/qegpu.h (libqegpu.so, built in-tree).

This is the same binding shape a JNI stub would have: plain
pointers and sizes, integer status codes, qe_last_error() for the message.
There is no fallback: if the library is missing, importing the symbols raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libqegpu.so")

QE_OK = 0
QE_ERR_INVALID_ARGUMENT = 1
QE_ERR_PROGRAM = 2
QE_ERR_CUDA = 3
QE_ERR_UNSUPPORTED = 4
QE_ERR_CAPACITY = 5
QE_ERR_JIT = 6
QE_ERR_STATE = 7

QE_RUN_NO_SELECTION = 1
QE_RUN_INTERPRETER = 2
QE_RUN_APPEND = 4
QE_RUN_NO_VALUES = 8
QE_STREAM_CHUNK_RELATIVE = 1
QE_AGG_SUPER_ROWS = 1 << 20
QE_NOALLOC_MAX_ROWS = 1 << 28
QE_NOALLOC_SCRATCH_BYTES = 532480
QE_GROUPBY_ATOMIC = 1
QE_GROUPBY_SHARED = 2
QE_GROUPBY_TWOPASS = 4

QE_GEN_UNIFORM_U40 = 0
QE_GEN_UNIFORM_INT = 1
QE_GEN_UNIFORM_REAL = 2
QE_GEN_CENTS = 3
QE_GEN_ROW_INDEX = 4


class QeError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"qegpu status {status}: {message}")
        self.status = status
        self.message = message


class ColDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("nullable", C.c_int32)]


class AggPartial(C.Structure):
    _fields_ = [("sum", C.c_double), ("count", C.c_int64), ("min", C.c_double), ("max", C.c_double)]


class PipelineDesc(C.Structure):
    _fields_ = [
        ("filter", C.c_void_p),
        ("proj", C.POINTER(C.c_void_p)),
        ("nproj", C.c_int32),
        ("agg_fns", C.POINTER(C.c_int32)),
        ("nagg", C.c_int32),
        ("row_begin", C.c_int64),
        ("row_end", C.c_int64),
        ("out_capacity_rows", C.c_int64),
        ("flags", C.c_uint32),
        ("sel_offset", C.c_int64),
    ]


class GroupByDesc(C.Structure):
    _fields_ = [
        ("filter", C.c_void_p),
        ("keys", C.POINTER(C.c_void_p)),
        ("nkeys", C.c_int32),
        ("inputs", C.POINTER(C.c_void_p)),
        ("agg_fns", C.POINTER(C.c_int32)),
        ("nagg", C.c_int32),
        ("row_begin", C.c_int64),
        ("row_end", C.c_int64),
        ("max_groups", C.c_int64),
        ("flags", C.c_uint32),
    ]


class StreamGen(C.Structure):
    _fields_ = [("kind", C.c_int32), ("seed", C.c_uint64), ("p0", C.c_double), ("p1", C.c_double), ("null_rate", C.c_double)]


class StreamDesc(C.Structure):
    _fields_ = [
        ("chunk_rows", C.c_int64),
        ("host_values", C.POINTER(C.c_void_p)),
        ("host_validity", C.POINTER(C.c_void_p)),
        ("npipelines", C.c_int32),
        ("pipelines", C.POINTER(PipelineDesc)),
        ("results", C.POINTER(C.c_void_p)),
        ("host_selection", C.POINTER(C.c_void_p)),
        ("host_out_values", C.POINTER(C.POINTER(C.c_void_p))),
        ("host_out_validity", C.POINTER(C.POINTER(C.c_void_p))),
        ("total_rows", C.c_int64),
        ("flags", C.c_uint32),
        ("gen", C.POINTER(StreamGen)),
        ("gen_row_offset", C.c_int64),
        ("chunk_kept", C.POINTER(C.POINTER(C.c_int64))),
    ]


class GroupPart(C.Structure):
    _fields_ = [("ngroups", C.c_int64), ("keys", C.c_void_p), ("key_null", C.c_void_p), ("first_row", C.c_void_p),
                ("partials", C.c_void_p)]


class AotDesc(C.Structure):
    _fields_ = [
        ("cols", C.POINTER(ColDesc)),
        ("ncols", C.c_int32),
        ("programs", C.POINTER(C.c_void_p)),
        ("nwords", C.POINTER(C.c_size_t)),
        ("has_filter", C.c_int32),
        ("nproj", C.c_int32),
        ("agg_fns", C.POINTER(C.c_int32)),
        ("nagg", C.c_int32),
        ("group_keys", C.c_int32),
        ("sel_hint", C.c_double),
        ("min_blocks", C.c_int32),
        ("seen_groups", C.c_int32),
        ("run_flags", C.c_uint32),
    ]


# every symbol include/qegpu.h declares: name -> (restype, argtypes)
_vp = C.c_void_p
SYMBOLS = {
    "qe_ctx_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "qe_ctx_destroy": (C.c_int, [_vp]),
    "qe_ctx_synchronize": (C.c_int, [_vp]),
    "qe_ctx_stream": (_vp, [_vp]),
    "qe_ctx_set_option": (C.c_int, [_vp, C.c_char_p, C.c_int64]),
    "qe_ctx_launch_count": (C.c_int64, [_vp]),
    "qe_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(_vp)]),
    "qe_host_free": (C.c_int, [_vp]),
    "qe_program_create": (C.c_int, [_vp, _vp, C.c_size_t, C.POINTER(_vp)]),
    "qe_program_destroy": (C.c_int, [_vp]),
    "qe_table_create": (C.c_int, [_vp, C.c_int, C.POINTER(ColDesc), C.c_int64, C.POINTER(_vp)]),
    "qe_table_upload": (C.c_int, [_vp, C.c_int, _vp, _vp, C.c_int64, C.c_int64]),
    "qe_table_device_ptr": (_vp, [_vp, C.c_int, C.c_int]),
    "qe_table_generate": (C.c_int, [_vp, C.c_int, C.c_int, C.c_uint64, C.c_double, C.c_double, C.c_double, C.c_int64]),
    "qe_table_destroy": (C.c_int, [_vp]),
    "qe_pipeline_run": (C.c_int, [_vp, _vp, _vp, C.POINTER(_vp), C.c_int, C.POINTER(C.c_int32), C.c_int, C.POINTER(_vp)]),
    "qe_pipeline_run_ex": (C.c_int, [_vp, _vp, C.POINTER(PipelineDesc), C.POINTER(_vp)]),
    "qe_pipeline_stream": (C.c_int, [_vp, _vp, C.POINTER(StreamDesc)]),
    "qe_result_num_rows": (C.c_int, [_vp, C.POINTER(C.c_int64)]),
    "qe_result_num_columns": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "qe_result_column_type": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "qe_result_column": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "qe_result_selection": (C.c_int, [_vp, _vp]),
    "qe_result_agg": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]),
    "qe_result_agg_partial": (C.c_int, [_vp, C.c_int, C.POINTER(AggPartial)]),
    "qe_result_device_ptr": (_vp, [_vp, C.c_int, C.c_int]),
    "qe_result_kernel_ms": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "qe_result_kernel_stats": (C.c_int, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "qe_result_debug_counters": (C.c_int, [_vp, _vp, C.c_size_t]),
    "qe_result_used_jit": (C.c_int, [_vp]),
    "qe_result_destroy": (C.c_int, [_vp]),
    "qe_agg_combine": (C.c_int, [C.c_int, C.POINTER(AggPartial), C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]),
    "qe_debug_generate": (C.c_int, [C.POINTER(ColDesc), C.c_int, C.POINTER(_vp), C.POINTER(C.c_size_t), C.c_int, C.c_int,
                                    C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "qe_debug_generate_groupby": (C.c_int, [C.POINTER(ColDesc), C.c_int, C.POINTER(_vp), C.POINTER(C.c_size_t), C.c_int, C.c_int,
                                            C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_size_t,
                                            C.POINTER(C.c_size_t)]),
    "qe_groupby_run": (C.c_int, [_vp, _vp, C.POINTER(GroupByDesc), C.POINTER(_vp)]),
    "qe_result_group_partial": (C.c_int, [_vp, C.c_int64, C.c_int, C.POINTER(AggPartial)]),
    "qe_result_group_first_rows": (C.c_int, [_vp, _vp]),
    "qe_result_group_variant": (C.c_int, [_vp]),
    "qe_result_group_threads": (C.c_int, [_vp]),
    "qe_result_group_copies": (C.c_int, [_vp]),
    "qe_ctx_trim": (C.c_int, [_vp]),
    "qe_jit_stats": (C.c_int, [C.POINTER(C.c_double)]),
    "qe_yardstick": (C.c_int, [_vp, C.c_int, C.c_int64, C.c_int, C.POINTER(C.c_float)]),
    "qe_result_agg_partials": (C.c_int, [_vp, C.c_int, C.POINTER(AggPartial), C.c_int64, C.POINTER(C.c_int64)]),
    "qe_aot_compile": (C.c_int, [C.POINTER(AotDesc), C.c_char_p, C.POINTER(C.c_int)]),
    "qe_agg_column_noalloc": (C.c_int, [C.c_int, _vp, _vp, C.c_int64, _vp, _vp]),
    "qe_partition_rows": (C.c_int, [C.c_int64, C.c_int, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "qe_mgpu_create": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(_vp)]),
    "qe_mgpu_destroy": (C.c_int, [_vp]),
    "qe_mgpu_num_devices": (C.c_int, [_vp]),
    "qe_mgpu_ctx": (_vp, [_vp, C.c_int]),
    "qe_mgpu_set_option": (C.c_int, [_vp, C.c_char_p, C.c_int64]),
    "qe_mtable_create": (C.c_int, [_vp, C.c_int, C.POINTER(ColDesc), C.c_int64, C.POINTER(_vp)]),
    "qe_mtable_upload": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "qe_mtable_generate": (C.c_int, [_vp, C.c_int, C.c_int, C.c_uint64, C.c_double, C.c_double, C.c_double]),
    "qe_mtable_part": (_vp, [_vp, C.c_int]),
    "qe_mtable_destroy": (C.c_int, [_vp]),
    "qe_mgpu_pipeline_run": (C.c_int, [_vp, _vp, C.POINTER(PipelineDesc), C.POINTER(_vp)]),
    "qe_mgpu_groupby_run": (C.c_int, [_vp, _vp, C.POINTER(GroupByDesc), C.POINTER(_vp)]),
    "qe_mgpu_pipeline_stream": (C.c_int, [_vp, _vp, C.POINTER(StreamDesc)]),
    "qe_groupby_merge": (C.c_int, [C.POINTER(GroupPart), C.c_int, C.c_int, C.POINTER(C.c_int32), C.c_int, C.c_int64,
                                   _vp, _vp, _vp, _vp, _vp, C.POINTER(C.c_int64)]),
    "qe_mresult_num_rows": (C.c_int, [_vp, C.POINTER(C.c_int64)]),
    "qe_mresult_num_columns": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "qe_mresult_part": (_vp, [_vp, C.c_int]),
    "qe_mresult_part_rows": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_int64)]),
    "qe_mresult_column": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "qe_mresult_selection": (C.c_int, [_vp, _vp]),
    "qe_mresult_agg": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]),
    "qe_mresult_destroy": (C.c_int, [_vp]),
    "qe_last_error": (C.c_char_p, []),
    "qe_version": (C.c_char_p, []),
}

_lib = None


def lib():
    """Load libqegpu.so.  Raises if it has not been built (python -m queryengine_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} is missing: build the CUDA extension with `python -m queryengine_b200.build` "
                "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


def check(status: int) -> None:
    if status != QE_OK:
        raise QeError(status, lib().qe_last_error().decode(errors="replace"))
