"""ctypes binding of libprestogpu.so (include/prestogpu.h).  There is no CPU fallback: if the library is
missing or no CUDA device is present, loading / context creation raises."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libprestogpu.so")

# ---- enums (include/prestogpu.h)
PG_OK, PG_NEED_MEMORY, PG_E_INSUFFICIENT_RESOURCES, PG_E_NUMERIC_OVERFLOW = 0, 1, 2, 3
PG_E_INVALID_ARGUMENT, PG_E_CUDA, PG_E_NCCL, PG_E_INTERNAL, PG_E_ILLEGAL_STATE = 4, 5, 6, 7, 8
BIGINT, INTEGER, DATE, DOUBLE, SHORT_DECIMAL, LONG_DECIMAL, VARCHAR, VARBINARY, BOOLEAN = range(1, 10)
MEM_HOST, MEM_DEVICE, MEM_HOST_PINNED = 0, 1, 2
AGG_COUNT, AGG_COUNT_COL, AGG_SUM, AGG_AVG, AGG_MIN, AGG_MAX = range(1, 7)
JOIN_INNER, JOIN_PROBE_OUTER = 0, 1
NOALLOC_MAX_ROWS = 4194304      # PG_NOALLOC_MAX_ROWS

STATUS_NAMES = {
    1: "NEED_MEMORY", 2: "GENERIC_INSUFFICIENT_RESOURCES", 3: "NUMERIC_VALUE_OUT_OF_RANGE", 4: "INVALID_ARGUMENT",
    5: "CUDA_ERROR", 6: "NCCL_ERROR", 7: "INTERNAL_ERROR", 8: "ILLEGAL_STATE",
}


class PgCol(C.Structure):
    _fields_ = [("type", C.c_int32), ("mem", C.c_int32), ("n", C.c_int64), ("values", C.c_void_p), ("nulls", C.c_void_p),
                ("offsets", C.c_void_p), ("precision", C.c_int32), ("scale", C.c_int32)]


class PgColOut(C.Structure):
    _fields_ = [("type", C.c_int32), ("mem", C.c_int32), ("capacity", C.c_int64), ("values", C.c_void_p), ("nulls", C.c_void_p),
                ("offsets", C.c_void_p), ("values_capacity_bytes", C.c_int64)]


class PgExpr(C.Structure):
    _fields_ = [("op", C.c_int32), ("type", C.c_int32), ("precision", C.c_int32), ("scale", C.c_int32), ("channel", C.c_int32),
                ("arg0", C.c_int32), ("arg1", C.c_int32), ("is_null", C.c_int32), ("lo", C.c_int64), ("hi", C.c_int64),
                ("bytes", C.c_void_p), ("nbytes", C.c_int32), ("reserved", C.c_int32)]


(EXPR_INPUT, EXPR_CONST, EXPR_ADD, EXPR_SUB, EXPR_MUL, EXPR_LT, EXPR_LE, EXPR_GT, EXPR_GE, EXPR_EQ, EXPR_NE, EXPR_AND, EXPR_OR, EXPR_NOT,
 EXPR_IS_NULL, EXPR_CAST, EXPR_HASH) = range(1, 18)


class PgAggSpec(C.Structure):
    _fields_ = [("kind", C.c_int32), ("in_type", C.c_int32), ("input_channel", C.c_int32), ("mask_channel", C.c_int32)]


STEP_SINGLE, STEP_PARTIAL, STEP_FINAL, STEP_INTERMEDIATE = 0, 1, 2, 3


class PrestoGpuError(RuntimeError):
    """Mirrors TrinoException(StandardErrorCode, message) (SURVEY.md §8b error conventions)."""

    def __init__(self, code, message):
        super().__init__("%s: %s" % (STATUS_NAMES.get(code, str(code)), message))
        self.code = code
        self.error_name = STATUS_NAMES.get(code, str(code))


_lib = None

P = C.c_void_p
I32, I64 = C.c_int32, C.c_int64
PI32, PI64 = C.POINTER(C.c_int32), C.POINTER(C.c_int64)
PCOL, PCOLOUT = C.POINTER(PgCol), C.POINTER(PgColOut)

_PROTOS = {
    "pg_ctx_create": (I32, [I32, C.POINTER(P)]),
    "pg_ctx_destroy": (I32, [P]),
    "pg_last_error": (C.c_char_p, [P]),
    "pg_ctx_set_stream": (I32, [P, P]),
    "pg_ctx_reserve_sms": (I32, [P, I32]),
    "pg_ctx_synchronize": (I32, [P]),
    "pg_ctx_launch_count": (I64, [P]),
    "pg_version": (C.c_char_p, []),
    "pg_ctx_set_timing": (I32, [P, I32]),
    "pg_ctx_kernel_time": (I32, [P, C.c_char_p, I32, C.POINTER(C.c_double), PI64]),
    "pg_ctx_kernel_report": (I32, [P, C.c_char_p, I64]),
    "pg_groupby_create": (I32, [P, PI32, I32, I32, I64, C.POINTER(P)]),
    "pg_groupby_destroy": (I32, [P]),
    "pg_groupby_get_group_ids": (I32, [P, PCOL, PCOL, I64, P, I32, PI64]),
    "pg_groupby_add_page": (I32, [P, PCOL, PCOL, I64]),
    "pg_groupby_contains": (I32, [P, PCOL, PCOL, I64, P, I32]),
    "pg_groupby_export_keys": (I32, [P, I64, I64, PCOLOUT, P, I32]),
    "pg_groupby_varwidth_bytes": (I32, [P, I32, PI64]),
    "pg_groupby_stats": (I32, [P, PI64, PI64, PI64, PI64]),
    "pg_groupby_set_memory_limit": (I32, [P, I64]),
    "pg_groupby_mark_distinct": (I32, [P, PCOL, PCOL, I64, P, I32]),
    "pg_groupby_semi_join": (I32, [P, PCOL, PCOL, I64, I32, P, P, I32]),
    "pg_agg_create": (I32, [P, I32, I32, C.POINTER(P)]),
    "pg_agg_destroy": (I32, [P]),
    "pg_agg_add_input": (I32, [P, P, I32, I64, PCOL, PCOL, I64]),
    "pg_agg_add_intermediate": (I32, [P, P, I32, I64, PCOL, I64]),
    "pg_agg_n_intermediate": (I32, [P, PI32]),
    "pg_agg_evaluate": (I32, [P, I64, I64, I32, PCOLOUT, PI64]),
    "pg_agg_device_bytes": (I32, [P, PI64]),
    "pg_hashagg_create": (I32, [P, PI32, PI32, I32, I32, C.POINTER(PgAggSpec), I32, I32, I64, C.POINTER(P)]),
    "pg_hashagg_destroy": (I32, [P]),
    "pg_hashagg_add_page": (I32, [P, PCOL, I32, I64]),
    "pg_hashagg_stats": (I32, [P, PI64, PI64, PI64, PI64]),
    "pg_hashagg_output_types": (I32, [P, PI32, I32, PI32]),
    "pg_hashagg_get_output": (I32, [P, I64, I64, PCOLOUT, PI64]),
    "pg_hashagg_varwidth_bytes": (I32, [P, I32, PI64]),
    "pg_hashagg_reset": (I32, [P]),
    "pg_hashagg_set_memory_limit": (I32, [P, I64]),
    "pg_hashagg_set_page_batching": (I32, [P, I64]),
    "pg_hashagg_add_pages": (I32, [P, PCOL, I32, I64, I64]),
    "pg_hashagg_get_output_packed": (I32, [P, I64, I64, P, I64, PI64, PI64]),
    "pg_hashagg_add_packed": (I32, [P, C.POINTER(C.c_void_p), I32]),
    "pg_hashagg_get_output_packed_device": (I32, [P, I64, I64, P, I64, PI64, PI64]),
    "pg_hashagg_add_packed_device": (I32, [P, P, I64, I32]),
    "pg_project_create": (I32, [P, C.POINTER(PgExpr), I32, I32, PI32, I32, C.POINTER(P)]),
    "pg_project_destroy": (I32, [P]),
    "pg_project_output_types": (I32, [P, PI32, PI32, PI32, I32, PI32]),
    "pg_project_process": (I32, [P, PCOL, I32, I64, PCOLOUT, PI64]),
    "pg_hashagg_set_input_processor": (I32, [P, P]),
    "pg_partition_create": (I32, [P, PCOL, I32, PCOL, I64, I32, I32, PI64, C.POINTER(P)]),
    "pg_partition_destroy": (I32, [P]),
    "pg_partition_scatter": (I32, [P, PCOL, PCOLOUT]),
    "pg_partition_scatter_peers": (I32, [P, PCOL, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), PI64]),
    "pg_partition_scatter_peers_all": (I32, [P, PCOL, I32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), PI64]),
    "pg_partition_ids": (I32, [P, P, I32]),
    "pg_partition_positions_noalloc": (I32, [P, P, I64, I32, I32, P, P, P]),
    "pg_joinbuild_create": (I32, [P, PI32, I32, PI32, I32, I32, I64, C.POINTER(P)]),
    "pg_joinbuild_destroy": (I32, [P]),
    "pg_joinbuild_add_page": (I32, [P, PCOL, I64]),
    "pg_joinbuild_finish": (I32, [P, C.POINTER(P)]),
    "pg_lookup_destroy": (I32, [P]),
    "pg_lookup_stats": (I32, [P, PI64, PI64, PI64, PI32]),
    "pg_lookup_probe": (I32, [P, PCOL, I32, PCOL, I64, I32, P, P, I32, I64, PI64]),
    "pg_lookup_gather": (I32, [P, P, I32, I64, PI32, I32, PCOLOUT]),
    "pg_lookup_track_visited": (I32, [P]),
    "pg_lookup_outer_positions": (I32, [P, P, I32, I64, PI64]),
    "pg_lookup_join_page": (I32, [P, PCOL, I32, PCOL, I64, I32, PI32, I32, P, I32, PCOLOUT, I64, PI64]),
    "pg_lookup_gather_var_bytes": (I32, [P, P, I32, I64, I32, PI64]),
}


def declared_symbols():
    """Every PG_API symbol declared in include/prestogpu.h (parsed from the header itself)."""
    import re
    hdr = os.path.join(HERE, "..", "include", "prestogpu.h")
    text = open(hdr).read()
    return sorted(set(re.findall(r"PG_API[^;(]*?\b(pg_[a-z0-9_]+)\s*\(", text)))


def lib():
    """Load libprestogpu.so (once).  Raises if it has not been built: the product has no CPU path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libprestogpu.so is not built (run `python presto-1_b200/build.py`); there is no CPU fallback")
    l = C.CDLL(LIB_PATH)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(l, name)
        fn.restype = res
        fn.argtypes = args
    _lib = l
    return l


def check(ctx_handle, rc):
    if rc != PG_OK:
        msg = lib().pg_last_error(ctx_handle)
        raise PrestoGpuError(rc, msg.decode() if msg else "")
