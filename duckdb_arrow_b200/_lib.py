"""ctypes binding of include/b2arrow.h.  The library is REQUIRED: import fails loudly when it is missing
and no code path in this package computes anything on the CPU."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# B2A_LIB: another build of the SAME library (kernel A/B experiments); there is still no fallback if it is missing
LIB_PATH = os.environ.get("B2A_LIB") or os.path.join(HERE, "libb2arrow.so")

B2A_NONE = 0xFFFFFFFFFFFFFFFF
VECTOR_SIZE = 2048
SCAN_HOST_ALIAS = 0x1
SCAN_DEVICE_ALIAS = 0x2

STATUS = {0: "OK", 1: "InvalidInputException", 2: "IOException", 3: "NotImplementedException",
          4: "ConversionException", 5: "CudaError", 6: "InternalException", 7: "OutOfMemoryException"}

TYPE_NAMES = ["INVALID", "SQLNULL", "BOOLEAN", "TINYINT", "SMALLINT", "INTEGER", "BIGINT", "UTINYINT", "USMALLINT",
              "UINTEGER", "UBIGINT", "HUGEINT", "UHUGEINT", "FLOAT", "DOUBLE", "DECIMAL", "VARCHAR", "BLOB", "DATE",
              "TIME", "TIMESTAMP", "TIMESTAMP_S", "TIMESTAMP_MS", "TIMESTAMP_NS", "TIMESTAMP WITH TIME ZONE",
              "INTERVAL", "UUID", "LIST", "STRUCT", "MAP", "ARRAY", "UNION", "TIME_TZ", "BIT", "VARINT", "JSON", "ENUM"]
T = {n: i for i, n in enumerate(TYPE_NAMES)}

OP_NAMES = ["NONE", "COPY", "BOOL", "STRING32", "STRING64", "STRING_VIEW", "BLOB_FIXED", "LIST32", "LIST64",
            "MUL_I32", "MUL_I64", "DIV_I64", "DATE64", "INTERVAL_MUL", "INTERVAL_DIV", "INTERVAL_MONTHS",
            "INTERVAL_MDN", "DECIMAL", "UUID", "DICT_INDEX", "VALIDITY", "REE", "UNION", "LISTVIEW32", "LISTVIEW64", "BOOL8"]


class B2AError(Exception):
    """Mirrors the DuckDB exception class the reference would have thrown (see b2a_status)."""

    def __init__(self, code: int, message: str):
        self.code = code
        self.kind = STATUS.get(code, "Error")
        super().__init__(f"{self.kind}: {message}")


class FieldInfo(C.Structure):
    _fields_ = [("name", C.c_char * 128), ("format", C.c_char * 48), ("type_id", C.c_int32), ("op", C.c_int32),
                ("out_width", C.c_int32), ("src_width", C.c_int32), ("n_children", C.c_int32),
                ("first_child", C.c_int32), ("next_sibling", C.c_int32), ("parent", C.c_int32),
                ("nullable", C.c_int32), ("n_buffers", C.c_int32), ("decimal_width", C.c_int32),
                ("decimal_scale", C.c_int32), ("param", C.c_int64), ("dict_id", C.c_int64),
                ("index_type_id", C.c_int32), ("index_width", C.c_int32)]


class IpcBuffer(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("size", C.c_uint64)]


class BatchInfo(C.Structure):
    _fields_ = [("rows", C.c_uint64), ("body_offset", C.c_uint64), ("body_length", C.c_uint64),
                ("meta_offset", C.c_uint64), ("buffer_bytes", C.c_uint64), ("compressed", C.c_uint32),
                ("pad", C.c_uint32)]


class ColumnOut(C.Structure):
    _fields_ = [("rows", C.c_uint64), ("null_count", C.c_uint64), ("data_off", C.c_uint64),
                ("validity_off", C.c_uint64), ("aux_off", C.c_uint64), ("chars_off", C.c_uint64),
                ("chars_len", C.c_uint64), ("stats_off", C.c_uint64), ("image_off", C.c_uint64)]


class Predicate(C.Structure):
    _fields_ = [("column", C.c_uint32), ("cmp", C.c_uint32), ("constant", C.c_int64)]


class Vector(C.Structure):
    pass


Vector._fields_ = [("type_id", C.c_int32), ("decimal_width", C.c_int32), ("decimal_scale", C.c_int32),
                   ("n_children", C.c_int32), ("param", C.c_int64), ("name", C.c_char_p), ("timezone", C.c_char_p),
                   ("data", C.c_void_p), ("validity", C.c_void_p), ("rows", C.c_uint64),
                   ("children", C.POINTER(Vector))]


class WriteOptions(C.Structure):
    _fields_ = [("row_group_size", C.c_uint64), ("n_kv_metadata", C.c_uint32), ("kv_keys", C.POINTER(C.c_char_p)),
                ("kv_values", C.POINTER(C.c_char_p)), ("flags", C.c_uint32)]


WRITE_LARGE_OFFSETS = 0x1  # B2A_WRITE_LARGE_OFFSETS
WRITE_LIST_VIEW = 0x2  # B2A_WRITE_LIST_VIEW
WRITE_STRING_VIEW = 0x4  # B2A_WRITE_STRING_VIEW
WRITE_LOSSLESS = 0x8  # B2A_WRITE_LOSSLESS


class HostStream(C.Structure):
    _fields_ = [("data", C.c_void_p), ("size", C.c_uint64), ("n_batches", C.c_uint64), ("writer", C.c_void_p),
                ("gpu_launches", C.c_uint64), ("priv", C.c_void_p)]


class MultiFileOptions(C.Structure):
    _fields_ = [("union_by_name", C.c_uint32), ("filename", C.c_uint32), ("hive_partitioning", C.c_uint32),
                ("hive_types_autocast", C.c_uint32), ("filename_column", C.c_char_p)]


class MfColumnInfo(C.Structure):
    _fields_ = [("name", C.c_char * 128), ("type_name", C.c_char * 256), ("type_id", C.c_int32),
                ("decimal_width", C.c_int32), ("decimal_scale", C.c_int32), ("kind", C.c_int32)]


class MfMapInfo(C.Structure):
    _fields_ = [("kind", C.c_int32), ("local_flat", C.c_int32), ("from_type_id", C.c_int32), ("to_type_id", C.c_int32),
                ("constant_is_null", C.c_int32), ("pad", C.c_int32), ("constant", C.c_char_p)]


MF_MAP_LOCAL, MF_MAP_CAST, MF_MAP_NULL, MF_MAP_CONSTANT = 0, 1, 2, 3
MF_COL_DATA, MF_COL_FILENAME, MF_COL_HIVE = 0, 1, 2

CMP = {"=": 0, "==": 0, "!=": 1, "<>": 1, "<": 2, "<=": 3, ">": 4, ">=": 5}


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m duckdb_arrow_b200.build` (needs nvcc). "
            "There is no CPU fallback for the Arrow IPC scan path.")
    lib = C.CDLL(LIB_PATH)
    vp, u64, u32, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
    P = C.POINTER
    sig = {
        "b2a_version": (C.c_char_p, []),
        "b2a_last_error": (C.c_char_p, []),
        "b2a_device_count": (i32, [P(i32)]),
        "b2a_reader_open_file": (i32, [C.c_char_p, P(vp)]),
        "b2a_reader_open_buffers": (i32, [P(IpcBuffer), u32, P(vp)]),
        "b2a_reader_close": (None, [vp]),
        "b2a_reader_num_fields": (i32, [vp, P(u32), P(u32)]),
        "b2a_reader_field": (i32, [vp, u32, P(FieldInfo)]),
        "b2a_reader_set_projection": (i32, [vp, P(C.c_char_p), u32]),
        "b2a_reader_num_batches": (i32, [vp, P(u64), P(u64)]),
        "b2a_reader_batch_info": (i32, [vp, u64, P(BatchInfo)]),
        "b2a_reader_image": (i32, [vp, P(vp), P(u64)]),
        "b2a_reader_schema_metadata": (i32, [vp, u32, P(C.c_char_p), P(C.c_char_p), P(u32)]),
        "b2a_scan_create": (i32, [vp, i32, u64, u64, P(vp)]),
        "b2a_scan_create_ex": (i32, [vp, i32, u64, u64, u32, P(vp)]),
        "b2a_scan_upload_bytes": (i32, [vp, P(u64)]),
        "b2a_scan_destroy": (None, [vp]),
        "b2a_scan_sizes": (i32, [vp, P(u64), P(u64), P(u64), P(u64), P(u64), P(u64)]),
        "b2a_scan_upload": (i32, [vp, vp]),
        "b2a_scan_set_device_image": (i32, [vp, vp]),
        "b2a_scan_decode": (i32, [vp, vp, u64, vp]),
        "b2a_scan_finish": (i32, [vp, vp]),
        "b2a_scan_download": (i32, [vp, vp, vp, vp]),
        "b2a_scan_column": (i32, [vp, u64, u32, P(ColumnOut)]),
        "b2a_scan_dictionary": (i32, [vp, u32, P(ColumnOut), P(u32)]),
        "b2a_scan_gpu_launches": (u64, [vp]),
        "b2a_scan_scratch": (i32, [vp, P(u64), P(u64), P(u64)]),
        "b2a_scan_download_scratch": (i32, [vp, vp, vp]),
        "b2a_scan_num_chunks": (i32, [vp, P(u64)]),
        "b2a_scan_filter": (i32, [vp, vp, P(Predicate), u32, vp, vp, vp, vp]),
        "b2a_scan_filter_no_alloc": (i32, [vp, vp, P(Predicate), u32, vp, vp, vp, vp]),
        "b2a_scan_sum_product": (i32, [vp, vp, u32, u32, vp, vp, vp, vp]),
        "b2a_scan_filter_sum_product": (i32, [vp, vp, P(Predicate), u32, u32, u32, vp, vp, vp]),
        "b2a_reader_used_footer": (i32, [vp, P(i32)]),
        "b2a_read_arrow_host": (i32, [vp, vp, u32]),
        "b2a_read_arrow_device": (i32, [vp, vp, u32, vp]),
        "b2a_read_arrow_host_scratch": (i32, [vp, vp, vp, u32]),
        "b2a_write_arrow_host": (i32, [P(Vector), u32, u64, P(WriteOptions), i32, P(HostStream)]),
        "b2a_host_stream_release": (None, [P(HostStream)]),
        "b2a_pinned_alloc": (i32, [P(vp), u64]),
        "b2a_pinned_free": (None, [vp]),
        "b2a_multifile_open": (i32, [P(C.c_char_p), u32, P(MultiFileOptions), P(vp)]),
        "b2a_multifile_close": (None, [vp]),
        "b2a_multifile_counts": (i32, [vp, P(u32), P(u32), P(u32)]),
        "b2a_multifile_column": (i32, [vp, u32, P(MfColumnInfo)]),
        "b2a_multifile_set_projection": (i32, [vp, P(C.c_char_p), u32]),
        "b2a_multifile_projected": (i32, [vp, u32, P(u32)]),
        "b2a_multifile_file": (i32, [vp, u32, P(C.c_char_p), P(vp)]),
        "b2a_multifile_map": (i32, [vp, u32, u32, P(MfMapInfo)]),
        "b2a_multifile_finalize_create": (i32, [vp, u32, vp, P(vp)]),
        "b2a_multifile_finalize_destroy": (None, [vp]),
        "b2a_multifile_finalize_sizes": (i32, [vp, P(u64), P(u64)]),
        "b2a_multifile_finalize_run": (i32, [vp, vp, vp, u64, u64, vp]),
        "b2a_multifile_finalize_finish": (i32, [vp, vp]),
        "b2a_multifile_finalize_column": (i32, [vp, u64, u32, P(ColumnOut)]),
        "b2a_multifile_finalize_gpu_launches": (u64, [vp]),
        "b2a_writer_create": (i32, [P(Vector), u32, u64, P(WriteOptions), i32, P(vp)]),
        "b2a_writer_destroy": (None, [vp]),
        "b2a_writer_measure": (i32, [vp, vp]),
        "b2a_writer_sizes": (i32, [vp, P(u64), P(u64), P(u64), P(u64)]),
        "b2a_writer_encode": (i32, [vp, vp, vp]),
        "b2a_writer_finish": (i32, [vp, vp]),
        "b2a_writer_headers": (i32, [vp, vp]),
        "b2a_writer_batch": (i32, [vp, u64, P(BatchInfo)]),
        "b2a_writer_gpu_launches": (u64, [vp]),
        "b2a_writer_schema_message": (i32, [vp, P(vp), P(u64)]),
        "b2a_writer_preamble": (i32, [vp, P(u64)]),
    }
    for name, (res, args) in sig.items():
        if not hasattr(lib, name):
            continue  # test_abi.py checks every declared symbol; optional here so partial builds still import
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(code: int) -> None:
    if code != 0:
        raise B2AError(code, lib.b2a_last_error().decode("utf-8", "replace"))
