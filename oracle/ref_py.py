"""Python driver of oracle/_ref/libref_oracle.so — the reference's OWN converters (DuckDB compiled from
/root/reference/duckdb by oracle/build_ref.py).  TEST INFRASTRUCTURE ONLY: imported by tests/, bench.py's CPU legs and
tools/ that generate fixtures / bench inputs; never by the product.

  scan_column(arr)            reference arrow_scan of one column -> Slab (same form as oracle_py.scan_column)
  scan_stream(reader)         timed CPU baseline: reference scan of a whole stream, all columns
  append(batch, flags)        reference ArrowAppender over the reference-scanned rows -> pyarrow StructArray (raw buffers)
  register(name, reader), query(sql)     DuckDB SQL over arrow_scan (Q6 of benchmark/lineitem.py:41)
  export(sql, batch_rows)     SELECT -> record batches cut like COPY ... (FORMAT ARROWS) cuts them
"""
from __future__ import annotations

import ctypes as C
import os

import pyarrow as pa

from . import oracle_py

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_ref", "libref_oracle.so")

FLAG_LARGE, FLAG_LIST_VIEW, FLAG_STRING_VIEW, FLAG_LOSSLESS = 1, 2, 4, 8

_lib = None


class RefError(Exception):
    def __init__(self, code, msg):
        self.code = code
        self.kind = msg.split(":")[0].strip() if ":" in msg else "?"
        super().__init__(msg)


def available() -> bool:
    return os.path.exists(LIB)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref is not built (python oracle/build_ref.py; needs /root/reference)")
        _lib = C.CDLL(LIB)
        _lib.ref_open.argtypes = [C.c_char_p, C.c_int]
        _lib.ref_scan_column.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.POINTER(oracle_py.OSlab)), C.c_char_p,
                                         C.c_int]
        _lib.ref_slab_free.argtypes = [C.POINTER(oracle_py.OSlab)]
        _lib.ref_scan_stream.restype = C.c_int64
        _lib.ref_scan_stream.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_char_p, C.c_int]
        _lib.ref_append.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
        _lib.ref_register_stream.argtypes = [C.c_char_p, C.c_void_p, C.c_char_p, C.c_int]
        _lib.ref_query.argtypes = [C.c_char_p, C.POINTER(C.c_char_p), C.c_char_p, C.c_int]
        _lib.ref_set_threads.argtypes = [C.c_int, C.c_char_p, C.c_int]
        _lib.ref_export_begin.restype = C.c_void_p
        _lib.ref_export_begin.argtypes = [C.c_char_p, C.c_uint32, C.c_void_p, C.c_char_p, C.c_int]
        _lib.ref_export_next.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_char_p, C.c_int]
        _lib.ref_export_end.argtypes = [C.c_void_p]
        _lib.ref_source_id.restype = C.c_char_p
        _lib.ref_file_open.restype = C.c_void_p
        _lib.ref_file_open.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        _lib.ref_file_read.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_char_p, C.c_int]
        _lib.ref_file_close.argtypes = [C.c_void_p]
        _lib.ref_load_extension.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int]
        _lib.ref_scan_arrow_ipc_view.argtypes = [C.c_char_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_uint32,
                                                 C.c_char_p, C.c_int]
        err = C.create_string_buffer(512)
        if _lib.ref_open(err, 512) != 0:
            raise RefError(4, err.value.decode())
    return _lib


_loaded_extensions = set()
_streams = []  # exported ArrowArrayStream structs referenced by registered views


def _export_stream(reader: pa.RecordBatchReader):
    mem = C.create_string_buffer(64)  # struct ArrowArrayStream: 5 pointers
    reader._export_to_c(C.addressof(mem))
    return mem


def _reader_of(obj) -> pa.RecordBatchReader:
    if isinstance(obj, pa.RecordBatchReader):
        return obj
    if isinstance(obj, pa.Table):
        return obj.to_reader()
    if isinstance(obj, pa.RecordBatch):
        return pa.RecordBatchReader.from_batches(obj.schema, [obj])
    raise TypeError(type(obj))


def scan_column(arr: pa.Array, name: str = "c") -> oracle_py.Slab:
    """The reference's arrow_scan over a one-column stream holding `arr` as one record batch."""
    L = lib()
    batch = pa.RecordBatch.from_arrays([arr], schema=pa.schema([pa.field(name, arr.type)]))
    mem = _export_stream(_reader_of(batch))
    out = C.POINTER(oracle_py.OSlab)()
    err = C.create_string_buffer(512)
    dict_len = len(arr.dictionary) if pa.types.is_dictionary(arr.type) else 0
    rc = L.ref_scan_column(C.addressof(mem), dict_len, C.byref(out), err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    slab = oracle_py.Slab(out)
    L.ref_slab_free(out)
    slab.arrow = arr
    return slab


def scan_stream(obj) -> tuple[int, int]:
    """-> (rows, checksum)"""
    L = lib()
    mem = _export_stream(_reader_of(obj))
    err = C.create_string_buffer(512)
    cs = C.c_uint64(0)
    rows = L.ref_scan_stream(C.addressof(mem), C.byref(cs), err, 512)
    if rows < 0:
        raise RefError(4, err.value.decode())
    return int(rows), int(cs.value)


def append(obj, flags: int = 0) -> pa.StructArray:
    """Rows of `obj` (RecordBatch / Table / reader) -> reference arrow_scan -> ONE DataChunk -> ArrowConverter::ToArrowArray."""
    L = lib()
    mem = _export_stream(_reader_of(obj))
    c_arr = oracle_py.ArrowArray()
    c_sch = oracle_py.ArrowSchema()
    err = C.create_string_buffer(512)
    rc = L.ref_append(C.addressof(mem), flags, C.addressof(c_arr), C.addressof(c_sch), err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    return pa.Array._import_from_c(C.addressof(c_arr), C.addressof(c_sch))


def register(name: str, obj) -> None:
    L = lib()
    mem = _export_stream(_reader_of(obj))
    _streams.append(mem)
    err = C.create_string_buffer(512)
    rc = L.ref_register_stream(name.encode(), C.addressof(mem), err, 512)
    if rc != 0:
        raise RefError(rc, err.value.decode())


def query(sql: str) -> list[list[str]]:
    L = lib()
    res = C.c_char_p()
    err = C.create_string_buffer(1024)
    rc = L.ref_query(sql.encode(), C.byref(res), err, 1024)
    if rc != 0:
        raise RefError(rc, err.value.decode())
    text = res.value.decode() if res.value else ""
    return [line.split("|") for line in text.splitlines()]


def set_threads(n: int) -> None:
    err = C.create_string_buffer(512)
    if lib().ref_set_threads(n, err, 512) != 0:
        raise RefError(4, err.value.decode())


def export(sql: str, batch_rows: int = 122_880, flags: int = 0):
    """Generator of pyarrow RecordBatches whose buffers are the reference appender's own (validity always
    materialised, arrow_appender.cpp:94)."""
    L = lib()
    c_sch = oracle_py.ArrowSchema()
    err = C.create_string_buffer(1024)
    h = L.ref_export_begin(sql.encode(), flags, C.addressof(c_sch), err, 1024)
    if not h:
        raise RefError(4, err.value.decode())
    try:
        schema = pa.Schema._import_from_c(C.addressof(c_sch))
        while True:
            c_arr = oracle_py.ArrowArray()
            rc = L.ref_export_next(h, batch_rows, C.addressof(c_arr), err, 1024)
            if rc < 0:
                raise RefError(4, err.value.decode())
            if rc == 0:
                break
            yield pa.RecordBatch._import_from_c(C.addressof(c_arr), schema)
    finally:
        L.ref_export_end(h)


def source_id() -> str:
    return lib().ref_source_id().decode()


class BufferedFile:
    """DuckDB's BufferedFileReader (the reference reads every IPC message through it, 4 KiB at a time)."""

    def __init__(self, path: str):
        err = C.create_string_buffer(512)
        self._h = lib().ref_file_open(path.encode(), err, 512)
        if not self._h:
            raise RefError(4, err.value.decode())

    def read_into(self, location: int, size: int, dst_address: int) -> None:
        err = C.create_string_buffer(512)
        if lib().ref_file_read(self._h, location, size, dst_address, err, 512) != 0:
            raise RefError(4, err.value.decode())

    def close(self):
        if self._h:
            lib().ref_file_close(self._h)
            self._h = None

    def __del__(self):
        self.close()


def load_extension(path: str, name: str = "nanoarrow") -> None:
    """dlopen + <name>_init(DatabaseInstance&) on the session's database (what LOAD does after its signature check)."""
    path = os.path.abspath(path)
    if path in _loaded_extensions:
        return
    err = C.create_string_buffer(1024)
    if lib().ref_load_extension(path.encode(), name.encode(), err, 1024) != 0:
        raise RefError(4, err.value.decode())
    _loaded_extensions.add(path)


def scan_arrow_ipc_view(name: str, buffers: list[tuple[int, int]]) -> None:
    """CREATE VIEW name AS FROM scan_arrow_ipc([{ptr, size}, ...]) — the POINTER values are built through the API, as
    DuckDB's clients do."""
    n = len(buffers)
    ptrs = (C.c_uint64 * n)(*[b[0] for b in buffers])
    sizes = (C.c_uint64 * n)(*[b[1] for b in buffers])
    err = C.create_string_buffer(1024)
    if lib().ref_scan_arrow_ipc_view(name.encode(), ptrs, sizes, n, err, 1024) != 0:
        raise RefError(4, err.value.decode())
