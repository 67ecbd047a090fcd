"""Python driver of the CPU oracle (oracle.c).  TEST INFRASTRUCTURE ONLY — see oracle.h.

Feeds pyarrow arrays to the oracle through the Arrow C data interface (the same structs the reference's
ArrowToDuckDB consumes) and returns the converted column as numpy slabs.  The Arrow type -> DuckDB type
mapping below restates duckdb/src/function/table/arrow/arrow_duck_schema.cpp:59-297."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np
import pyarrow as pa

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")

T = {n: i for i, n in enumerate(
    ["INVALID", "SQLNULL", "BOOLEAN", "TINYINT", "SMALLINT", "INTEGER", "BIGINT", "UTINYINT", "USMALLINT", "UINTEGER",
     "UBIGINT", "HUGEINT", "UHUGEINT", "FLOAT", "DOUBLE", "DECIMAL", "VARCHAR", "BLOB", "DATE", "TIME", "TIMESTAMP",
     "TIMESTAMP_SEC", "TIMESTAMP_MS", "TIMESTAMP_NS", "TIMESTAMP_TZ", "INTERVAL", "UUID", "LIST", "STRUCT", "MAP",
     "ARRAY", "UNION", "TIME_TZ", "BIT", "VARINT", "JSON", "ENUM"])}
WIDTH = {"BOOLEAN": 1, "TINYINT": 1, "UTINYINT": 1, "SMALLINT": 2, "USMALLINT": 2, "INTEGER": 4, "UINTEGER": 4,
         "FLOAT": 4, "DATE": 4, "BIGINT": 8, "UBIGINT": 8, "DOUBLE": 8, "TIME": 8, "TIMESTAMP": 8, "TIMESTAMP_SEC": 8,
         "TIMESTAMP_MS": 8, "TIMESTAMP_NS": 8, "TIMESTAMP_TZ": 8, "HUGEINT": 16, "UHUGEINT": 16, "UUID": 16,
         "INTERVAL": 16, "VARCHAR": 16, "BLOB": 16, "LIST": 16, "MAP": 16, "SQLNULL": 0, "STRUCT": 0, "ARRAY": 0, "DECIMAL": 0,
         "UNION": 1, "TIME_TZ": 8, "BIT": 16, "VARINT": 16, "JSON": 16}
SIZE_NORMAL, SIZE_SUPER, SIZE_FIXED, SIZE_VIEW = 0, 1, 2, 3
DT = {"ms": 0, "us": 1, "ns": 2, "s": 3, "days": 4, "months": 5, "mdn": 6}


class OType(C.Structure):
    pass


OType._fields_ = [("id", C.c_int32), ("width", C.c_int32), ("size_type", C.c_int32), ("datetime", C.c_int32),
                  ("bit_width", C.c_int32), ("index_type", C.c_int32), ("fixed_size", C.c_int64),
                  ("n_children", C.c_int32), ("has_dictionary", C.c_int32), ("children", C.POINTER(C.POINTER(OType))),
                  ("run_end_encoded", C.c_int32), ("bool8", C.c_int32),
                  ("list_view", C.c_int32)]


class ArrowArray(C.Structure):  # Arrow C data interface; OArray in oracle.h is its prefix
    pass


ArrowArray._fields_ = [("length", C.c_int64), ("null_count", C.c_int64), ("offset", C.c_int64),
                       ("n_buffers", C.c_int64), ("n_children", C.c_int64), ("buffers", C.POINTER(C.c_void_p)),
                       ("children", C.POINTER(C.POINTER(ArrowArray))), ("dictionary", C.POINTER(ArrowArray)),
                       ("release", C.c_void_p), ("private_data", C.c_void_p)]


class ArrowSchema(C.Structure):
    _fields_ = [("format", C.c_char_p), ("name", C.c_char_p), ("metadata", C.c_void_p), ("flags", C.c_int64),
                ("n_children", C.c_int64), ("children", C.c_void_p), ("dictionary", C.c_void_p),
                ("release", C.c_void_p), ("private_data", C.c_void_p)]


class OSlab(C.Structure):
    pass


OSlab._fields_ = [("rows", C.c_uint64), ("cap", C.c_uint64), ("width", C.c_int32), ("type_id", C.c_int32),
                  ("data", C.POINTER(C.c_uint8)), ("valid", C.POINTER(C.c_uint8)), ("n_children", C.c_int32),
                  ("all_null", C.c_int32), ("children", C.POINTER(C.POINTER(OSlab))), ("sel", C.POINTER(C.c_uint32)),
                  ("dict", C.POINTER(OSlab))]

_lib = None


def build() -> str:
    subprocess.check_call(["make", "-s", "-C", HERE])
    return LIB


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(os.path.join(HERE, "oracle.c")):
            build()
        _lib = C.CDLL(LIB)
        _lib.oracle_scan_column.restype = C.c_int
        _lib.oracle_scan_column.argtypes = [C.c_void_p, C.POINTER(OType), C.c_int64, C.POINTER(C.POINTER(OSlab)),
                                            C.c_char_p, C.c_int]
        _lib.oracle_slab_free.argtypes = [C.POINTER(OSlab)]
        _lib.oracle_scan_batch_checksum.restype = C.c_uint64
        _lib.oracle_scan_batch_checksum.argtypes = [C.c_void_p, C.POINTER(C.POINTER(OType)), C.c_int]
        _lib.oracle_read_arrow_batch_checksum.restype = C.c_uint64
        _lib.oracle_read_arrow_batch_checksum.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p,
                                                          C.POINTER(C.POINTER(OType)), C.c_int]
        _lib.oracle_validate_full.restype = C.c_int
        _lib.oracle_validate_full.argtypes = [C.c_void_p, C.POINTER(OType)]
    return _lib


class OracleError(Exception):
    def __init__(self, code, msg):
        self.code = code
        self.kind = {1: "ConversionException", 2: "NotImplementedException", 3: "InvalidInputException"}.get(code, "?")
        super().__init__(f"{self.kind}: {msg}")


_keep = []  # ctypes objects referenced by pointer from OType trees


def _decimal_width(precision: int) -> int:
    return 2 if precision <= 4 else 4 if precision <= 9 else 8 if precision <= 18 else 16


def make_type(t: pa.DataType) -> OType:
    """GetTypeFromFormat / GetArrowLogicalType restated on pyarrow types."""
    o = OType()
    kids = []

    def simple(name, **kw):
        o.id = T[name]
        o.width = WIDTH[name]
        for k, v in kw.items():
            setattr(o, k, v)

    if pa.types.is_dictionary(t):
        inner = make_type(t.value_type)
        inner.has_dictionary = 1
        it = t.index_type
        inner.index_type = T[{"int8": "TINYINT", "int16": "SMALLINT", "int32": "INTEGER", "int64": "BIGINT",
                              "uint8": "UTINYINT", "uint16": "USMALLINT", "uint32": "UINTEGER",
                              "uint64": "UBIGINT"}[str(it)]]
        return inner
    if pa.types.is_run_end_encoded(t):
        # arrow_duck_schema.cpp:276-297: the column has the type of its values; children = {run_ends, values}
        inner = make_type(t.value_type)
        kids = [make_type(t.run_end_type), make_type(t.value_type)]
        arr = (C.POINTER(OType) * 2)(*[C.pointer(k) for k in kids])
        _keep.extend(kids)
        _keep.append(arr)
        inner.children = arr
        inner.n_children = 2
        inner.run_end_encoded = 1
        return inner
    if isinstance(t, pa.BaseExtensionType):
        # ArrowType::GetTypeFromSchema (arrow_duck_schema.cpp:372-386) + the default registrations
        # (arrow_type_extension.cpp:366-390)
        name, st = getattr(t, "extension_name", ""), t.storage_type
        vendor, type_name = getattr(t, "vendor_name", ""), getattr(t, "type_name", "")
        inner = make_type(st)
        if name == "arrow.uuid" and st == pa.binary(16):
            simple("UUID")
            return o
        if name == "arrow.bool8" and st == pa.int8():
            simple("BOOLEAN", bool8=1)
            return o
        if name == "arrow.json":
            if not (pa.types.is_string(st) or pa.types.is_large_string(st) or pa.types.is_string_view(st)):
                raise OracleError(3, f'Arrow extension type "{st}" not supported for arrow.json')
            inner.id = T["JSON"]
            return inner
        if name == "arrow.opaque" and vendor == "DuckDB":
            if type_name in ("hugeint", "uhugeint") and st == pa.binary(16):
                simple(type_name.upper())
                return o
            if type_name == "time_tz" and st == pa.binary(8):
                simple("TIME_TZ")
                return o
            if type_name in ("bit", "varint"):
                if not (pa.types.is_binary(st) or pa.types.is_large_binary(st)):
                    raise OracleError(3, f'Arrow extension type "{st}" not supported for {type_name}')
                inner.id = T[type_name.upper()]
                return inner
        return inner
    if pa.types.is_null(t):
        simple("SQLNULL")
    elif pa.types.is_boolean(t):
        simple("BOOLEAN")
    elif pa.types.is_integer(t):
        simple({"int8": "TINYINT", "int16": "SMALLINT", "int32": "INTEGER", "int64": "BIGINT", "uint8": "UTINYINT",
                "uint16": "USMALLINT", "uint32": "UINTEGER", "uint64": "UBIGINT"}[str(t)])
    elif pa.types.is_float32(t):
        simple("FLOAT")
    elif pa.types.is_float64(t):
        simple("DOUBLE")
    elif pa.types.is_decimal(t):
        if t.precision > 38 or t.bit_width > 128:
            raise OracleError(2, f"Unsupported Internal Arrow Type for Decimal d:{t.precision},{t.scale},{t.bit_width}")
        simple("DECIMAL", bit_width=t.bit_width)
        o.width = _decimal_width(t.precision)
    elif pa.types.is_string(t):
        simple("VARCHAR", size_type=SIZE_NORMAL)
    elif pa.types.is_large_string(t):
        simple("VARCHAR", size_type=SIZE_SUPER)
    elif pa.types.is_string_view(t):
        simple("VARCHAR", size_type=SIZE_VIEW)
    elif pa.types.is_binary(t):
        simple("BLOB", size_type=SIZE_NORMAL)
    elif pa.types.is_large_binary(t):
        simple("BLOB", size_type=SIZE_SUPER)
    elif pa.types.is_fixed_size_binary(t):
        simple("BLOB", size_type=SIZE_FIXED, fixed_size=t.byte_width)
    elif pa.types.is_date32(t):
        simple("DATE", datetime=DT["days"])
    elif pa.types.is_date64(t):
        simple("DATE", datetime=DT["ms"])
    elif pa.types.is_time(t):
        simple("TIME", datetime=DT[t.unit])
    elif pa.types.is_timestamp(t):
        if t.tz is None:
            simple({"s": "TIMESTAMP_SEC", "ms": "TIMESTAMP_MS", "us": "TIMESTAMP", "ns": "TIMESTAMP_NS"}[t.unit])
        else:
            simple("TIMESTAMP_TZ", datetime=DT[t.unit])
    elif pa.types.is_duration(t):
        simple("INTERVAL", datetime=DT[t.unit])
    elif pa.types.is_interval(t):
        name = str(t)
        simple("INTERVAL", datetime=DT["mdn"] if "month_day_nano" in name else DT["days"] if "day_time" in name
               else DT["months"])
    elif pa.types.is_list(t) or pa.types.is_large_list(t):
        simple("LIST", size_type=SIZE_SUPER if pa.types.is_large_list(t) else SIZE_NORMAL)
        kids = [make_type(t.value_type)]
    elif pa.types.is_list_view(t) or pa.types.is_large_list_view(t):
        simple("LIST", size_type=SIZE_SUPER if pa.types.is_large_list_view(t) else SIZE_NORMAL, list_view=1)
        kids = [make_type(t.value_type)]
    elif pa.types.is_map(t):
        simple("MAP", size_type=SIZE_NORMAL)
        kids = [make_type(pa.struct([pa.field("key", t.key_type), pa.field("value", t.item_type)]))]
    elif pa.types.is_fixed_size_list(t):
        simple("ARRAY", fixed_size=t.list_size)
        kids = [make_type(t.value_type)]
    elif pa.types.is_struct(t):
        if t.num_fields == 0:
            raise OracleError(3, "Attempted to convert a STRUCT with no fields to DuckDB which is not supported")
        simple("STRUCT")
        kids = [make_type(t.field(i).type) for i in range(t.num_fields)]
    elif pa.types.is_union(t):
        # arrow_duck_schema.cpp:235-259: sparse only; the declared type codes are ignored
        if t.mode != "sparse":
            raise OracleError(2, 'Unsupported Internal Arrow Type: "d" Union')
        if t.num_fields == 0:
            raise OracleError(3, "Attempted to convert a UNION with no fields to DuckDB which is not supported")
        simple("UNION")
        kids = [make_type(t.field(i).type) for i in range(t.num_fields)]
    else:
        raise OracleError(2, f"Unsupported Internal Arrow Type {t}")
    if kids:
        arr = (C.POINTER(OType) * len(kids))(*[C.pointer(k) for k in kids])
        _keep.extend(kids)
        _keep.append(arr)
        o.children = arr
        o.n_children = len(kids)
    return o


class Slab:
    """numpy view of one converted column (copied out of the C slab)."""

    def __init__(self, p):
        s = p.contents
        self.rows = int(s.rows)
        self.width = int(s.width)
        self.type_id = int(s.type_id)
        self.all_null = bool(s.all_null)
        self.valid = np.ctypeslib.as_array(s.valid, shape=(self.rows,)).astype(bool) if self.rows else np.zeros(0, bool)
        if self.width and self.rows:
            self.data = np.ctypeslib.as_array(s.data, shape=(self.rows * self.width,)).copy().reshape(self.rows, self.width)
        else:
            self.data = np.zeros((self.rows, self.width), np.uint8)
        self.sel = None
        self.dict = None
        if s.sel:
            self.sel = np.ctypeslib.as_array(s.sel, shape=(self.rows,)).copy() if self.rows else np.zeros(0, np.uint32)
            self.dict = Slab(s.dict) if s.dict else None
        self.children = [Slab(s.children[i]) for i in range(s.n_children)]


def scan_column(arr: pa.Array, parent_offset: int = 0) -> Slab:
    """Convert one Arrow array the way the reference's scan would, 2 048 rows at a time."""
    L = lib()
    typ = make_type(arr.type)
    c_arr = ArrowArray()
    c_sch = ArrowSchema()
    arr._export_to_c(C.addressof(c_arr), C.addressof(c_sch))
    out = C.POINTER(OSlab)()
    err = C.create_string_buffer(256)
    try:
        rc = L.oracle_scan_column(C.addressof(c_arr), C.byref(typ), parent_offset, C.byref(out), err, 256)
        if rc != 0:
            raise OracleError(rc, err.value.decode())
        slab = Slab(out)
        L.oracle_slab_free(out)
        slab.arrow = arr  # keeps the character buffers alive: string_t pointers refer to them
        return slab
    finally:
        for st, ty in ((c_arr, ArrowArray), (c_sch, ArrowSchema)):
            if st.release:
                C.CFUNCTYPE(None, C.POINTER(ty))(st.release)(C.byref(st))


def scan_batches_checksum(batches: list[pa.RecordBatch]) -> int:
    """Timed CPU baseline body: full reference-style scan of every column of every batch."""
    L = lib()
    total = 0
    if not batches:
        return 0
    types = [make_type(f.type) for f in batches[0].schema]
    tarr = (C.POINTER(OType) * len(types))(*[C.pointer(t) for t in types])
    for b in batches:
        c_arr = ArrowArray()
        c_sch = ArrowSchema()
        b._export_to_c(C.addressof(c_arr), C.addressof(c_sch))
        total += L.oracle_scan_batch_checksum(C.addressof(c_arr), tarr, len(types))
        for st, ty in ((c_arr, ArrowArray), (c_sch, ArrowSchema)):
            if st.release:
                C.CFUNCTYPE(None, C.POINTER(ty))(st.release)(C.byref(st))
    return total


def validate_full(arr: pa.Array) -> bool:
    """nanoarrow FULL validation of one array (True = valid)."""
    L = lib()
    typ = make_type(arr.type)
    c_arr, c_sch = ArrowArray(), ArrowSchema()
    arr._export_to_c(C.addressof(c_arr), C.addressof(c_sch))
    try:
        return L.oracle_validate_full(C.addressof(c_arr), C.byref(typ)) == 0
    finally:
        for st, ty in ((c_arr, ArrowArray), (c_sch, ArrowSchema)):
            if st.release:
                C.CFUNCTYPE(None, C.POINTER(ty))(st.release)(C.byref(st))


def stream_messages(buf, max_batches: int | None = None):
    """[(body address, body bytes, RecordBatch)] of an in-memory IPC stream (pyarrow does the framing, untimed).
    The returned list keeps the message bodies alive."""
    import pyarrow.ipc as ipc
    reader = ipc.MessageReader.open_stream(pa.py_buffer(buf) if not isinstance(buf, pa.Buffer) else buf)
    schema, out = None, []
    for msg in reader:
        if msg.type == "schema":
            schema = ipc.read_schema(msg)
        elif msg.type == "record batch":
            out.append((msg.body.address, msg.body.size, ipc.read_record_batch(msg, schema), msg))
            if max_batches and len(out) >= max_batches:
                break
    return out


def read_arrow_checksum(messages) -> int:
    """Timed CPU baseline body: per record-batch message DecodeBody (4 KiB bounce copy) + FULL validation + scan."""
    L = lib()
    total = 0
    if not messages:
        return 0
    types = [make_type(f.type) for f in messages[0][2].schema]
    tarr = (C.POINTER(OType) * len(types))(*[C.pointer(t) for t in types])
    for addr, size, b, _ in messages:
        c_arr, c_sch = ArrowArray(), ArrowSchema()
        b._export_to_c(C.addressof(c_arr), C.addressof(c_sch))
        r = L.oracle_read_arrow_batch_checksum(addr, size, C.addressof(c_arr), tarr, len(types))
        for st, ty in ((c_arr, ArrowArray), (c_sch, ArrowSchema)):
            if st.release:
                C.CFUNCTYPE(None, C.POINTER(ty))(st.release)(C.byref(st))
        if r == 0xFFFFFFFFFFFFFFFF:
            raise OracleError(3, "Arrow IPC: offsets are not monotonically increasing")
        total += r
    return total & 0xFFFFFFFFFFFFFFFF

