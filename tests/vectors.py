"""Test tooling: DuckDB vectors (host numpy form) built from pyarrow arrays, and their upload to the device as
duckdb_arrow_b200.DeviceVector.  This is an independent construction of the writer's INPUT (it does not go through
the product's reader), so writer parity does not lean on reader parity.

HostVector layout = DuckDB's flat vector layout: fixed-width values, 1 byte per BOOLEAN,
string_t{u32 len; 12 inline bytes | 4 prefix bytes + pointer}, list_entry_t{u64 offset, u64 length}, validity as a
bool array (None = all valid)."""
from __future__ import annotations

import numpy as np
import pyarrow as pa

VS = 2048
STRING_TYPES = ("VARCHAR", "BLOB", "BIT", "VARINT", "JSON")  # string_t vectors


class HostVector:
    def __init__(self, type: str, rows: int, data: np.ndarray | None = None, validity: np.ndarray | None = None,
                 children=(), name="", decimal=(0, 0), param=0, timezone=None, str_start=None, str_len=None, heap=None):
        self.type, self.rows, self.data, self.validity = type, rows, data, validity
        self.children, self.name, self.decimal, self.param, self.timezone = list(children), name, decimal, param, timezone
        # VARCHAR / BLOB: byte range of every string inside `heap` (the string_t records in `data` are derived)
        self.str_start, self.str_len, self.heap = str_start, str_len, heap

    @property
    def width(self):
        return 0 if self.data is None or self.rows == 0 else self.data.size // self.rows


def _validity(arr: pa.Array):
    if arr.null_count == 0:
        return None
    return np.asarray(arr.is_valid().to_numpy(zero_copy_only=False), dtype=bool)


def _values(arr: pa.Array, dtype) -> np.ndarray:
    buf = arr.buffers()[1]
    if buf is None:
        return np.zeros(len(arr), dtype)
    return np.frombuffer(buf, dtype=dtype)[arr.offset: arr.offset + len(arr)].copy()


def make_string_t(start: np.ndarray, length: np.ndarray, heap: np.ndarray, ptr_base: int) -> np.ndarray:
    """string_t(const char*, uint32_t), duckdb/src/include/duckdb/common/types/string_type.hpp:43-64"""
    n = len(length)
    rec = np.zeros((n, 16), np.uint8)
    rec[:, 0:4] = length.astype(np.uint32).view(np.uint8).reshape(n, 4)
    pad = np.concatenate([heap, np.zeros(16, np.uint8)])
    idx = start[:, None] + np.arange(12)[None, :]
    first12 = pad[np.minimum(idx, len(pad) - 1)]
    inline = length <= 12
    mask = np.arange(12)[None, :] < length[:, None]
    rec[inline, 4:16] = np.where(mask, first12, 0)[inline]
    long_ = ~inline
    rec[long_, 4:8] = first12[long_, 0:4]
    ptr = (np.uint64(ptr_base) + start.astype(np.uint64))
    rec[long_, 8:16] = ptr.view(np.uint8).reshape(n, 8)[long_]
    return rec.reshape(-1)


def _and_mask(hv: HostVector, mask: np.ndarray) -> None:
    """parent NULLs -> child NULLs, down through row-aligned children (STRUCT fields, ARRAY elements)"""
    if hv.type == "SQLNULL":
        return
    hv.validity = mask.copy() if hv.validity is None else (hv.validity & mask)
    if hv.type in STRING_TYPES and hv.str_len is not None:
        hv.str_len = np.where(hv.validity, hv.str_len, 0)
    if hv.type == "STRUCT":
        for k in hv.children:
            _and_mask(k, mask)
    elif hv.type == "ARRAY":
        _and_mask(hv.children[0], np.repeat(mask, hv.param))


def from_arrow(arr: pa.Array, name: str = "") -> HostVector:
    """Arrow array -> the DuckDB vector the reference's scan would produce for it (logical content)."""
    if isinstance(arr, pa.ChunkedArray):
        arr = arr.combine_chunks()
    t, n = arr.type, len(arr)
    valid = _validity(arr)
    fixed = {pa.int8(): ("TINYINT", np.int8), pa.int16(): ("SMALLINT", np.int16), pa.int32(): ("INTEGER", np.int32),
             pa.int64(): ("BIGINT", np.int64), pa.uint8(): ("UTINYINT", np.uint8), pa.uint16(): ("USMALLINT", np.uint16),
             pa.uint32(): ("UINTEGER", np.uint32), pa.uint64(): ("UBIGINT", np.uint64), pa.float32(): ("FLOAT", np.float32),
             pa.float64(): ("DOUBLE", np.float64), pa.date32(): ("DATE", np.int32), pa.time64("us"): ("TIME", np.int64),
             pa.timestamp("us"): ("TIMESTAMP", np.int64), pa.timestamp("s"): ("TIMESTAMP_S", np.int64),
             pa.timestamp("ms"): ("TIMESTAMP_MS", np.int64), pa.timestamp("ns"): ("TIMESTAMP_NS", np.int64)}
    if t in fixed:
        ty, dt = fixed[t]
        return HostVector(ty, n, _values(arr, dt).view(np.uint8), valid, name=name)
    if pa.types.is_timestamp(t) and t.tz is not None and t.unit == "us":
        return HostVector("TIMESTAMP WITH TIME ZONE", n, _values(arr, np.int64).view(np.uint8), valid, name=name,
                          timezone=t.tz)
    if pa.types.is_null(t):
        return HostVector("SQLNULL", n, None, None, name=name)
    if pa.types.is_boolean(t):
        b = np.asarray(arr.fill_null(False).to_numpy(zero_copy_only=False), dtype=np.uint8)
        return HostVector("BOOLEAN", n, b, valid, name=name)
    if pa.types.is_decimal(t):
        w = t.precision
        raw = _values(arr, np.dtype((np.void, t.bit_width // 8))).view(np.uint8).reshape(n, t.bit_width // 8)
        pw = 2 if w <= 4 else 4 if w <= 9 else 8 if w <= 18 else 16  # DuckDB physical width
        if pw <= raw.shape[1]:
            data = np.ascontiguousarray(raw[:, :pw])
        else:  # sign-extend
            neg = (raw[:, -1] & 0x80) != 0
            data = np.zeros((n, pw), np.uint8)
            data[:, :raw.shape[1]] = raw
            data[neg, raw.shape[1]:] = 0xFF
        return HostVector("DECIMAL", n, data.reshape(-1), valid, name=name, decimal=(w, t.scale))
    if pa.types.is_string(t) or pa.types.is_binary(t):
        offs = _values_offsets(arr, np.int32)
        heap = np.frombuffer(arr.buffers()[2], np.uint8).copy() if arr.buffers()[2] is not None else np.zeros(0, np.uint8)
        start = offs[:-1].astype(np.int64)
        length = (offs[1:] - offs[:-1]).astype(np.int64)
        if valid is not None:
            length = np.where(valid, length, 0)
        hv = HostVector("VARCHAR" if pa.types.is_string(t) else "BLOB", n, None, valid, name=name, str_start=start,
                        str_len=length, heap=heap)
        return hv
    if pa.types.is_list(t) or pa.types.is_map(t):
        offs = _values_offsets(arr, np.int32).astype(np.uint64)
        ent = np.zeros((n, 2), np.uint64)
        ent[:, 0] = offs[:-1]
        ent[:, 1] = offs[1:] - offs[:-1]
        if pa.types.is_map(t):
            child = HostVector("STRUCT", len(arr.keys), None, None,
                               [from_arrow(arr.keys, "key"), from_arrow(arr.items, "value")], name="entries")
            return HostVector("MAP", n, ent.view(np.uint8).reshape(-1), valid, [child], name=name)
        return HostVector("LIST", n, ent.view(np.uint8).reshape(-1), valid, [from_arrow(arr.values, "l")], name=name)
    if pa.types.is_fixed_size_list(t):
        child = from_arrow(arr.values, "")
        if valid is not None:  # the scan broadcasts a NULL array row over its k child rows (arrow_conversion.cpp:313-323)
            _and_mask(child, np.repeat(valid, t.list_size))
        return HostVector("ARRAY", n, None, valid, [child], name=name, param=t.list_size)
    if pa.types.is_struct(t):
        kids = [from_arrow(arr.field(i), t.field(i).name) for i in range(t.num_fields)]
        if valid is not None:  # the scan ANDs the struct's NULLs into every child (arrow_conversion.cpp:1143-1151)
            for k in kids:
                _and_mask(k, valid)
        return HostVector("STRUCT", n, None, valid, kids, name=name)
    if pa.types.is_dictionary(t):  # a DuckDB ENUM: codes sized by the dictionary (EnumTypeInfo::DictType) + its values
        nvals = len(arr.dictionary)
        dt = np.uint8 if nvals <= 0xFF else np.uint16 if nvals <= 0xFFFF else np.uint32
        idx = arr.indices
        codes = _values(idx, np.dtype(idx.type.to_pandas_dtype())).astype(dt)
        return HostVector("ENUM", n, codes.view(np.uint8), valid, [from_arrow(arr.dictionary.cast(pa.string()), "dictionary")],
                          name=name)
    if isinstance(t, pa.BaseExtensionType):  # the extension types of arrow_type_extension.cpp:366-390
        st = arr.storage
        ext = t.extension_name
        if ext == "arrow.bool8":  # ArrowBool8::ArrowToDuck: every row, NULL slots included
            return HostVector("BOOLEAN", n, (_values(st, np.int8) != 0).astype(np.uint8), valid, name=name)
        if ext == "arrow.uuid":  # big-endian bytes -> hugeint_t{lower, upper ^ sign}; NULL slots are not converted
            raw = _values(st, np.dtype((np.void, 16))).view(np.uint8).reshape(n, 16)
            be = raw.view(">u8").reshape(n, 2)
            out = np.zeros((n, 2), np.uint64)
            out[:, 0] = be[:, 1]
            out[:, 1] = be[:, 0] ^ np.uint64(1 << 63)
            if valid is not None:
                out[~valid] = 0
            return HostVector("UUID", n, out.view(np.uint8).reshape(-1), valid, name=name)
        if ext == "arrow.json":
            hv = from_arrow(st, name)
            hv.type = "JSON"
            return hv
        if ext == "arrow.opaque" and t.vendor_name == "DuckDB":
            kind = {"hugeint": "HUGEINT", "uhugeint": "UHUGEINT", "time_tz": "TIME_TZ", "bit": "BIT",
                    "varint": "VARINT"}[t.type_name]
            if kind in ("BIT", "VARINT"):
                hv = from_arrow(st, name)
                hv.type = kind
                return hv
            w = st.type.byte_width
            return HostVector(kind, n, _values(st, np.dtype((np.void, w))).view(np.uint8).reshape(-1), valid, name=name)
        raise NotImplementedError(str(t))
    if str(t) == "month_day_nano_interval":
        raw = _values(arr, np.dtype([("m", np.int32), ("d", np.int32), ("ns", np.int64)]))
        out = np.zeros(n, np.dtype([("m", np.int32), ("d", np.int32), ("us", np.int64)]))
        out["m"], out["d"], out["us"] = raw["m"], raw["d"], raw["ns"] // 1000
        return HostVector("INTERVAL", n, out.view(np.uint8).reshape(-1), valid, name=name)
    raise NotImplementedError(str(t))


def uuid_vector(values: list, name: str = "") -> HostVector:
    """uuid.UUID | None -> DuckDB UUID vector: hugeint_t{u64 lower; i64 upper} with the top bit of `upper` flipped
    (UUID::FromString, duckdb/src/common/types/uuid.cpp:8-68)"""
    n = len(values)
    data = np.zeros((n, 2), np.uint64)
    valid = np.array([v is not None for v in values], bool)
    for i, v in enumerate(values):
        if v is not None:
            data[i, 0] = v.int & ((1 << 64) - 1)
            data[i, 1] = (v.int >> 64) ^ (1 << 63)
    return HostVector("UUID", n, data.view(np.uint8).reshape(-1), None if valid.all() else valid, name=name)


def time_tz_vector(micros: np.ndarray, offset_s: np.ndarray, valid: np.ndarray | None, name: str = "") -> HostVector:
    """dtime_tz_t(time, offset): bits = micros << 24 | (MAX_OFFSET - offset) (datetime.hpp: OFFSET_BITS = 24,
    MAX_OFFSET = 16 * 60 * 60 - 1)"""
    bits = (micros.astype(np.uint64) << np.uint64(24)) | (57599 - offset_s.astype(np.int64)).astype(np.uint64)
    return HostVector("TIME_TZ", len(micros), bits.view(np.uint8).reshape(-1), valid, name=name)


def union_vector(tags: np.ndarray, members: list[HostVector], valid: np.ndarray | None, name: str = "") -> HostVector:
    """DuckDB UNION vector = STRUCT{tag UTINYINT, members...}; here (like on the read side) the node's own data are
    the tags"""
    return HostVector("UNION", len(tags), tags.astype(np.uint8), valid, members, name=name)


def _values_offsets(arr, dtype):
    buf = arr.buffers()[1]
    if buf is None:
        return np.zeros(len(arr) + 1, dtype)
    return np.frombuffer(buf, dtype=dtype)[arr.offset: arr.offset + len(arr) + 1].copy()


def table_vectors(table: pa.Table) -> list[HostVector]:
    return [from_arrow(table.column(i), table.schema.field(i).name) for i in range(table.num_columns)]


def to_device(hv: HostVector, device: int = 0):
    """Upload a HostVector tree; VARCHAR string_t pointers are patched to the device address of the uploaded heap."""
    import torch
    import duckdb_arrow_b200 as b2a
    dev = f"cuda:{device}"
    keep = []
    data = None
    if hv.type in STRING_TYPES:
        heap = torch.from_numpy(np.concatenate([hv.heap, np.zeros(16, np.uint8)])).to(dev)
        keep.append(heap)
        rec = make_string_t(hv.str_start, hv.str_len, hv.heap, heap.data_ptr())
        data = torch.from_numpy(rec).to(dev) if rec.size else torch.zeros(16, dtype=torch.uint8, device=dev)
    elif hv.data is not None:
        data = torch.from_numpy(np.ascontiguousarray(hv.data)).to(dev) if hv.data.size else \
            torch.zeros(16, dtype=torch.uint8, device=dev)
    validity = None
    if hv.validity is not None:
        bits = np.packbits(hv.validity, bitorder="little")
        words = np.zeros(((hv.rows + 63) // 64) * 8, np.uint8)
        words[: bits.size] = bits
        validity = torch.from_numpy(words).to(dev)
    kids = [to_device(c, device) for c in hv.children]
    return b2a.DeviceVector(hv.type, hv.rows, data, validity, kids, hv.name, hv.decimal, hv.param, hv.timezone, keep)
