"""Logical (Python value) views used to pin the oracle against golden vectors:
  slab_to_py      : oracle slab -> list of Python values (what DuckDB would show for the vector)
  column_to_py    : product ColumnView -> the same (used by a few GPU tests)
  golden_json     : arrow-testing integration JSON (*.json.gz) -> expected DuckDB values, with the unit
                    conversions of arrow_conversion.cpp restated as plain formulas."""
from __future__ import annotations

import ctypes
import gzip
import json
import struct

import numpy as np
import pyarrow as pa


def _trunc_div(a: int, b: int) -> int:
    q = abs(a) // abs(b)
    return q if (a >= 0) == (b >= 0) else -q


def _int(raw: bytes, signed=True) -> int:
    return int.from_bytes(raw, "little", signed=signed)


def _string(rec: np.ndarray) -> bytes:
    ln = _int(rec[0:4].tobytes(), False)
    if ln <= 12:
        return rec[4:4 + ln].tobytes()
    ptr = _int(rec[8:16].tobytes(), False)
    return ctypes.string_at(ptr, ln)


def _storage_type(t):
    while isinstance(t, pa.BaseExtensionType) or pa.types.is_run_end_encoded(t):
        t = t.value_type if pa.types.is_run_end_encoded(t) else t.storage_type
    return t


def _scalar(rec: np.ndarray, t: pa.DataType):
    raw = rec.tobytes()
    if pa.types.is_boolean(t):
        return bool(raw[0])
    if pa.types.is_integer(t):
        return _int(raw, pa.types.is_signed_integer(t))
    if pa.types.is_float32(t):
        return struct.unpack("<f", raw)[0]
    if pa.types.is_float64(t):
        return struct.unpack("<d", raw)[0]
    if pa.types.is_decimal(t) or pa.types.is_date(t) or pa.types.is_time(t) or pa.types.is_timestamp(t):
        return _int(raw, True)
    if pa.types.is_duration(t) or pa.types.is_interval(t):
        return (_int(raw[0:4]), _int(raw[4:8]), _int(raw[8:16]))
    if pa.types.is_string(t) or pa.types.is_large_string(t) or pa.types.is_binary(t) or pa.types.is_large_binary(t) \
            or pa.types.is_fixed_size_binary(t) or pa.types.is_string_view(t):
        return _string(rec)
    raise NotImplementedError(str(t))


def slab_to_py(slab, t: pa.DataType) -> list:
    t = _storage_type(t)
    n = slab.rows
    if pa.types.is_dictionary(t):
        if slab.dict is None:  # a dictionary column that arrived flattened (the reference does that to MAP values)
            return slab_to_py(slab, t.value_type)
        vals = slab_to_py(slab.dict, t.value_type)
        return [vals[int(slab.sel[i])] if int(slab.sel[i]) < len(vals) else None for i in range(n)]
    if pa.types.is_null(t):
        return [None] * n
    valid = slab.valid
    if pa.types.is_struct(t):
        kids = [slab_to_py(slab.children[i], t.field(i).type) for i in range(t.num_fields)]
        return [tuple(k[i] for k in kids) if valid[i] else None for i in range(n)]
    if pa.types.is_union(t):
        # DuckDB UNION = STRUCT{tag, members...}: the logical value of a valid row is the tagged member's value
        kids = [slab_to_py(slab.children[i], t.field(i).type) for i in range(t.num_fields)]
        tags = slab.data.ravel()
        return [kids[int(tags[i])][i] if valid[i] else None for i in range(n)]
    if pa.types.is_list(t) or pa.types.is_large_list(t) or pa.types.is_map(t) or pa.types.is_list_view(t) \
            or pa.types.is_large_list_view(t):
        if pa.types.is_map(t):
            ct = pa.struct([pa.field("key", t.key_type), pa.field("value", t.item_type)])
        else:
            ct = t.value_type
        child = slab_to_py(slab.children[0], ct)
        ent = slab.data.copy().view(np.uint64).reshape(n, 2)
        # list_entry_t offsets are relative to the child vector of the row's 2 048-row chunk; the chunks' child vectors
        # (sum of the chunk's lengths each) lie end to end in the child slab
        out, cum, base = [], 0, 0
        for i in range(n):
            if i % 2048 == 0:
                base = cum
            off, ln = int(ent[i, 0]), int(ent[i, 1])
            out.append(child[base + off: base + off + ln] if valid[i] else None)
            cum += ln
        return out
    if pa.types.is_fixed_size_list(t):
        k = t.list_size
        child = slab_to_py(slab.children[0], t.value_type)
        return [child[i * k:(i + 1) * k] if valid[i] else None for i in range(n)]
    return [_scalar(slab.data[i], t) if valid[i] else None for i in range(n)]


class _ViewSlab:
    """adapts a product ColumnView to the slab interface of slab_to_py (host string pointers required)"""

    def __init__(self, cv, lo=0, hi=None, ranges=None):
        if ranges is None:
            ranges = [(lo, cv.rows if hi is None else hi)]
        self.rows = sum(b - a for a, b in ranges)
        self.valid = np.concatenate([cv.valid[a:b] for a, b in ranges]) if ranges else cv.valid[0:0]
        self.data = np.concatenate([cv.data[a:b] for a, b in ranges]) if ranges else cv.data[0:0]
        self.sel = None
        self.dict = None
        self.children = []
        if cv.is_dictionary:
            self.sel = self.data.copy().view(np.uint32).ravel()
            d = cv.dictionary()
            self.dict = _ViewSlab(d, 0, d.rows)
            return
        kids = cv.children()
        ft = cv.field.type
        whole = len(ranges) == 1 and ranges[0] == (0, cv.rows)
        if ft in ("LIST", "MAP") and cv.field.op.startswith("LISTVIEW") and self.rows:
            if not whole:
                raise NotImplementedError("nested list view slices")
            win = cv.chunk_child_windows()
            self.children = [_ViewSlab(kids[0], ranges=[(int(a), int(a + b)) for a, b in win])]
        elif ft in ("LIST", "MAP") and self.rows:
            if not whole:
                raise NotImplementedError("nested list slices")
            starts = cv.chunk_child_starts()
            self.children = [_ViewSlab(kids[0], int(starts[0]), int(starts[-1]))]
        elif ft in ("LIST", "MAP"):
            self.children = [_ViewSlab(kids[0], 0, 0)]
        elif ft in ("STRUCT", "UNION"):
            self.children = [_ViewSlab(k, ranges=ranges) for k in kids]
        elif ft == "ARRAY":
            k = cv.field.param
            self.children = [_ViewSlab(kids[0], ranges=[(a * k, b * k) for a, b in ranges])]


def column_to_py(cv, t: pa.DataType) -> list:
    return slab_to_py(_ViewSlab(cv), t)


# --------------------------------------------------------------------------------------------- golden JSON
_TIME_MUL = {"SECOND": 1_000_000, "MILLISECOND": 1_000, "MICROSECOND": 1}


def _golden_values(col: dict, field: dict, dictionaries: dict) -> list:
    n = col["count"]
    ty = field["type"]
    name = ty["name"]
    validity = col.get("VALIDITY", [1] * n)
    if "dictionary" in field:
        dvals = dictionaries[field["dictionary"]["id"]]
        return [dvals[int(i)] if v else None for i, v in zip(col["DATA"], validity)]
    if name == "null":
        return [None] * n
    kids_f = field.get("children", [])
    kids_c = col.get("children", [])
    if name == "struct":
        kids = [_golden_values(c, f, dictionaries) for c, f in zip(kids_c, kids_f)]
        return [tuple(k[i] for k in kids) if validity[i] else None for i in range(n)]
    if name in ("list", "largelist", "map"):
        child = _golden_values(kids_c[0], kids_f[0], dictionaries)
        off = [int(o) for o in col["OFFSET"]]
        return [child[off[i]:off[i + 1]] if validity[i] else None for i in range(n)]
    if name == "fixedsizelist":
        k = ty["listSize"]
        child = _golden_values(kids_c[0], kids_f[0], dictionaries)
        return [child[i * k:(i + 1) * k] if validity[i] else None for i in range(n)]
    data = col.get("DATA", [])

    def conv(x):
        if name == "bool":
            return bool(x)
        if name == "int" or name == "decimal":
            return int(x)
        if name == "floatingpoint":
            return float(np.float32(x)) if ty["precision"] == "SINGLE" else float(x)
        if name in ("utf8", "largeutf8"):
            return x.encode("utf-8")
        if name in ("binary", "largebinary", "fixedsizebinary"):
            return bytes.fromhex(x)
        if name == "date":
            return int(x) if ty["unit"] == "DAY" else _trunc_div(int(x), 86_400_000)
        if name == "time":
            return _trunc_div(int(x), 1000) if ty["unit"] == "NANOSECOND" else int(x) * _TIME_MUL[ty["unit"]]
        if name == "timestamp":
            if not ty.get("timezone"):
                return int(x)
            return _trunc_div(int(x), 1000) if ty["unit"] == "NANOSECOND" else int(x) * _TIME_MUL[ty["unit"]]
        if name == "duration":
            us = _trunc_div(int(x), 1000) if ty["unit"] == "NANOSECOND" else int(x) * _TIME_MUL[ty["unit"]]
            return (0, 0, us)
        if name == "interval":
            if ty["unit"] == "YEAR_MONTH":
                return (int(x), 0, 0)
            if ty["unit"] == "MONTH_DAY_NANO":
                return (int(x["months"]), int(x["days"]), _trunc_div(int(x["nanoseconds"]), 1000))
        raise NotImplementedError(name)

    return [conv(data[i]) if validity[i] else None for i in range(n)]


def golden_json(path: str):
    """-> list of batches, each a list of (field name, expected values)"""
    with gzip.open(path, "rt") as f:
        j = json.load(f)
    fields = j["schema"]["fields"]

    def find_dict_field(fs, did):
        for f in fs:
            if f.get("dictionary", {}).get("id") == did:
                return f
            r = find_dict_field(f.get("children", []), did)
            if r:
                return r
        return None

    dictionaries = {}
    for d in j.get("dictionaries", []):
        f = dict(find_dict_field(fields, d["id"]))
        f.pop("dictionary")
        dictionaries[d["id"]] = _golden_values(d["data"]["columns"][0], f, dictionaries)
    out = []
    for b in j["batches"]:
        out.append([(f["name"], _golden_values(c, f, dictionaries)) for c, f in zip(b["columns"], fields)])
    return out


def same(a, b) -> bool:
    """equality that treats NaN == NaN and compares nested containers"""
    if isinstance(a, float) and isinstance(b, float):
        return a == b or (a != a and b != b)
    if isinstance(a, (list, tuple)) and isinstance(b, (list, tuple)):
        return len(a) == len(b) and all(same(x, y) for x, y in zip(a, b))
    return a == b
