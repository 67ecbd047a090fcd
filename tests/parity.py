"""Parity helpers: compare the product's decoded vectors (ColumnView over the arena written by the CUDA kernels)
with the CPU oracle's slabs (oracle/oracle.c, fed by pyarrow's own IPC decoding — an independent slicer).

Canonical comparison form: validity = first `rows` bits; fixed-width = bytes of valid rows;
string_t = (length, inline bytes | prefix + pointer offset from the column's character buffer);
list = list_entry_t[] of every row + the child over the range the chunks cover; dictionary = sel_t[] + values."""
from __future__ import annotations

import os
import sys

import numpy as np
import pyarrow as pa

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import oracle_py as O  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
VS = 2048


def _storage(arr: pa.Array) -> pa.Array:
    while isinstance(arr, pa.ExtensionArray):
        arr = arr.storage
    return arr


def _char_base(arr: pa.Array) -> int:
    """address of the character buffer the oracle's string_t pointers point into"""
    arr = _storage(arr)
    if pa.types.is_run_end_encoded(arr.type):  # the flattened strings still point into the values' character buffer
        arr = _storage(arr.values)
    t = arr.type
    if pa.types.is_fixed_size_binary(t):
        return arr.buffers()[1].address
    return arr.buffers()[2].address if arr.buffers()[2] is not None else 0


def _string_canon(data16: np.ndarray, base: int):
    lens = data16[:, 0:4].copy().view(np.uint32).ravel()
    ptr = data16[:, 8:16].copy().view(np.uint64).ravel().astype(np.int64)
    return lens, data16[:, 4:16], ptr - np.int64(base)


def compare_column(cv, slab: O.Slab, arr: pa.Array, path: str = "col", lo: int = 0, hi: int | None = None):
    """cv: duckdb_arrow_b200.ColumnView (product);  slab: oracle result for `arr`;  [lo,hi): the rows of the product
    column that correspond to the oracle slab (children of lists are decoded over their whole node)."""
    arr = _storage(arr)
    hi = cv.rows if hi is None else hi
    n = hi - lo
    assert n == slab.rows, f"{path}: rows {n} != oracle {slab.rows}"
    ftype = cv.field.type if not cv.is_dictionary else "DICT"
    valid = cv.valid[lo:hi]
    # a DuckDB dictionary vector has no mask of its own: NULL rows select the dictionary's extra NULL slot
    # (SetSelectionVector, arrow_conversion.cpp:1227-1348), so row validity is compared through sel + dictionary
    if ftype != "DICT":
        assert np.array_equal(valid, slab.valid), \
            f"{path}: validity differs at rows {np.nonzero(valid != slab.valid)[0][:8]}"
    if cv.co.stats_off != 0xFFFFFFFFFFFFFFFF and lo == 0 and hi == cv.rows:
        assert cv.valid_count == int(cv.valid.sum()), f"{path}: device popcount {cv.valid_count} != {int(cv.valid.sum())}"
    if ftype == "SQLNULL":
        assert slab.all_null or n == 0
        return
    if ftype == "DICT":
        sel = cv.data[lo:hi].copy().view(np.uint32).ravel()
        assert np.array_equal(sel, slab.sel), f"{path}: selection vector differs"
        dv = cv.dictionary()
        dict_arr = _storage(arr.dictionary)
        dslab = slab.dict
        # the oracle's dictionary slab carries the extra NULL slot only if the reference added it
        compare_values(dv, dslab, dict_arr, path + ".dict", 0, len(dict_arr), cv.field.type)
        if dslab.rows > len(dict_arr):
            assert not dslab.valid[len(dict_arr)]
            words = dv.validity_words(extra_rows=1)
            bit = (int(words[len(dict_arr) // 64]) >> (len(dict_arr) % 64)) & 1
            assert bit == 0, f"{path}: dictionary NULL slot is not invalid"
        return
    compare_values(cv, slab, arr, path, lo, hi, ftype)


def compare_values(cv, slab, arr, path, lo, hi, ftype):
    n = hi - lo
    ovalid = slab.valid[:n]
    if ftype in ("VARCHAR", "BLOB", "JSON", "BIT", "VARINT"):
        base = cv.ds.ptr_origin + int(cv.co.chars_off)
        plen, pin, prel = _string_canon(cv.data[lo:hi], base)
        olen, oin, orel = _string_canon(slab.data[:n], _char_base(arr))
        v = ovalid
        assert np.array_equal(plen[v], olen[v]), f"{path}: string lengths differ"
        short = v & (olen <= 12)
        long_ = v & (olen > 12)
        assert np.array_equal(pin[short], oin[short]), f"{path}: inlined bytes differ"
        assert np.array_equal(pin[long_][:, :4], oin[long_][:, :4]), f"{path}: prefixes differ"
        if pa.types.is_string_view(_storage(arr).type) or pa.types.is_binary_view(_storage(arr).type):
            # SetVectorStringView (arrow_conversion.cpp:425-452): pointer = data buffer [index] + offset.  The product
            # reports view pointers relative to the stream image (chars_off = 0), the oracle's are relative to pyarrow's
            # first data buffer: with ONE data buffer the two differ by that buffer's (8-byte aligned) position in the
            # image, the same for every row — and the bytes pointed at are compared as well
            # (several data buffers: pyarrow's lie anywhere in memory, only the bytes pointed at can be compared)
            if len(_storage(arr).buffers()) <= 3:
                delta = prel[long_] - orel[long_]
                assert delta.size == 0 or (np.all(delta == delta[0]) and delta[0] % 8 == 0), \
                    f"{path}: view pointers differ"
            import ctypes
            pp = cv.data[lo:hi][long_][:, 8:16].copy().view(np.uint64).ravel()
            op = slab.data[:n][long_][:, 8:16].copy().view(np.uint64).ravel()
            for a, b, ln in list(zip(pp.tolist(), op.tolist(), olen[long_].tolist()))[:: max(1, len(pp) // 500)]:
                assert ctypes.string_at(a, ln) == ctypes.string_at(b, ln), f"{path}: pointed-to bytes differ"
            return
        assert np.array_equal(prel[long_], orel[long_]), f"{path}: string pointers differ"
        return
    if ftype == "LIST" and cv.field.op.startswith("LISTVIEW"):
        assert lo == 0 and hi == cv.rows, "list views are compared whole"
        assert np.array_equal(cv.data[lo:hi], slab.data[:n]), f"{path}: list_entry_t differ"
        child_cv, child_arr = cv.children()[0], _storage(arr).values
        win = cv.chunk_child_windows() if n else np.zeros((0, 2), np.uint64)
        offs, sizes = _view_offsets_sizes(arr)
        done = 0
        for k in range((n + VS - 1) // VS):
            o, z = offs[k * VS:(k + 1) * VS], sizes[k * VS:(k + 1) * VS]
            lowest = min([int(o[0])] + [int(a) for a, b in zip(o, z) if b != 0])
            assert (int(win[k, 0]), int(win[k, 1])) == (lowest, int(z.sum())), f"{path}: child window of chunk {k} differs"
            c_lo, c_hi = lowest, lowest + int(z.sum())
            compare_child(child_cv, _slab_slice(slab.children[0], done, done + c_hi - c_lo), child_arr, f"{path}[{k}]",
                          c_lo, c_hi)
            done += c_hi - c_lo
        assert done == slab.children[0].rows
        return
    if ftype in ("LIST", "MAP"):
        assert np.array_equal(cv.data[lo:hi], slab.data[:n]), f"{path}: list_entry_t differ"
        child_cv = cv.children()[0]
        if lo == 0 and hi == cv.rows and n > 0:
            starts = cv.chunk_child_starts()
            c_lo, c_hi = int(starts[0]), int(starts[-1])
            offs = _list_offsets(arr)
            exp = np.array([offs[min(k * VS, n)] for k in range((n + VS - 1) // VS + 1)], dtype=np.uint64)
            assert np.array_equal(starts, exp), f"{path}: per-chunk child starts differ"
        else:
            offs = _list_offsets(arr)
            c_lo, c_hi = (int(offs[0]), int(offs[n])) if n else (0, 0)
        child_arr = _list_values(arr)
        sub = child_arr.slice(c_lo, c_hi - c_lo) if False else child_arr
        compare_child(child_cv, slab.children[0], sub, path + "[]", c_lo, c_hi)
        return
    if ftype == "UNION":
        p, o = cv.data[lo:hi], slab.data[:n]
        bad = np.nonzero((p != o).any(axis=1) & ovalid)[0]
        assert bad.size == 0, f"{path}: union tags differ at rows {bad[:8]}"
    if ftype in ("STRUCT", "UNION"):
        for i, ccv in enumerate(cv.children()):
            compare_child(ccv, slab.children[i], arr.field(i) if hasattr(arr, "field") else arr, f"{path}.{ccv.field.name}",
                          lo, hi)
        return
    if ftype == "ARRAY":
        k = cv.field.param
        compare_child(cv.children()[0], slab.children[0], _list_values(arr), path + "[k]", lo * k, hi * k)
        return
    # fixed width
    p, o = cv.data[lo:hi], slab.data[:n]
    assert p.shape == o.shape, f"{path}: width {p.shape} != {o.shape}"
    bad = np.nonzero((p != o).any(axis=1) & ovalid)[0]
    assert bad.size == 0, f"{path}: values differ at rows {bad[:8]} ({ftype})"


def compare_child(ccv, cslab, carr, path, lo, hi):
    carr = _storage(carr)
    if ccv.is_dictionary:
        n = hi - lo
        assert n == cslab.rows
        sel = ccv.data[lo:hi].copy().view(np.uint32).ravel()
        assert np.array_equal(sel, cslab.sel), f"{path}: selection vector differs"
        return
    assert hi - lo == cslab.rows, f"{path}: child rows {hi - lo} != oracle {cslab.rows}"
    v = ccv.valid[lo:hi]
    assert np.array_equal(v, cslab.valid), f"{path}: validity differs at {np.nonzero(v != cslab.valid)[0][:8]}"
    if ccv.field.type == "SQLNULL":
        return
    compare_values(ccv, cslab, carr, path, lo, hi, ccv.field.type)


class _SlabSlice:
    pass


def _slab_slice(slab, a, b):
    """rows [a, b) of an oracle slab (flat / struct / fixed-size-list / dictionary children only)"""
    s = _SlabSlice()
    s.rows = b - a
    s.width, s.type_id, s.all_null = slab.width, slab.type_id, slab.all_null
    s.valid = slab.valid[a:b]
    s.data = slab.data[a:b]
    s.sel = slab.sel[a:b] if slab.sel is not None else None
    s.dict = slab.dict
    if slab.children and slab.type_id == O.T["ARRAY"]:
        k = slab.children[0].rows // slab.rows if slab.rows else 0
        s.children = [_slab_slice(slab.children[0], a * k, b * k)]
    elif slab.children and slab.type_id in (O.T["STRUCT"], O.T["UNION"]):
        s.children = [_slab_slice(c, a, b) for c in slab.children]
    elif slab.children:
        raise NotImplementedError("slice of a list slab")
    else:
        s.children = []
    return s


def _view_offsets_sizes(arr):
    arr = _storage(arr)
    dt = np.int64 if pa.types.is_large_list_view(arr.type) else np.int32
    bufs = arr.buffers()
    n = len(arr)
    return (np.frombuffer(bufs[1], dtype=dt)[arr.offset: arr.offset + n].astype(np.int64),
            np.frombuffer(bufs[2], dtype=dt)[arr.offset: arr.offset + n].astype(np.int64))


def _list_offsets(arr):
    arr = _storage(arr)
    t = arr.type
    dt = np.int64 if pa.types.is_large_list(t) else np.int32
    buf = arr.buffers()[1]
    if buf is None:
        return np.zeros(1, np.int64)
    return np.frombuffer(buf, dtype=dt)[arr.offset: arr.offset + len(arr) + 1].astype(np.int64)


def _list_values(arr):
    arr = _storage(arr)
    if pa.types.is_map(arr.type):
        # the map's child is the "entries" struct (keys, items)
        return pa.StructArray.from_arrays([arr.keys, arr.items], names=["key", "value"])
    return arr.values


def compare_scan_with_oracle(ds, batches: list[pa.RecordBatch], columns: list[str] | None = None):
    """ds: DecodedScan of the same stream pyarrow decoded into `batches`."""
    reader = ds.reader
    out_cols = reader.output_columns()
    assert len(batches) == ds.scan.batch_hi - ds.scan.batch_lo
    for bi, batch in enumerate(batches):
        for f in out_cols:
            top_index = [c.index for c in reader.columns].index(f.index)
            arr = batch.column(top_index)
            slab = O.scan_column(arr)
            compare_column(ds.column(ds.scan.batch_lo + bi, f.index), slab, arr, f"batch{bi}.{f.name}")
