"""Seeded input tables for the writer tests (the types DuckDB's ArrowAppender can write), and a helper that rebuilds
pyarrow arrays from (nodes, buffers) as produced by the oracle / found in a record batch."""
from __future__ import annotations

import decimal

import numpy as np
import pyarrow as pa


def _nullify(rng, values: list, p=0.15):
    return [None if rng.random() < p else v for v in values]


def scalar_table(n=5000, seed=1) -> pa.Table:
    rng = np.random.default_rng(seed)
    words = ["", "a", "bc", "twelve bytes", "thirteen byte", "a considerably longer string value €µ",
             "x" * 100, "mid-size str"]
    cols = {
        "i8": pa.array(_nullify(rng, rng.integers(-128, 128, n).tolist()), pa.int8()),
        "i16": pa.array(rng.integers(-2**15, 2**15, n), pa.int16()),
        "i32": pa.array(_nullify(rng, rng.integers(-2**31, 2**31, n).tolist()), pa.int32()),
        "i64": pa.array(_nullify(rng, rng.integers(-2**62, 2**62, n).tolist()), pa.int64()),
        "u8": pa.array(rng.integers(0, 256, n), pa.uint8()),
        "u16": pa.array(rng.integers(0, 2**16, n), pa.uint16()),
        "u32": pa.array(_nullify(rng, rng.integers(0, 2**32, n).tolist()), pa.uint32()),
        "u64": pa.array(rng.integers(0, 2**63, n).astype(np.uint64), pa.uint64()),
        "f32": pa.array(_nullify(rng, rng.random(n).astype(np.float32).tolist()), pa.float32()),
        "f64": pa.array(rng.random(n), pa.float64()),
        "b": pa.array(_nullify(rng, (rng.random(n) < 0.5).tolist()), pa.bool_()),
        "b_nonull": pa.array(rng.random(n) < 0.3, pa.bool_()),
        "date": pa.array(_nullify(rng, rng.integers(0, 20000, n).tolist()), pa.date32()),
        "ts_us": pa.array(_nullify(rng, rng.integers(0, 2**50, n).tolist()), pa.timestamp("us")),
        "ts_s": pa.array(rng.integers(0, 2**31, n), pa.timestamp("s")),
        "ts_ms": pa.array(rng.integers(0, 2**41, n), pa.timestamp("ms")),
        "ts_ns": pa.array(rng.integers(0, 2**60, n), pa.timestamp("ns")),
        "ts_tz": pa.array(_nullify(rng, rng.integers(0, 2**50, n).tolist()), pa.timestamp("us", tz="UTC")),
        "time": pa.array(rng.integers(0, 86400_000_000, n), pa.time64("us")),
        "dec4": pa.array(_nullify(rng, [decimal.Decimal(int(x)).scaleb(-1) for x in rng.integers(-999, 1000, n)]),
                         pa.decimal32(4, 1)),
        "dec9": pa.array([decimal.Decimal(int(x)).scaleb(-2) for x in rng.integers(-10**8, 10**8, n)],
                         pa.decimal32(9, 2)),
        "dec18": pa.array(_nullify(rng, [decimal.Decimal(int(x)).scaleb(-3) for x in rng.integers(-10**17, 10**17, n)]),
                          pa.decimal64(18, 3)),
        "dec38": pa.array(_nullify(rng, [decimal.Decimal(int(x) * 10**15 + 7).scaleb(-4)
                                         for x in rng.integers(-10**17, 10**17, n)]), pa.decimal128(38, 4)),
        "s": pa.array(_nullify(rng, [words[i] for i in rng.integers(0, len(words), n)]), pa.string()),
        "s_nonull": pa.array([words[i] + str(i) for i in rng.integers(0, len(words), n)], pa.string()),
        "bin": pa.array(_nullify(rng, [bytes(rng.integers(0, 256, int(l)).astype(np.uint8))
                                       for l in rng.integers(0, 40, n)]), pa.binary()),
        "nul": pa.nulls(n),
        "iv": pa.array(_nullify(rng, [pa.MonthDayNano([int(a), int(b), int(c) * 1000]) for a, b, c in
                                      zip(rng.integers(-100, 100, n), rng.integers(-1000, 1000, n),
                                          rng.integers(-10**12, 10**12, n))]), pa.month_day_nano_interval()),
    }
    return pa.table(cols)


def nested_table(n=3000, seed=2) -> pa.Table:
    rng = np.random.default_rng(seed)

    def ilist(maxlen=5, p=0.15):
        return [None if rng.random() < p else rng.integers(-1000, 1000, int(rng.integers(0, maxlen))).tolist()
                for _ in range(n)]

    strs = ["", "k", "key-two", "a key longer than twelve", "zz"]
    cols = {
        "l_i64": pa.array(ilist(), pa.list_(pa.int64())),
        "l_str": pa.array([None if rng.random() < 0.1 else
                           [None if rng.random() < 0.2 else strs[int(j)] for j in rng.integers(0, len(strs), int(rng.integers(0, 4)))]
                           for _ in range(n)], pa.list_(pa.string())),
        "ll": pa.array([None if rng.random() < 0.1 else
                        [None if rng.random() < 0.2 else rng.integers(0, 9, int(rng.integers(0, 3))).tolist()
                         for _ in range(int(rng.integers(0, 4)))] for _ in range(n)], pa.list_(pa.list_(pa.int32()))),
        "st": pa.array([None if rng.random() < 0.1 else {"a": int(i) if i % 7 else None, "b": strs[i % len(strs)]}
                        for i in range(n)], pa.struct([("a", pa.int32()), ("b", pa.string())])),
        "l_st": pa.array([None if rng.random() < 0.1 else
                          [{"x": int(j), "y": None if j % 3 == 0 else float(j)} for j in range(int(rng.integers(0, 4)))]
                          for _ in range(n)], pa.list_(pa.struct([("x", pa.int64()), ("y", pa.float64())]))),
        "m": pa.array([None if rng.random() < 0.1 else
                       [(f"k{j}", None if j % 4 == 0 else int(j)) for j in range(int(rng.integers(0, 4)))]
                       for _ in range(n)], pa.map_(pa.string(), pa.int64())),
        "arr": pa.array([None if rng.random() < 0.15 else [int(i), None if i % 5 == 0 else int(i + 1), int(i + 2)]
                         for i in range(n)], pa.list_(pa.int32(), 3)),
        "l_arr": pa.array([None if rng.random() < 0.1 else
                           [None if rng.random() < 0.2 else [int(j), int(j) + 1] for j in range(int(rng.integers(0, 4)))]
                           for _ in range(n)], pa.list_(pa.list_(pa.int16(), 2))),
        "st_l": pa.array([{"v": None if i % 6 == 0 else list(range(i % 4)), "w": bool(i % 2)} for i in range(n)],
                         pa.struct([("v", pa.list_(pa.int8())), ("w", pa.bool_())])),
    }
    return pa.table(cols)


def lossless_table(n=4000, seed=12):
    """One column of every type arrow_lossless_conversion changes (arrow_converter.cpp:130-279) plus the two that are
    extension types whatever the setting (UHUGEINT, VARINT; :387-393), as the Arrow extension arrays the reference scans
    into BOOLEAN / UUID / HUGEINT / UHUGEINT / TIME_TZ / BIT / VARINT vectors; a struct and a list carry them nested."""
    rng = np.random.default_rng(seed)

    def fixed(width, null_p=0.2):
        raw = rng.integers(0, 256, (n, width)).astype(np.uint8)
        raw[0], raw[1 % n] = 0, 255
        mask = rng.random(n) < null_p
        st = pa.Array.from_buffers(pa.binary(width), n, [None, pa.py_buffer(raw.tobytes())])
        return st, mask

    def with_nulls(st, mask):
        valid = np.packbits(~mask, bitorder="little").tobytes()
        bufs = st.buffers()
        return pa.Array.from_buffers(st.type, len(st), [pa.py_buffer(valid)] + bufs[1:], int(mask.sum()))

    def opaque(st, type_name):
        return pa.ExtensionArray.from_storage(pa.opaque(st.type, type_name, "DuckDB"), st)

    cols = {}
    st, m = fixed(16)
    cols["huge"] = opaque(with_nulls(st, m), "hugeint")
    st, m = fixed(16)
    cols["uhuge"] = opaque(with_nulls(st, m), "uhugeint")
    st, m = fixed(16)
    cols["id"] = pa.ExtensionArray.from_storage(pa.uuid(), with_nulls(st, m))
    st, _ = fixed(16)
    cols["id_nn"] = pa.ExtensionArray.from_storage(pa.uuid(), st)
    micros = rng.integers(0, 86_400_000_000, n).astype(np.uint64)
    offs = rng.integers(0, 2 * 57599, n).astype(np.uint64)  # dtime_tz_t: micros << 24 | (MAX_OFFSET - offset)
    bits = ((micros << np.uint64(24)) | offs).astype("<u8")
    st = pa.Array.from_buffers(pa.binary(8), n, [None, pa.py_buffer(bits.tobytes())])
    cols["ttz"] = opaque(with_nulls(st, rng.random(n) < 0.2), "time_tz")
    flags = rng.integers(0, 3, n).astype(np.int8)  # bool8: any non-zero byte is true
    cols["flag"] = pa.ExtensionArray.from_storage(pa.bool8(), with_nulls(pa.array(flags, pa.int8()), rng.random(n) < 0.2))
    # BIT: first byte = number of padding bits (0..7), padding bits set; VARINT: 3-byte header + magnitude
    bit_vals = []
    for _ in range(n):
        nb = int(rng.integers(1, 20))
        pad = int(rng.integers(0, 8))
        body = bytearray(rng.integers(0, 256, nb).astype(np.uint8).tobytes())
        body[0] |= (0xFF << (8 - pad)) & 0xFF if pad else 0
        bit_vals.append(bytes([pad]) + bytes(body))
    cols["bits"] = opaque(pa.array(_nullify(rng, bit_vals), pa.binary()), "bit")
    var_vals = []
    for _ in range(n):
        nb = int(rng.integers(1, 24))
        mag = bytearray(rng.integers(0, 256, nb).astype(np.uint8).tobytes())
        if nb > 1 and mag[0] == 0:
            mag[0] = 1
        head = (0x800000 | nb).to_bytes(3, "big")  # positive: header = byte count with the top bit set
        var_vals.append(head + bytes(mag))
    cols["vint"] = opaque(pa.array(_nullify(rng, var_vals), pa.binary()), "varint")
    cols["plain"] = pa.array(_nullify(rng, rng.integers(-1000, 1000, n).tolist()), pa.int32())
    cols["s"] = pa.array(_nullify(rng, ["v%d" % i if i % 3 else "a string well past twelve bytes %d" % i
                                         for i in range(n)]), pa.string())
    nested = pa.StructArray.from_arrays([cols["id"], cols["flag"], cols["bits"]], names=["id", "flag", "bits"])
    cols["st"] = nested
    lens = rng.integers(0, 4, n)
    lo = np.zeros(n + 1, np.int32)
    np.cumsum(lens, out=lo[1:])
    take = pa.array(rng.integers(0, n, int(lo[-1])), pa.int64())
    cols["l_id"] = pa.ListArray.from_arrays(pa.array(lo, pa.int32()), cols["id"].take(take))
    cols["l_flag"] = pa.ListArray.from_arrays(pa.array(lo, pa.int32()), cols["flag"].take(take))
    return pa.table(cols)


def rare_columns(n=5000, seed=4):
    """The writable types Arrow has no 1:1 array for (so they cannot come from vectors.from_arrow): UUID (written as its
    36-character string), TIME_TZ (time64[us], offset dropped), BIT (the internal bytes as binary) and sparse UNIONs
    with flat members -> (HostVectors, the table pyarrow must read back)"""
    import uuid
    from vectors import HostVector, from_arrow, time_tz_vector, union_vector, uuid_vector
    rng = np.random.default_rng(seed)
    raw = rng.integers(0, 256, (n, 16)).astype(np.uint8)
    raw[0], raw[1 % n] = 0, 255
    ids = [None if rng.random() < 0.15 else uuid.UUID(bytes=raw[i].tobytes()) for i in range(n)]
    ids_nn = [uuid.UUID(bytes=raw[::-1][i].tobytes()) for i in range(n)]
    micros = rng.integers(0, 86_400_000_000, n)
    offs = rng.integers(-57599, 57600, n)
    tz_valid = rng.random(n) >= 0.2
    bits = from_arrow(pa.array(_nullify(rng, [bytes(rng.integers(0, 256, int(l)).astype(np.uint8))
                                                for l in rng.integers(1, 30, n)]), pa.binary()), "bits")
    bits_expected = pa.array([None if not v else bits.heap[s:s + l].tobytes() for v, s, l in
                              zip(np.ones(n, bool) if bits.validity is None else bits.validity, bits.str_start,
                                  bits.str_len)], pa.binary())
    bits.type = "BIT"
    # union(num BIGINT, str VARCHAR, flag BOOLEAN, d DECIMAL(4,1)): NULL union rows, NULL member values
    tags = rng.integers(0, 4, n)
    u_valid = rng.random(n) >= 0.1
    words = ["", "u", "union member", "a union member longer than twelve bytes"]
    m_num = pa.array(_nullify(rng, rng.integers(-2**40, 2**40, n).tolist()), pa.int64())
    m_str = pa.array(_nullify(rng, [words[i] for i in rng.integers(0, len(words), n)]), pa.string())
    m_flag = pa.array(_nullify(rng, (rng.random(n) < 0.5).tolist()), pa.bool_())
    m_dec = pa.array(_nullify(rng, [decimal.Decimal(int(x)).scaleb(-1) for x in rng.integers(-999, 1000, n)]),
                     pa.decimal32(4, 1))
    members = [m_num, m_str, m_flag, m_dec]
    names = ["num", "str", "flag", "d"]
    union = union_vector(tags, [from_arrow(a, nm) for a, nm in zip(members, names)], u_valid, "u")
    exp_union = [None if not u_valid[i] else members[int(tags[i])][i].as_py() for i in range(n)]
    # the same union below a LIST (the members are appended through the list's child selection)
    lens = rng.integers(0, 4, n)
    l_valid = rng.random(n) >= 0.1
    starts = np.minimum(np.cumsum(lens) - lens, max(n - 4, 0))
    ent = np.stack([starts, lens], 1).astype(np.uint64)
    inner = union_vector(tags, [from_arrow(a, nm) for a, nm in zip(members, names)], u_valid, "l")
    l_union = HostVector("LIST", n, ent.view(np.uint8).reshape(-1), None if l_valid.all() else l_valid, [inner],
                         name="l_u")
    exp_l_union = [None if not l_valid[i] else exp_union[int(starts[i]): int(starts[i] + lens[i])] for i in range(n)]
    cols = [uuid_vector(ids, "id"), uuid_vector(ids_nn, "id_nn"),
            time_tz_vector(micros, offs, None if tz_valid.all() else tz_valid, "ttz"), bits, union, l_union]
    expected = {
        "id": [None if v is None else str(v) for v in ids],
        "id_nn": [str(v) for v in ids_nn],
        "ttz": pa.array(np.where(tz_valid, micros, 0), pa.int64(), mask=~tz_valid).cast(pa.time64("us")).to_pylist(),
        "bits": bits_expected.to_pylist(),
        "u": exp_union,
        "l_u": exp_l_union,
    }
    types = [pa.string(), pa.string(), pa.time64("us"), pa.binary(),
             pa.sparse_union([pa.field(nm, a.type) for a, nm in zip(members, names)]),
             pa.list_(pa.sparse_union([pa.field(nm, a.type) for a, nm in zip(members, names)]))]
    return cols, expected, types


def arrays_from_buffers(types: list[pa.DataType], nodes: list, buffers: list[bytes]) -> list[pa.Array]:
    """Rebuild pyarrow arrays from pre-order (length, null_count) nodes and buffers (validity first)."""
    ni, bi = [0], [0]

    def take_buf():
        b = buffers[bi[0]]
        bi[0] += 1
        return pa.py_buffer(b) if len(b) else None

    def build(t: pa.DataType) -> pa.Array:
        length, nulls = nodes[ni[0]]
        ni[0] += 1
        if pa.types.is_null(t):
            return pa.nulls(length)
        if pa.types.is_union(t):  # sparse, V5: type ids only
            ids = take_buf() or pa.py_buffer(b"")
            kids = [build(t.field(i).type) for i in range(t.num_fields)]
            return pa.UnionArray.from_buffers(t, length, [None, ids], children=kids)
        validity = take_buf()
        if pa.types.is_struct(t):
            kids = [build(t.field(i).type) for i in range(t.num_fields)]
            return pa.StructArray.from_buffers(t, length, [validity], nulls, children=kids)
        if pa.types.is_list(t) or pa.types.is_large_list(t) or pa.types.is_map(t):
            offs = take_buf() or pa.py_buffer(b"\0" * 8)
            if pa.types.is_map(t):
                child = build(pa.struct([pa.field("key", t.key_type, nullable=False), pa.field("value", t.item_type)]))
            else:
                child = build(t.value_type)
            return pa.Array.from_buffers(t, length, [validity, offs], nulls, children=[child])
        if pa.types.is_list_view(t) or pa.types.is_large_list_view(t):
            offs, sizes = take_buf() or pa.py_buffer(b""), take_buf() or pa.py_buffer(b"")
            child = build(t.value_type)
            return pa.Array.from_buffers(t, length, [validity, offs, sizes], nulls, children=[child])
        if pa.types.is_string_view(t):  # validity, views, the one variadic data buffer
            views, data = take_buf() or pa.py_buffer(b""), take_buf() or pa.py_buffer(b"")
            return pa.Array.from_buffers(t, length, [validity, views, data], nulls)
        if pa.types.is_fixed_size_list(t):
            child = build(t.value_type)
            return pa.Array.from_buffers(t, length, [validity], nulls, children=[child])
        if (pa.types.is_string(t) or pa.types.is_binary(t) or pa.types.is_large_string(t)
                or pa.types.is_large_binary(t)):
            offs = take_buf() or pa.py_buffer(b"\0" * 8)
            data = take_buf() or pa.py_buffer(b"")
            return pa.Array.from_buffers(t, length, [validity, offs, data], nulls)
        data = take_buf() or pa.py_buffer(b"")
        return pa.Array.from_buffers(t, length, [validity, data], nulls)

    out = [build(t) for t in types]
    assert ni[0] == len(nodes) and bi[0] == len(buffers), "nodes / buffers not fully consumed"
    return out


def writer_type(t: pa.DataType, large: bool = False, view: bool = False, sview: bool = False) -> pa.DataType:
    """The Arrow type the reference's converter gives the same DuckDB type under SET arrow_large_buffer_size = true
    (arrow_converter.cpp:179-196,262-293: utf8 -> large_utf8, binary -> large_binary, list -> large_list) and / or SET
    arrow_output_list_view = true (:282-287: list -> list_view / large_list_view); map, struct and fixed-size list keep
    their kind, their children change."""
    def rec(x):
        return writer_type(x, large, view, sview)
    if pa.types.is_string(t):  # produce_arrow_string_view wins over the offset size (arrow_converter.cpp:189-197)
        return pa.string_view() if sview else pa.large_string() if large else t
    if pa.types.is_binary(t):
        return pa.large_binary() if large else t
    if pa.types.is_list(t):
        f = t.value_field.with_type(rec(t.value_type))
        if view:
            return pa.large_list_view(f) if large else pa.list_view(f)
        return pa.large_list(f) if large else pa.list_(f)
    if pa.types.is_fixed_size_list(t):
        return pa.list_(t.value_field.with_type(rec(t.value_type)), t.list_size)
    if pa.types.is_map(t):
        return pa.map_(rec(t.key_type), t.item_field.with_type(rec(t.item_type)))
    if pa.types.is_struct(t):
        return pa.struct([t.field(i).with_type(rec(t.field(i).type)) for i in range(t.num_fields)])
    if pa.types.is_union(t):
        return pa.sparse_union([t.field(i).with_type(rec(t.field(i).type)) for i in range(t.num_fields)], t.type_codes)
    return t


def large_type(t: pa.DataType) -> pa.DataType:
    return writer_type(t, large=True)
