"""Test tooling: the oracle scan's slab (pinned to the reference's scan, tests/test_ref_scan.py) -> tests/vectors.HostVector, i.e.
the writer's input exactly as the reference's scan leaves it — NULLs of parents pushed into children the way the reference
does it per child type, which tests/vectors.from_arrow only approximates for BLOB / nested ARRAY children."""
import ctypes
import numpy as np
import pyarrow as pa
from oracle import oracle_py as O
from vectors import HostVector

NAME = {v: k for k, v in O.T.items()}
RENAME = {"TIMESTAMP_SEC": "TIMESTAMP_S", "TIMESTAMP_TZ": "TIMESTAMP WITH TIME ZONE"}
VS = 2048


def slab_vector(slab, t, name="", bounds=None):
    if isinstance(t, pa.BaseExtensionType) or pa.types.is_dictionary(t) or pa.types.is_string_view(t) or pa.types.is_union(t):
        raise NotImplementedError(str(t))
    ty = NAME[slab.type_id]
    ty = RENAME.get(ty, ty)
    n = slab.rows
    if bounds is None:
        bounds = list(range(0, n, VS)) + [n]
    valid = None if slab.valid.all() else slab.valid.copy()
    if ty == "SQLNULL":
        return HostVector("SQLNULL", n, None, None, name=name)
    if ty in ("VARCHAR", "BLOB"):
        d = slab.data
        lens = d[:, 0:4].copy().view(np.uint32).ravel().astype(np.int64)
        lens = np.where(slab.valid, lens, 0)
        parts, start, pos = [], np.zeros(n, np.int64), 0
        for i in range(n):
            ln = int(lens[i])
            start[i] = pos
            if ln == 0:
                continue
            if ln <= 12:
                parts.append(d[i, 4:4 + ln].tobytes())
            else:
                parts.append(ctypes.string_at(int(d[i, 8:16].copy().view(np.uint64)[0]), ln))
            pos += ln
        heap = np.frombuffer(b"".join(parts), np.uint8).copy() if parts else np.zeros(0, np.uint8)
        return HostVector(ty, n, None, valid, name=name, str_start=start, str_len=lens, heap=heap)
    if ty in ("LIST", "MAP"):
        ent = slab.data.copy().view(np.uint64).reshape(n, 2)
        # per-chunk offsets -> offsets into the concatenated child
        base = 0
        out = ent.copy()
        child_bounds = [0]
        for lo, hi in zip(bounds[:-1], bounds[1:]):
            out[lo:hi, 0] = ent[lo:hi, 0] + base
            base += int(ent[lo:hi, 1].sum())
            child_bounds.append(base)
        ct = t.value_type if not pa.types.is_map(t) else pa.struct([pa.field("key", t.key_type), pa.field("value", t.item_type)])
        child = slab_vector(slab.children[0], ct, "entries" if ty == "MAP" else "l", child_bounds)
        return HostVector(ty, n, out.view(np.uint8).reshape(-1), valid, [child], name=name)
    if ty == "ARRAY":
        return HostVector("ARRAY", n, None, valid, [slab_vector(slab.children[0], t.value_type, "", [b * t.list_size for b in bounds])], name=name, param=t.list_size)
    if ty == "STRUCT":
        kids = [slab_vector(slab.children[i], t.field(i).type, t.field(i).name, bounds) for i in range(t.num_fields)]
        return HostVector("STRUCT", n, None, valid, kids, name=name)
    data = slab.data.copy().reshape(-1)
    if ty == "DECIMAL":
        return HostVector("DECIMAL", n, data, valid, name=name, decimal=(t.precision, t.scale))
    if ty == "TIMESTAMP WITH TIME ZONE":
        return HostVector(ty, n, data, valid, name=name, timezone=t.tz)
    return HostVector(ty, n, data, valid, name=name)


def table_vectors_from_oracle_scan(table):
    return [slab_vector(O.scan_column(table.column(i).combine_chunks()), table.schema.field(i).type, table.schema.field(i).name)
            for i in range(table.num_columns)]
