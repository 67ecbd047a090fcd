#!/usr/bin/env python
"""The differential fuzzes behind DESIGN section 7, as they were run by hand in the build container (they need oracle/_ref, the
reference's DuckDB compiled from /root/reference; CPU only).  A slice of each runs in the CPU test suite with fixed seeds.

  python tools/fuzz_reference.py scan   SEED0 COUNT   random nested arrays: oracle/oracle.c vs the reference's arrow scan
  python tools/fuzz_reference.py append SEED0 COUNT   random nested tables: oracle_write.py vs the reference's ArrowAppender
  python tools/fuzz_reference.py types  SEED0 COUNT   random Arrow types: the product's reader + binder vs the reference's DESCRIBE
  python tools/fuzz_reference.py plan   SEED0 COUNT   random nested, compressed streams through the product's reader + planner

`scan` with dictionaries (the default here) can abort the process: two shapes make the REFERENCE overflow a buffer (DESIGN
section 7) — run it in chunks (e.g. 50 seeds per process).  Every mode prints one line per difference and a summary."""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402
import pyarrow as pa  # noqa: E402
import pyarrow.ipc as ipc  # noqa: E402

import fuzz_arrays  # noqa: E402


def fuzz_scan(seed0, count):
    from test_ref_scan import check
    fuzz_arrays.ALLOW_DICTIONARY = True
    res = {"ok": 0, "error": 0, "diff": 0}
    for seed in range(seed0, seed0 + count):
        rng = np.random.default_rng(seed)
        n = int(rng.choice([1, 63, 64, 65, 100, 2047, 2048, 2049, 3000, 5000]))
        arr = fuzz_arrays.gen(rng, n)
        off = int(rng.choice([0, 0, 1, 7, 64, 2048])) if n > 2100 else int(rng.choice([0, 0, 1]))
        if off and off < len(arr):
            arr = arr.slice(off)
        try:
            res[check(arr, f"seed {seed}: {arr.type}")] += 1
        except AssertionError as e:
            res["diff"] += 1
            print("DIFF seed", seed, arr.type, len(arr), str(e)[:300])
    print(res)


def fuzz_append(seed0, count):
    import test_ref_append as TA
    from oracle import oracle_py as O, oracle_write as OW, ref_py as R
    from scan_vectors import slab_vector
    ok = skipped = loose = 0
    for seed in range(seed0, seed0 + count):
        rng = np.random.default_rng(seed)
        n = int(rng.choice([1, 100, 2047, 2048, 2049, 3000]))
        table = pa.table({f"c{i}": fuzz_arrays.gen(rng, n) for i in range(int(rng.integers(1, 4)))})
        rgs = int(rng.choice([1000, 2048, 5000]))
        done = False
        for flags in (0, R.FLAG_LARGE, R.FLAG_LIST_VIEW, R.FLAG_STRING_VIEW):
            kw = dict(large=bool(flags & R.FLAG_LARGE), view=bool(flags & R.FLAG_LIST_VIEW),
                      sview=bool(flags & R.FLAG_STRING_VIEW), lossless=False)
            for row in range(0, n, rgs):
                part = pa.table({nm: table.column(nm).combine_chunks().slice(row, min(rgs, n - row)) for nm in table.column_names})
                try:
                    cols = [slab_vector(O.scan_column(part.column(i).chunk(0)), part.schema.field(i).type,
                                        part.schema.field(i).name) for i in range(part.num_columns)]
                except NotImplementedError:
                    skipped += 1
                    done = True
                    break
                (nb, nodes, slices, body), = OW.serialize_batches(cols, part.num_rows, part.num_rows, **kw)
                ref = R.append(part, flags)
                rnodes, rbufs, notes = [], [], []
                for i in range(ref.type.num_fields):
                    TA.ref_nodes_buffers(ref.field(i), rnodes, rbufs, notes, ref.type.field(i).name)
                bad = None if rnodes == nodes else "nodes"
                for k, ((o, ln), rb) in enumerate(zip(slices, rbufs)):
                    if bad:
                        break
                    same, how = (False, "") if len(rb) != ln else TA._same_buffer(rb, body[o:o + ln], notes[k])
                    if not same:
                        bad = notes[k][:2]
                    loose += how == "valid rows"
                if bad:
                    print("DIFF seed", seed, "flags", flags, "rows", row, bad, [str(c.type) for c in table.columns])
                    done = True
                    break
                ok += 1
            if done:
                break
    print({"batches_equal": ok, "tables_skipped": skipped, "buffers_compared_on_valid_rows_only": loose})


def fuzz_types(seed0, count):
    import duckdb_arrow_b200 as b2a
    from oracle import ref_py as R
    tmp = tempfile.mkdtemp()
    ok = both = 0
    for seed in range(seed0, seed0 + count):
        t = fuzz_arrays.random_type(np.random.default_rng(seed))
        path = os.path.join(tmp, f"t{seed}.arrows")
        try:
            table = pa.table({"c": fuzz_arrays.empty_array(t)})
            with ipc.new_stream(path, table.schema) as w:
                w.write_table(table)
        except (pa.ArrowNotImplementedError, pa.ArrowInvalid, pa.ArrowTypeError):
            continue
        try:
            mine = b2a.read_arrow([path]).types[0]
        except b2a.B2AError as e:
            mine = "ERR " + str(e).split(": ", 1)[-1]
        try:
            R.register(f"t_types_{seed}", table)
            ref = R.query(f"DESCRIBE t_types_{seed}")[0][1]
        except R.RefError as e:
            ref = "ERR " + str(e).split(": ", 1)[-1]
        if mine == ref:
            ok += 1
        elif mine.startswith("ERR") and ref.startswith("ERR") and mine[:60] == ref[:60]:
            both += 1
        else:
            print("DIFF seed", seed, str(t)[:90], "| here:", mine[:90], "| reference:", ref[:90])
    print({"same_type": ok, "same_error": both})


def fuzz_plan(seed0, count):
    import duckdb_arrow_b200 as b2a
    fuzz_arrays.ALLOW_DICTIONARY = True
    ok, refused = 0, {}
    for seed in range(seed0, seed0 + count):
        rng = np.random.default_rng(seed)
        n = int(rng.choice([1, 63, 64, 65, 100, 2047, 2048, 2049, 3000, 5000]))
        table = pa.table({f"c{i}": fuzz_arrays.gen(rng, n) for i in range(int(rng.integers(1, 4)))})
        comp = [None, "zstd", "lz4"][int(rng.integers(0, 3))]
        sink = pa.BufferOutputStream()
        with ipc.new_stream(sink, table.schema, options=ipc.IpcWriteOptions(compression=comp)) as w:
            w.write_table(table, max_chunksize=int(rng.choice([500, 2048, 4000])))
        try:
            r = b2a.ArrowIPCReader.open_buffers([sink.getvalue().to_pybytes()])
            assert r.num_rows == n
            assert b2a.ArrowScan(r, device=-1).n_tiles >= 0  # device -1: plan only
            ok += 1
        except b2a.B2AError as e:
            refused[str(e)[:70]] = refused.get(str(e)[:70], 0) + 1
    print({"planned": ok, "refused": refused})


if __name__ == "__main__":
    mode, seed0, count = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    {"scan": fuzz_scan, "append": fuzz_append, "types": fuzz_types, "plan": fuzz_plan}[mode](seed0, count)
    sys.stdout.flush()
    os._exit(0)  # (the reference library's threads can keep the interpreter from exiting)
