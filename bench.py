#!/usr/bin/env python
"""bench.py — read_arrow throughput on TPC-H lineitem (BASELINE.json metric), B200 arm and reference (CPU) arm.

Workload (BASELINE.json configs[1]): lineitem SF10 as the reference would write it — DuckDB dbgen (the reference's own
tpch extension, compiled into oracle/_ref) -> the reference's ArrowAppender -> IPC stream of 122 880-row batches with
the appender's validity bitmaps (tools/dbgen_lineitem.py).  ONE stream; with --gpus N its record batches are sharded
contiguously over the ranks (strong scaling), nothing but row counts / aggregates crosses GPUs.

A "step" = one full read_arrow scan (all 16 columns, every record batch of the rank's shard).
  value        rows/s, whole job, stream image already resident in HBM (CUDA events, max over ranks)
  roofline     algorithmic bytes of one b2a_decode_kernel launch / its CUDA-event duration vs the measured HBM peak
  e2e          same scan through the C ABI from the FILE: mmap'ed stream -> pinned staging -> H2D -> decode -> D2H of the
               DuckDB vectors into pinned host memory, all inside the timed region (sub-legs: pinned image, Q6)
  legs         BASELINE configs[2..4]: filters + Q6 aggregate (sharded), multi-file glob with dictionary strings and
               union_by_name, COPY TO (writer) — each with its own roofline block
  cpu_baseline the reference's own arrow_scan (oracle/_ref) behind its BufferedFileReader, on a bounded sample

python bench.py [--gpus N] [--steps K] [--warmup W] [--sf 10] [--impl reference]
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

BATCH_ROWS = 122_880
# TPC-H Q6 (benchmark/lineitem.py:41) as pushed-down ConstantFilters over the decoded integer vectors
Q6_PREDS = [("l_shipdate", ">=", 8766), ("l_shipdate", "<", 9131), ("l_discount", ">=", 5), ("l_discount", "<=", 7),
            ("l_quantity", "<", 2400)]
Q6_COLUMNS = ["l_quantity", "l_extendedprice", "l_discount", "l_shipdate"]
Q6_SQL = ("SELECT sum(l_extendedprice * l_discount) FROM {} WHERE l_shipdate >= DATE '1994-01-01' AND "
          "l_shipdate < DATE '1995-01-01' AND l_discount BETWEEN 0.05 AND 0.07 AND l_quantity < 24")
Q6_KNOWN = {0.01: "1193053.2253", 1.0: "123141078.2283"}  # test/nodejs/arrow_test.js:424, benchmark/lineitem.py:41


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            p = [x.strip() for x in l.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                mx.append(float(p[1]))
            except ValueError:
                continue
            for n, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        # the sampler also sees the idle gaps between legs: the clock under load is the upper half of the samples
        load = sorted(sm)[len(sm) // 2:] if sm else []
        return {"sm_mhz": float(np.median(load)) if load else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------------- data
def stream_path(sf: float):
    """-> (path of the lineitem IPC stream, "dbgen" | "synthetic").  dbgen needs oracle/_ref (the reference's DuckDB)."""
    from tools import dbgen_lineitem as D
    if D.available():
        return D.lineitem_path(sf, BATCH_ROWS, log=log), "dbgen"
    # no compiled reference in this tree: numpy-drawn lineitem written by pyarrow (no validity bitmaps)
    import fcntl
    import pyarrow as pa
    from tools import lineitem
    rows = int(6_001_215 * sf) if sf not in (1.0, 10.0) else lineitem.ROWS_PER_SF[int(sf)]
    path = os.path.join(D.data_dir(), f"lineitem_synth_sf{sf:g}.arrows")
    with open(path + ".lock", "w") as lk:
        fcntl.flock(lk, fcntl.LOCK_EX)
        if not os.path.exists(path):
            with pa.OSFile(path + ".tmp", "wb") as sink:
                lineitem.write_stream(sink, rows, seed=42, unique=16)
            os.replace(path + ".tmp", path)
    return path, "synthetic"


def build_stream(rows: int, seed: int, pin: bool, sf: float | None = None):
    """The lineitem stream as ONE host tensor (tests / tools): the dbgen stream of the scale factor with `rows` rows
    when oracle/_ref is present, else the numpy-drawn stand-in.  -> (torch uint8 tensor, stream bytes)"""
    import torch
    from tools import dbgen_lineitem as D
    sf_of = {v: k for k, v in D.ROWS.items()}
    if D.available() and (sf is not None or rows in sf_of):
        path = D.lineitem_path(sf if sf is not None else sf_of[rows], BATCH_ROWS, log=log)
        size = os.path.getsize(path)
        host = torch.empty(size + 64, dtype=torch.uint8, pin_memory=pin)
        with open(path, "rb") as f:
            f.readinto(host.numpy()[:size])
        return host, size
    import pyarrow as pa
    from tools import lineitem
    cap = int(rows * 150) + (1 << 20)
    host = torch.empty(cap, dtype=torch.uint8, pin_memory=pin)
    sink = pa.FixedSizeBufferWriter(pa.py_buffer(host.numpy()))
    lineitem.write_stream(sink, rows, seed=seed, unique=16)
    return host, sink.tell()


def batch_index(path):
    from tools import dbgen_lineitem as D
    return D.message_index(path)


# ---------------------------------------------------------------------------------------------------------- CPU arm
def _ref_scan_messages(path, schema, msgs, validate=True):
    """The reference's read_arrow of the given record-batch messages on ONE thread:
       a1  IPCFileStreamReader: every message through DuckDB's BufferedFileReader (4 KiB bounce buffer), fresh
           allocation per message (src/ipc/stream_reader/ipc_file_stream_reader.cpp:47-138)      [reference, oracle/_ref]
       a2  slicing by pyarrow's IPC reader (nanoarrow is not on disk) + nanoarrow's FULL validation of the offsets
           (base_stream_reader.cpp:117-140)                                                       [port, oracle.c]
       a4-a9  ArrowScanFunction / ArrowToDuckDB over every 2 048-row chunk, all columns           [reference, oracle/_ref]"""
    import pyarrow as pa
    import pyarrow.ipc as ipc
    from oracle import oracle_py, ref_py
    f = ref_py.BufferedFile(path)
    rows = 0
    try:
        for off, meta, body in msgs:
            buf = np.empty(meta + body, np.uint8)
            f.read_into(off, meta + body, buf.ctypes.data)
            batch = ipc.read_record_batch(ipc.read_message(pa.py_buffer(buf)), schema)
            if validate:
                for col in batch.columns:
                    if pa.types.is_string(col.type) or pa.types.is_binary(col.type):
                        if not oracle_py.validate_full(col):
                            raise RuntimeError("offsets not monotone")
            r, _ = ref_py.scan_stream(batch)
            rows += r
    finally:
        f.close()
    return rows


def cpu_reference_rate(path, schema, msgs, seconds, threads):
    """rows/s of the reference read_arrow over `msgs`; threads > 1 = the messages dealt out as if they were a glob of
    `threads` files (the reference scans ONE thread per file, src/file_scanner/arrow_file_scan.cpp:35-41)."""
    from concurrent.futures import ThreadPoolExecutor
    parts = [msgs[i::threads] for i in range(threads)]
    parts = [p for p in parts if p]
    passes, rows, t0 = 0, 0, time.time()
    with ThreadPoolExecutor(len(parts)) as ex:
        while True:
            rows += sum(ex.map(lambda p: _ref_scan_messages(path, schema, p), parts))
            passes += 1
            if time.time() - t0 >= seconds:
                break
    dt = time.time() - t0
    return rows / dt, f"{len(msgs)} batches x {passes} passes ({rows} rows) in {dt:.1f}s on {len(parts)} threads"


def cpu_q6(path, schema, msgs, threads):
    """The reference's Q6 over read_arrow: messages read whole (a1), the four projected columns sliced (projection
    pushdown, read_arrow.cpp:44-50), arrow_scan + DuckDB's filter and aggregate above it.  -> (answer, seconds)"""
    import pyarrow as pa
    import pyarrow.ipc as ipc
    from oracle import ref_py
    proj = pa.schema([schema.field(n) for n in Q6_COLUMNS])

    def gen():
        f = ref_py.BufferedFile(path)
        try:
            for off, meta, body in msgs:
                buf = np.empty(meta + body, np.uint8)
                f.read_into(off, meta + body, buf.ctypes.data)
                b = ipc.read_record_batch(ipc.read_message(pa.py_buffer(buf)), schema)
                yield pa.RecordBatch.from_arrays([b.column(n) for n in Q6_COLUMNS], schema=proj)
        finally:
            f.close()

    ref_py.set_threads(threads)
    t0 = time.time()
    ref_py.register("b2a_bench_lineitem", pa.RecordBatchReader.from_batches(proj, gen()))
    ans = ref_py.query(Q6_SQL.format("b2a_bench_lineitem"))[0][0]
    dt = time.time() - t0
    ref_py.set_threads(1)
    return ans, dt


def cpu_port_rate(path, max_batches, seconds, threads):
    """Fallback when oracle/_ref is absent: the oracle port of the reference's read_arrow (kind "port")."""
    import pyarrow as pa
    from oracle import oracle_py as O
    O.lib()
    mm = pa.memory_map(path)
    msgs = O.stream_messages(mm.read_buffer(), max_batches)
    rows = sum(len(m[2]) for m in msgs)
    from concurrent.futures import ThreadPoolExecutor
    parts = [p for p in (msgs[i::threads] for i in range(threads)) if p]
    passes, t0 = 0, time.time()
    with ThreadPoolExecutor(len(parts)) as ex:
        while True:
            list(ex.map(O.read_arrow_checksum, parts))
            passes += 1
            if time.time() - t0 >= seconds:
                break
    dt = time.time() - t0
    return rows * passes / dt, f"{len(msgs)} batches ({rows} rows) x {passes} passes in {dt:.1f}s on {len(parts)} threads"


def reference_arm(args, path, data_kind, config, steps, warmup):
    from oracle import ref_py
    cores = os.cpu_count() or 1
    sample_batches = 64
    per = max(2.0, min(20.0, 100.0 / (steps + warmup + 2)))
    if ref_py.available():
        ref_py.lib()  # opens the DuckDB instance once, before any worker thread touches it
        from oracle import oracle_py
        oracle_py.lib()
        schema, idx = batch_index(path)
        msgs = idx[:sample_batches]
        kind = "reference"
        what = ("the reference's own converters compiled from /root/reference/duckdb (oracle/_ref): BufferedFileReader "
                "message reads (a1) + pyarrow slicing with the port of nanoarrow's FULL offsets validation (a2) + "
                "ArrowScanFunction/ArrowToDuckDB over every chunk of all 16 columns (a4-a9); ")

        def rate(seconds, threads):
            return cpu_reference_rate(path, schema, msgs, seconds, threads)
    else:
        kind = "port"
        what = "oracle port of the reference read_arrow (oracle/_ref not built); "

        def rate(seconds, threads):
            return cpu_port_rate(path, sample_batches, seconds, threads)
    for _ in range(warmup):
        rate(1.0, cores)
    vals, sample = [], ""
    t0 = time.time()
    for _ in range(steps):
        v, sample = rate(per, cores)
        vals.append(v)
    dt = (time.time() - t0) / steps
    v = float(np.mean(vals))
    one, one_sample = rate(per, 1)
    q6 = None
    if kind == "reference":
        try:
            ans, q6_s = cpu_q6(path, schema, msgs, cores)
            ans1, q6_s1 = cpu_q6(path, schema, msgs, 1)
            q6_rows = len(msgs) * BATCH_ROWS
            q6 = {"value": q6_rows / q6_s, "unit": "rows/s", "threads": cores, "seconds": q6_s, "answer_on_sample": ans,
                  "single_thread_value": q6_rows / q6_s1,
                  "what": "SELECT sum(l_extendedprice*l_discount) ... (benchmark/lineitem.py:41) through DuckDB over "
                          "arrow_scan of the projected columns, messages read through BufferedFileReader; "
                          f"{len(msgs)} batches"}
        except Exception as e:  # noqa: BLE001
            q6 = {"error": repr(e)}
    print(json.dumps({
        "impl": "reference", "metric": "read_arrow rows/s (lineitem, all columns)", "value": v, "unit": "rows/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": data_kind,
        "config": config,
        "cpu_baseline": {"value": v, "unit": "rows/s", "cores": cores, "kind": kind,
                         "sample": what + "the stream's messages dealt out over all host threads as if they were a "
                                   "glob of files (the reference gives ONE file one scan thread), " + sample,
                         "single_file_value": one, "single_file_cores": 1,
                         "single_file_note": "what the reference does on ONE file (arrow_file_scan.cpp:35-41); " + one_sample,
                         "q6": q6},
        "e2e": {"value": v, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ---------------------------------------------------------------------------------------------------------- B200 arm
class Ctx:
    pass


def _barrier(ctx):
    ctx.torch.cuda.synchronize()
    if ctx.dist:
        ctx.dist.barrier()
    ctx.torch.cuda.synchronize()


def _max_over_ranks(ctx, x: float) -> float:
    t = ctx.torch.tensor([x], dtype=ctx.torch.float64, device=ctx.dev)
    if ctx.dist:
        ctx.dist.all_reduce(t, op=ctx.dist.ReduceOp.MAX)
    return float(t.item())


def _sum_over_ranks(ctx, x: float) -> float:
    t = ctx.torch.tensor([x], dtype=ctx.torch.float64, device=ctx.dev)
    if ctx.dist:
        ctx.dist.all_reduce(t, op=ctx.dist.ReduceOp.SUM)
    return float(t.item())


def _time_steps(ctx, fn, steps, warmup):
    """CUDA events on the current stream around `steps` calls of fn(), after `warmup`; -> (ms per step max over ranks,
    mean of the per-step event times of this rank)."""
    torch = ctx.torch
    stream = torch.cuda.current_stream()
    for _ in range(warmup):
        fn()
    _barrier(ctx)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    a0, z0 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record(stream)
    for a, z in evs:
        a.record(stream)
        fn()
        z.record(stream)
    z0.record(stream)
    _barrier(ctx)
    return _max_over_ranks(ctx, a0.elapsed_time(z0)) / steps, float(np.mean([a.elapsed_time(z) for a, z in evs]))


def _i128(lo_hi) -> int:
    lo, hi = [int(x) for x in lo_hi]
    return (hi << 64) | (lo & 0xFFFFFFFFFFFFFFFF)


def _combine_q6(ctx, total_t, sum_t, rows):
    """The ONE collective of the path: all_reduce(SUM) of {rows, selected, 128-bit sum as four 32-bit limbs in int64
    lanes} on the device; read back once."""
    torch = ctx.torch
    lo, hi = sum_t[0], sum_t[1]
    m = 0xFFFFFFFF
    lanes = torch.stack([torch.tensor(rows, dtype=torch.int64, device=ctx.dev), total_t.reshape(()).to(torch.int64),
                         lo & m, (lo >> 32) & m, hi & m, (hi >> 32) & m])
    if ctx.dist:
        ctx.dist.all_reduce(lanes, op=ctx.dist.ReduceOp.SUM)
    v = [int(x) for x in lanes.cpu().tolist()]
    u = (v[2] + (v[3] << 32) + (v[4] << 64) + (v[5] << 96)) & ((1 << 128) - 1)
    s = u - (1 << 128) if u >> 127 else u
    return v[0], v[1], s


def _decimal_str(v: int, scale: int) -> str:
    sign = "-" if v < 0 else ""
    v = abs(v)
    return f"{sign}{v // 10 ** scale}.{v % 10 ** scale:0{scale}d}"


def _q6_device_variant(ctx, path, steps, device_alias, fused=False):
    """decode -> filter -> sum(a*b) over this rank's shard as one CUDA graph.  device_alias: B2A_SCAN_DEVICE_ALIAS — the
    four fixed-width columns are read in place in the device image (the reference's zero-copy DirectConversion for device
    consumers): nothing to decode, the graph holds the filter and the aggregate only.  fused: b2a_scan_filter_sum_product —
    count and sum in ONE kernel, no selection vectors in between."""
    torch, b2a = ctx.torch, ctx.b2a
    reader = b2a.ArrowIPCReader.open_file(path)
    reader.set_projection(Q6_COLUMNS)
    lo, hi = ctx.P.batch_range(reader.num_batches, ctx.rank, ctx.world)
    scan = b2a.ArrowScan(reader, device=ctx.local_rank, batch_lo=lo, batch_hi=hi, device_alias=device_alias)
    my_rows = sum(reader.batch_info(b).rows for b in range(lo, hi))
    image = reader.image()
    d_image = torch.empty(scan.image_bytes + 256, dtype=torch.uint8, device=ctx.dev)
    d_image[:scan.image_bytes].copy_(torch.from_numpy(image[scan.image_lo:scan.image_lo + scan.image_bytes]))
    scan.set_device_image(d_image)
    d_out = torch.empty(max(scan.out_bytes, 256), dtype=torch.uint8, device=ctx.dev)
    res = torch.zeros(2, dtype=torch.int64, device=ctx.dev)
    scan.decode(d_out)
    if fused:                                         # first call builds + caches the per-chunk column tables
        total_t, res = scan.filter_sum_product(Q6_PREDS, "l_extendedprice", "l_discount", d_out)
        fout = (None, None, total_t)
    else:
        fout = scan.filter(Q6_PREDS, d_out)
        scan.sum_product_async("l_extendedprice", "l_discount", fout[0], fout[1], d_out, res)
    scan.finish()
    decodes = scan.n_tiles > 0

    def body():
        if decodes:
            scan.decode(d_out)
        if fused:
            scan.filter_sum_product(Q6_PREDS, "l_extendedprice", "l_discount", d_out, out=(fout[2], res))
        else:
            scan.filter(Q6_PREDS, d_out, out=fout)
            scan.sum_product_async("l_extendedprice", "l_discount", fout[0], fout[1], d_out, res)

    graph = None
    side = torch.cuda.Stream()
    try:
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            body()
            side.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = g
    except Exception as e:  # noqa: BLE001 - capture is an optimisation; the stream-launched body is the same work
        log(f"[bench] rank {ctx.rank}: CUDA graph capture failed ({e!r}); launching the kernels per step")
        torch.cuda.synchronize()
    fn = graph.replay if graph is not None else body
    ms, _ = _time_steps(ctx, fn, steps, 3)
    rows, selected, total = _combine_q6(ctx, fout[2], res, my_rows)
    scan.finish()
    # decode (if any) + the predicate columns (4 + 8 + 8 B/row) + selection entries and the two factors of the selected rows
    alg = (scan.algorithmic_bytes if decodes else 0) + my_rows * 20 + int(fout[2].item()) * ((0 if fused else 4) + 16)
    kernels = (["b2a_decode_kernel"] if decodes else []) + \
        (["b2a_filter_sum_kernel"] if fused else ["b2a_filter_kernel", "b2a_sum_product_kernel"])
    traffic, traffic_source = None, None
    if not decodes:  # the filter kernel is (nearly) the whole step: its DRAM bytes from the builder's ncu capture
        try:
            tf = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["b2a_filter_kernel"]
            traffic = int((tf["dram_bytes_read"] + tf["dram_bytes_write"]) / tf["rows_per_launch"] * my_rows)
            traffic_source = ("profiles/traffic.json: dram__bytes_read+write of b2a_filter_kernel (ncu --set full, SF10), per "
                              "row, x the rows of this launch (not measured in this run)")
        except Exception:  # noqa: BLE001
            pass
    r = {"ms_per_step": ms, "value": rows / (ms / 1e3), "rows": rows, "selected": selected, "sum": _decimal_str(total, 4),
         "cuda_graph": graph is not None, "kernels_per_step": len(kernels),
         "roofline": {"bound": "hbm", "kernels": kernels, "achieved": alg / (ms / 1e3) / 1e9, "peak": ctx.peak,
                      "unit": "GB/s", "frac": alg / (ms / 1e3) / 1e9 / ctx.peak,
                      "algorithmic_bytes_per_step_rank0": int(alg), "traffic": traffic,
                      "traffic_source": traffic_source}}
    del d_image, d_out
    scan.close()
    reader.close()
    return r


def leg_filter_q6(ctx, args, path, steps):
    """BASELINE configs[2]: read_arrow with pushed-down filters + the Q6 aggregate, batches of ONE stream sharded over the
    ranks, one CUDA graph per rank, no host read-back inside the timed loop, one all_reduce after it.  `value`: the four
    columns read in place in the device image (B2A_SCAN_DEVICE_ALIAS — they are DirectConversion columns, which the
    reference does not copy either); `materialised`: the same with the vectors copied into the arena first (what rounds
    1 and 2 reported before the flag existed)."""
    fused = _q6_device_variant(ctx, path, steps, device_alias=True, fused=True)
    two = _q6_device_variant(ctx, path, steps, device_alias=True)
    mat = _q6_device_variant(ctx, path, steps, device_alias=False)
    # `value` = the faster of the two zero-copy forms (one fused kernel wins where the step is launch-bound, i.e. on a small
    # shard; filter + aggregate as two kernels at full occupancy wins on a whole SF10), both reported
    zero, variant = (fused, "fused") if fused["ms_per_step"] <= two["ms_per_step"] else (two, "two_kernels")
    keys = ("value", "ms_per_step", "kernels_per_step", "cuda_graph", "roofline")
    out = {"metric": "read_arrow + pushed-down Q6 filters + sum(l_extendedprice*l_discount), rows/s (device-resident)",
           "value": zero["value"], "unit": "rows/s", "ms_per_step": zero["ms_per_step"], "rows": zero["rows"],
           "selected": zero["selected"], "selectivity": zero["selected"] / max(1, zero["rows"]), "sum": zero["sum"],
           "scaling": "strong", "cuda_graph": zero["cuda_graph"], "kernels_per_step": zero["kernels_per_step"],
           "variant": variant,
           "zero_copy": "B2A_SCAN_DEVICE_ALIAS: fixed-width columns are consumed in place in the device image",
           "collective": "one all_reduce(SUM) of 6 int64 lanes after the timed loop (rows, selected, 128-bit sum limbs)",
           "roofline": zero["roofline"],
           "zero_copy_fused": {k: fused[k] for k in keys},
           "zero_copy_two_kernels": {k: two[k] for k in keys},
           "materialised": {k: mat[k] for k in keys},
           "same_answer_every_way": bool(fused["sum"] == mat["sum"] == two["sum"]
                                         and fused["selected"] == mat["selected"] == two["selected"]),
           "config": {"workload": f"lineitem SF{args.sf:g}, columns {Q6_COLUMNS}, predicates {Q6_PREDS}"}}
    known = Q6_KNOWN.get(float(args.sf))
    if known is not None and ctx.data_kind == "dbgen":
        out["known_answer"] = known
        out["known_answer_ok"] = bool(out["sum"] == known)
    return out


def leg_e2e_q6(ctx, args, path, steps):
    """Q6 end to end from the FILE: only the buffers of the 4 projected columns cross PCIe (b2a_read_arrow_device:
    staged H2D behind the decode launches), filter + aggregate on the device, 24 bytes come back."""
    torch, b2a = ctx.torch, ctx.b2a
    reader = b2a.ArrowIPCReader.open_file(path)
    reader.set_projection(Q6_COLUMNS)
    lo, hi = ctx.P.batch_range(reader.num_batches, ctx.rank, ctx.world)
    scan = b2a.ArrowScan(reader, device=ctx.local_rank, batch_lo=lo, batch_hi=hi)
    my_rows = sum(reader.batch_info(b).rows for b in range(lo, hi))
    d_out = torch.empty(scan.out_bytes, dtype=torch.uint8, device=ctx.dev)
    res = torch.zeros(2, dtype=torch.int64, device=ctx.dev)
    back = torch.empty(3, dtype=torch.int64, pin_memory=True)
    fout = None

    def one():
        nonlocal fout
        scan.read_device(d_out)
        fout = scan.filter(Q6_PREDS, d_out, out=fout)
        scan.sum_product_async("l_extendedprice", "l_discount", fout[0], fout[1], d_out, res)
        back[:2].copy_(res, non_blocking=True)
        back[2:].copy_(fout[2], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return _i128(back[:2].tolist()), int(back[2])

    one()
    one()
    _barrier(ctx)
    t0 = time.perf_counter()
    for _ in range(steps):
        total, selected = one()
    dt = _max_over_ranks(ctx, time.perf_counter() - t0) / steps
    rows = int(_sum_over_ranks(ctx, my_rows))
    out = {"value": rows / dt, "unit": "rows/s", "ms_per_step": dt * 1e3, "steps": steps,
           "h2d_bytes_per_step": int(scan.upload_bytes), "d2h_bytes_per_step": 24,
           "rank0_selected": selected, "rank0_sum": _decimal_str(total, 4),
           "what": "b2a_read_arrow_device on a 4-column projection of the mmap'ed file (staged H2D of the projected "
                   "buffers only, decode per slice) + b2a_scan_filter (5 predicates) + b2a_scan_sum_product; 24 bytes "
                   "read back per step"}
    scan.finish()
    del d_out
    scan.close()
    reader.close()
    return out


def leg_copy(ctx, args, scan, reader, d_out, my_rows, steps):
    """BASELINE configs[4]: COPY lineitem TO (FORMAT ARROWS) — the rank's decoded DuckDB vectors (HBM) -> the complete
    IPC stream of its shard (HBM): pass 1 (create + measure) by wall clock, pass 2 (b2a_encode_kernel) by CUDA events."""
    torch, b2a = ctx.torch, ctx.b2a
    cols = []
    for f in reader.columns:
        parts = []
        for b in range(scan.batch_lo, scan.batch_hi):
            co = scan.column(b, f.index)
            parts.append(d_out[co.data_off: co.data_off + co.rows * f.out_width])
        cols.append(b2a.DeviceVector(f.type, my_rows, torch.cat(parts), None, name=f.name,
                                     decimal=(f.decimal_width, f.decimal_scale)))
    torch.cuda.synchronize()

    def new_writer():
        return b2a.ArrowWriter(cols, my_rows, BATCH_ROWS, device=ctx.local_rank)

    w = new_writer()
    w.measure()
    d_stream = torch.empty(w.stream_bytes, dtype=torch.uint8, device=ctx.dev)
    ms, enc_ms = _time_steps(ctx, lambda: w.encode(d_stream), steps, 3)
    w.finish()
    m = []
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        w2 = new_writer()
        w2.measure()
        torch.cuda.synchronize()
        m.append((time.perf_counter() - t0) * 1e3)
        w2.close()
    measure_ms = _max_over_ranks(ctx, float(np.median(m)))
    rows = int(_sum_over_ranks(ctx, my_rows))
    alg = w.algorithmic_bytes
    traffic, traffic_source = None, None
    try:
        te = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["b2a_encode_kernel"]
        traffic = int((te["dram_bytes_read"] + te["dram_bytes_write"]) / 59986052 * my_rows)
        traffic_source = ("profiles/traffic.json: dram__bytes_read+write of b2a_encode_kernel from the builder's ncu --set "
                          "full capture on SF10, per row, x the rows of this launch (not measured in this run)")
    except Exception:  # noqa: BLE001
        pass
    out = {"metric": "COPY lineitem TO (FORMAT ARROWS) rows/s (device-resident vectors -> IPC stream in HBM)",
           "value": rows / ((ms + measure_ms) / 1e3), "unit": "rows/s", "encode_only_value": rows / (ms / 1e3),
           "encode_ms": ms, "measure_ms": measure_ms, "rows": rows, "n_batches_rank0": int(w.n_batches),
           "stream_bytes_rank0": int(w.stream_bytes), "scaling": "strong",
           "roofline": {"bound": "hbm", "kernel": "b2a_encode_kernel", "achieved": alg / (enc_ms / 1e3) / 1e9,
                        "peak": ctx.peak, "unit": "GB/s", "frac": alg / (enc_ms / 1e3) / 1e9 / ctx.peak,
                        "kernel_ms": enc_ms, "algorithmic_bytes_per_launch": int(alg), "traffic": traffic,
                        "traffic_source": traffic_source,
                        "algorithmic_bytes_note": "input vectors (156 B/row) + characters of the strings longer than 12 "
                                                  "bytes fetched from the string heap + every byte of the produced bodies"},
           "config": {"workload": f"lineitem SF{args.sf:g} COPY TO arrows, row_group_size {BATCH_ROWS}"}}
    w.close()
    del d_stream
    # the SET produce_arrow_string_view = true variant of the same COPY (ArrowVarcharToStringViewData)
    try:
        wv = b2a.ArrowWriter(cols, my_rows, BATCH_ROWS, device=ctx.local_rank, produce_arrow_string_view=True)
        wv.measure()
        d_sv = torch.empty(wv.stream_bytes, dtype=torch.uint8, device=ctx.dev)
        ms_v, enc_v = _time_steps(ctx, lambda: wv.encode(d_sv), min(steps, 3), 3)
        wv.finish()
        out["string_view"] = {"encode_ms": ms_v, "algorithmic_bytes_per_launch": int(wv.algorithmic_bytes),
                              "achieved_gbs": wv.algorithmic_bytes / (enc_v / 1e3) / 1e9,
                              "frac": wv.algorithmic_bytes / (enc_v / 1e3) / 1e9 / ctx.peak,
                              "vs_offsets_writer": ms_v / ms}
        wv.close()
        del d_sv
    except Exception as e:  # noqa: BLE001
        out["string_view"] = {"error": repr(e)}
    del cols
    return out


def _multifile_inputs(ctx, path, n_files, rows_per_file):
    """Two schema variants of a (SF100/64)-row lineitem file with the four low-cardinality string columns dictionary
    encoded (the reference cannot read dictionary batches: base_stream_reader.cpp:87-96): variant B has another column
    order and an INTEGER l_linenumber, so union_by_name maps by name and casts on the device.  Built from the dbgen
    stream's own batches (cycled), once per box."""
    import fcntl
    import pyarrow as pa
    import pyarrow.ipc as ipc
    from tools import dbgen_lineitem as D
    d = os.path.join(D.data_dir(), f"multifile_{rows_per_file}")
    glob_dir = os.path.join(d, "glob")
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, ".lock"), "w") as lk:
        fcntl.flock(lk, fcntl.LOCK_EX)
        if not os.path.exists(os.path.join(d, ".done")):
            dict_cols = {"l_returnflag": pa.int8(), "l_linestatus": pa.int8(), "l_shipinstruct": pa.int32(),
                         "l_shipmode": pa.int32()}
            src = ipc.open_stream(pa.memory_map(path))
            base = src.schema
            sch_a = pa.schema([pa.field(f.name, pa.dictionary(dict_cols[f.name], pa.string())) if f.name in dict_cols
                               else f for f in base])
            order_b = [base.names[i] for i in (15, 3, 0, 1, 2, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14)]
            sch_b = pa.schema([pa.field("l_linenumber", pa.int32()) if n == "l_linenumber" else sch_a.field(n)
                               for n in order_b])
            # one dictionary per column for the whole file (a new dictionary per batch would be a replacement)
            dicts = {"l_returnflag": pa.array(["A", "N", "R"]), "l_linestatus": pa.array(["F", "O"]),
                     "l_shipinstruct": pa.array(["COLLECT COD", "DELIVER IN PERSON", "NONE", "TAKE BACK RETURN"]),
                     "l_shipmode": pa.array(["AIR", "FOB", "MAIL", "RAIL", "REG AIR", "SHIP", "TRUCK"])}
            import pyarrow.compute as pc
            batches = []
            for i, b in enumerate(src):
                if i >= 16:
                    break
                arrs = []
                for f in base:
                    c = b.column(f.name)
                    if f.name in dict_cols:
                        idx = pc.index_in(c, value_set=dicts[f.name]).cast(dict_cols[f.name])
                        c = pa.DictionaryArray.from_arrays(idx, dicts[f.name])
                    arrs.append(c)
                batches.append(pa.RecordBatch.from_arrays(arrs, schema=sch_a))
            fa, fb = os.path.join(d, "variant_a.arrows"), os.path.join(d, "variant_b.arrows")
            with pa.OSFile(fa, "wb") as oa, pa.OSFile(fb, "wb") as ob:
                with ipc.new_stream(oa, sch_a) as wa, ipc.new_stream(ob, sch_b) as wb:
                    left, k = rows_per_file, 0
                    while left > 0:
                        b = batches[k % len(batches)]
                        k += 1
                        if b.num_rows > left:
                            b = b.slice(0, left)
                        left -= b.num_rows
                        wa.write_batch(b)
                        cols = [b.column(n) for n in order_b]
                        cols[1] = cols[1].cast(pa.int32())
                        wb.write_batch(pa.RecordBatch.from_arrays(cols, schema=sch_b))
            os.makedirs(glob_dir, exist_ok=True)
            for i in range(64):
                p = os.path.join(glob_dir, f"lineitem_{i:03d}.arrows")
                if not os.path.lexists(p):
                    os.symlink(fb if i % 2 else fa, p)
            open(os.path.join(d, ".done"), "w").close()
    return sorted(os.path.join(glob_dir, f) for f in os.listdir(glob_dir))[:n_files]


def leg_multifile(ctx, args, path, steps):
    """BASELINE configs[3]: multi-file glob of lineitem files (SF100/64 rows each), union_by_name, dictionary-encoded
    strings; files of the glob sharded over the ranks (8 per GPU), device-resident, launches round-robin over 4 streams."""
    torch, b2a = ctx.torch, ctx.b2a
    from duckdb_arrow_b200._lib import check, lib
    files_per_gpu, rows_per_file = args.files_per_gpu, int(600_037_902 // 64 * min(1.0, args.sf / 10.0))
    paths = _multifile_inputs(ctx, path, files_per_gpu * ctx.world, rows_per_file)
    mine = ctx.P.file_shard(paths, ctx.rank, ctx.world)
    mfs = b2a.api.MultiFileScan(paths, device=ctx.local_rank, union_by_name=True)   # global bind: union over ALL files
    cur = torch.cuda.current_stream()
    files, alg, my_rows = [], 0, 0
    for p in mine:
        f = mfs.files[paths.index(p)]
        sc = b2a.ArrowScan(f.reader, device=ctx.local_rank)
        sc.upload()
        d_o = torch.empty(sc.out_bytes, dtype=torch.uint8, device=ctx.dev)
        fin = C.c_void_p()
        check(lib.b2a_multifile_finalize_create(mfs._h, f.index, sc._h, C.byref(fin)))
        fb, nc = C.c_uint64(), C.c_uint64()
        check(lib.b2a_multifile_finalize_sizes(fin, C.byref(fb), C.byref(nc)))
        d_f = torch.empty(max(fb.value, 256), dtype=torch.uint8, device=ctx.dev)
        files.append((sc, d_o, fin, d_f, nc.value))
        alg += sc.algorithmic_bytes + 2 * fb.value   # casts: source vector read + result written (same width class)
        my_rows += f.reader.num_rows
    side = [torch.cuda.Stream(device=ctx.dev) for _ in range(4)]

    def step():
        for st in side:
            st.wait_stream(cur)
        for i, (sc, d_o, fin, d_f, n_casts) in enumerate(files):
            st = side[i % len(side)]
            with torch.cuda.stream(st):
                sc.decode(d_o)
                if n_casts:
                    check(lib.b2a_multifile_finalize_run(fin, C.c_void_p(d_o.data_ptr()), C.c_void_p(d_f.data_ptr()), 0, 0,
                                                         C.c_void_p(st.cuda_stream)))
        for st in side:
            cur.wait_stream(st)

    ms, _ = _time_steps(ctx, step, steps, 3)
    rows = int(_sum_over_ranks(ctx, my_rows))
    out = {"metric": "multi-file read_arrow rows/s (glob, union_by_name, dictionary strings; device-resident)",
           "value": rows / (ms / 1e3), "unit": "rows/s", "ms_per_step": ms, "rows": rows, "files": len(paths),
           "files_per_gpu": files_per_gpu, "rows_per_file": rows_per_file, "scaling": "weak",
           "roofline": {"bound": "hbm", "kernels": ["b2a_decode_kernel", "b2a_cast_kernel"],
                        "achieved": alg / (ms / 1e3) / 1e9, "peak": ctx.peak, "unit": "GB/s",
                        "frac": alg / (ms / 1e3) / 1e9 / ctx.peak, "algorithmic_bytes_per_step_rank0": int(alg),
                        "traffic": None},
           "config": {"workload": "lineitem files of SF100/64 rows, two schema variants (column order, INTEGER vs BIGINT "
                                  "l_linenumber), 4 dictionary-encoded string columns, union_by_name"}}
    for sc, d_o, fin, d_f, _ in files:
        sc.finish()
        check(lib.b2a_multifile_finalize_finish(fin, C.c_void_p(cur.cuda_stream)))
        lib.b2a_multifile_finalize_destroy(fin)
        sc.close()
    mfs.close()
    return out


def b200_arm(args, path, data_kind, config, steps, warmup):
    import torch
    import duckdb_arrow_b200 as b2a
    from duckdb_arrow_b200 import parallel as P
    from duckdb_arrow_b200._lib import check, lib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the scan path has no CPU fallback")
    ctx = Ctx()
    ctx.torch, ctx.b2a, ctx.P = torch, b2a, P
    ctx.rank = int(os.environ.get("RANK", "0"))
    ctx.world = int(os.environ.get("WORLD_SIZE", "1"))
    ctx.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    ctx.dev = f"cuda:{ctx.local_rank}"
    ctx.data_kind = data_kind
    ctx.peak, peak_kind = measured_peak()
    torch.cuda.set_device(ctx.local_rank)
    numa = P.bind_to_device_numa_node(ctx.local_rank) if os.environ.get("B2A_NUMA_BIND", "1") != "0" else {}
    log(f"[bench] rank {ctx.rank}: numa {numa}, {len(os.sched_getaffinity(0))} cpus")
    ctx.dist = None
    if ctx.world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device(ctx.dev))
        ctx.dist = dist
    rank, world = ctx.rank, ctx.world

    reader = b2a.ArrowIPCReader.open_file(path)                      # mmap of the file: what read_arrow('x.arrows') opens
    if args.columns:
        reader.set_projection(args.columns.split(","))
        config["columns"] = args.columns
    lo, hi = P.batch_range(reader.num_batches, rank, world)
    scan = b2a.ArrowScan(reader, device=ctx.local_rank, batch_lo=lo, batch_hi=hi)
    my_rows = sum(reader.batch_info(b).rows for b in range(lo, hi))
    total_rows = reader.num_rows
    config.update({"rows": total_rows, "batches": reader.num_batches, "rows_this_rank0": my_rows})
    log(f"[bench] rank {rank}: batches [{lo},{hi}) = {my_rows} rows, image {scan.image_bytes / 1e9:.2f} GB, out "
        f"{scan.out_bytes / 1e9:.2f} GB, algorithmic {scan.algorithmic_bytes / 1e9:.2f} GB "
        f"({scan.algorithmic_bytes / max(1, my_rows):.1f} B/row), {scan.n_tasks} tasks, {scan.n_tiles} tiles")
    image = reader.image()
    d_image = torch.empty(scan.image_bytes + 256, dtype=torch.uint8, device=ctx.dev)
    d_image[:scan.image_bytes].copy_(torch.from_numpy(image[scan.image_lo:scan.image_lo + scan.image_bytes]))
    d_out = torch.empty(scan.out_bytes, dtype=torch.uint8, device=ctx.dev)
    scan.set_device_image(d_image)

    # ---- device-resident decode (the headline `value`)
    sampler = ClockSampler(ctx.local_rank)
    sampler.start()
    launches0 = scan.gpu_launches
    ms_per_step, kernel_ms = _time_steps(ctx, lambda: scan.decode(d_out), steps, warmup)
    scan.finish()
    launches = (scan.gpu_launches - launches0) * steps // (steps + warmup)
    value = total_rows / (ms_per_step / 1e3)
    log(f"[bench] decode: {ms_per_step:.3f} ms/step (max over ranks), rank {rank} kernel {kernel_ms:.3f} ms")

    legs = {}
    if not args.profile and not args.no_legs:
        k = max(3, min(steps, 10))
        for name, fn in (("filter_q6", lambda: leg_filter_q6(ctx, args, path, k)),
                         ("copy", lambda: leg_copy(ctx, args, scan, reader, d_out, my_rows, min(k, 5))),
                         ("multifile", lambda: leg_multifile(ctx, args, path, k))):
            if name in args.skip_legs.split(","):
                continue
            try:
                t0 = time.time()
                if name == "multifile":   # frees the big buffers of the main scan first? no: 180 GB is plenty
                    pass
                legs[name] = fn()
                log(f"[bench] leg {name}: {legs[name]['value'] / 1e9:.2f} G rows/s, {legs[name]['ms_per_step'] if 'ms_per_step' in legs[name] else legs[name]['encode_ms']:.3f} ms "
                    f"({time.time() - t0:.1f}s)")
            except Exception as e:  # noqa: BLE001 - a failing leg must not take the headline line with it
                legs[name] = {"error": repr(e)}
                log(f"[bench] leg {name} FAILED: {e!r}")
                torch.cuda.synchronize()
            _barrier(ctx)
        if world > 1:
            # every rank scanning the WHOLE stream (what round 1 reported): weak-scaling sub-leg
            try:
                sc_all = b2a.ArrowScan(reader, device=ctx.local_rank)
                di = torch.empty(sc_all.image_bytes + 256, dtype=torch.uint8, device=ctx.dev)
                di[:sc_all.image_bytes].copy_(torch.from_numpy(image[sc_all.image_lo:sc_all.image_lo + sc_all.image_bytes]))
                do = torch.empty(sc_all.out_bytes, dtype=torch.uint8, device=ctx.dev)
                sc_all.set_device_image(di)
                ms_all, _ = _time_steps(ctx, lambda: sc_all.decode(do), max(3, min(steps, 5)), 3)
                sc_all.finish()
                legs["per_gpu_full_scan"] = {"value": total_rows * world / (ms_all / 1e3), "unit": "rows/s",
                                             "ms_per_step": ms_all, "scaling": "weak",
                                             "what": "every rank decodes the whole SF stream (no sharing)"}
                sc_all.close()
                del di, do
            except Exception as e:  # noqa: BLE001
                legs["per_gpu_full_scan"] = {"error": repr(e)}
            _barrier(ctx)

    # ---- e2e: the FILE -> host vectors through the C ABI (headline e2e), + pinned image, + Q6 with device consumers
    e2e = None
    if not args.no_e2e and not args.profile:
        e_steps = max(2, min(steps, 5))

        def run_e2e(rd, host_alias, blo, bhi, reopen_path=None):
            """reopen_path: every step is the WHOLE call a user makes — open the file (mmap, footer / message index),
            plan the scan, run it, close both; only the pinned output arena is kept across steps."""
            sc = b2a.ArrowScan(rd, device=ctx.local_rank, batch_lo=blo, batch_hi=bhi, host_alias=host_alias)
            h_out = torch.empty(sc.out_bytes, dtype=torch.uint8, pin_memory=True)
            launches = [0]

            def one():
                if reopen_path is None:
                    check(lib.b2a_read_arrow_host(sc._h, C.c_void_p(h_out.data_ptr()), args.slices))
                    return
                r2 = b2a.ArrowIPCReader.open_file(reopen_path)
                s2 = b2a.ArrowScan(r2, device=ctx.local_rank, batch_lo=blo, batch_hi=bhi, host_alias=host_alias)
                check(lib.b2a_read_arrow_host(s2._h, C.c_void_p(h_out.data_ptr()), args.slices))
                launches[0] += s2.gpu_launches
                s2.close()
                r2.close()

            one()
            one()
            _barrier(ctx)
            l0 = sc.gpu_launches + launches[0]
            t0 = time.perf_counter()
            for _ in range(e_steps):
                one()
            torch.cuda.synchronize()
            dt = _max_over_ranks(ctx, time.perf_counter() - t0) / e_steps
            res = {"value": total_rows / dt, "unit": "rows/s", "h2d_bytes_per_step": int(sc.upload_bytes),
                   "d2h_bytes_per_step": int(sc.out_bytes), "steps": e_steps, "ms_per_step": dt * 1e3,
                   "launches_per_step": (sc.gpu_launches + launches[0] - l0) // e_steps,
                   "includes_open_and_plan": reopen_path is not None}
            sc.close()
            del h_out
            return res

        e2e = run_e2e(reader, True, lo, hi, reopen_path=path)
        e2e["what"] = ("read_arrow('lineitem.arrows') through the C ABI, the whole call every step (open + index the file, "
                       "plan, scan, close): mmap'ed file -> host threads pack the buffers that "
                       "need conversion into a pinned ring -> H2D -> b2a_scatter_kernel + b2a_decode_kernel -> D2H of the "
                       "converted vectors (pinned); fixed-width columns alias the host image (reference DirectConversion, "
                       "arrow_conversion.cpp:454-461); the reference arm reads the same file through BufferedFileReader")
        log(f"[bench] e2e file: {e2e['ms_per_step']:.1f} ms/step")
        # the shard as a stream of its own in PINNED memory (what scan_arrow_ipc gets from a client that pins)
        try:
            _, idx = batch_index(path)
            first = idx[0][0]
            a, z = idx[lo][0], idx[hi - 1][0] + idx[hi - 1][1] + idx[hi - 1][2]
            size = first + (z - a) + 8
            pinned = torch.empty(size, dtype=torch.uint8, pin_memory=True)
            pn = pinned.numpy()
            pn[:first] = image[:first]
            pn[first:first + (z - a)] = image[a:z]
            pn[first + (z - a):] = np.frombuffer(b"\xff\xff\xff\xff\0\0\0\0", np.uint8)
            rd2 = b2a.ArrowIPCReader.open_buffers([(pinned.data_ptr(), size)])
            sub = run_e2e(rd2, True, 0, rd2.num_batches)
            e2e["pinned_image"] = {k: sub[k] for k in ("value", "ms_per_step", "h2d_bytes_per_step", "d2h_bytes_per_step")}
            full = run_e2e(rd2, False, 0, rd2.num_batches)
            e2e["pinned_image_full_copy"] = {k: full[k] for k in ("value", "ms_per_step", "h2d_bytes_per_step",
                                                                  "d2h_bytes_per_step")}
            rd2.close()
            del pinned, pn
        except Exception as e:  # noqa: BLE001
            e2e["pinned_image"] = {"error": repr(e)}
        try:
            e2e["q6"] = leg_e2e_q6(ctx, args, path, e_steps)
            log(f"[bench] e2e q6: {e2e['q6']['ms_per_step']:.1f} ms/step")
        except Exception as e:  # noqa: BLE001
            e2e["q6"] = {"error": repr(e)}
            torch.cuda.synchronize()

    # keep the GPU busy for a moment so that the 100 ms nvidia-smi sampler sees clocks under load
    t_busy = time.time()
    while time.time() - t_busy < 1.0 and not args.profile:
        for _ in range(20):
            scan.decode(d_out)
        torch.cuda.synchronize()
    clocks = sampler.stop()
    if rank != 0:
        if ctx.dist:
            ctx.dist.destroy_process_group()
        return

    achieved = scan.algorithmic_bytes / (kernel_ms / 1e3) / 1e9
    traffic, traffic_source = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            tj = json.load(open(tp))
            per_row = tj.get("b2a_decode_kernel_dram_bytes_per_row")
            if per_row:
                traffic = int(per_row * my_rows)
                traffic_source = ("profiles/traffic.json: dram__bytes_read+write of b2a_decode_kernel from the builder's "
                                  "ncu --set full capture, per row, x the rows of this launch (not measured in this run)")
        except Exception:
            traffic = None
    cpu = {"value": 0.0, "unit": "rows/s", "cores": 1, "kind": "reference", "sample": "skipped"}
    if not args.profile and world == 1:
        from oracle import ref_py
        if ref_py.available():
            schema, idx = batch_index(path)
            v, sample = cpu_reference_rate(path, schema, idx[:32], args.cpu_seconds, 1)
            cpu = {"value": v, "unit": "rows/s", "cores": 1, "kind": "reference",
                   "sample": "reference read_arrow of ONE file on one thread (what the reference does, "
                             "arrow_file_scan.cpp:35-41): BufferedFileReader + arrow_scan from oracle/_ref, FULL "
                             "validation port; " + sample, "host_cores_available": os.cpu_count()}
        else:
            v, sample = cpu_port_rate(path, 32, args.cpu_seconds, 1)
            cpu = {"value": v, "unit": "rows/s", "cores": 1, "kind": "port", "sample": sample,
                   "host_cores_available": os.cpu_count()}
    out = {
        "metric": "read_arrow rows/s (lineitem, all columns)", "value": value, "unit": "rows/s",
        "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": data_kind, "config": config,
        "gb_per_s": value * scan.algorithmic_bytes / max(1, my_rows) / 1e9,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": ctx.peak, "unit": "GB/s", "frac": achieved / ctx.peak,
                     "traffic": traffic, "traffic_source": traffic_source, "peak_kind": peak_kind,
                     "kernel": "b2a_decode_kernel", "kernel_ms": kernel_ms,
                     "algorithmic_bytes_per_launch": int(scan.algorithmic_bytes),
                     "rows_per_launch": my_rows, "frac_of_nominal_8TBs": achieved / 8000.0},
        "cpu_baseline": cpu, "e2e": e2e, "legs": legs, "gpu_launches": int(launches), "clocks": clocks,
    }
    print(json.dumps(out))
    if ctx.dist:
        ctx.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--sf", type=float, default=10.0)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-legs", action="store_true")
    ap.add_argument("--skip-legs", default="", help="comma-separated: filter_q6,copy,multifile")
    ap.add_argument("--files-per-gpu", type=int, default=8)
    ap.add_argument("--slices", type=int, default=0, help="pipeline slices of the host path (0 = library default)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--columns", default="", help="comma-separated projection (experiments only; the metric is all 16)")
    ap.add_argument("--profile", action="store_true", help="short run for ncu: decode only")
    args = ap.parse_args()
    steps = max(1, args.steps)
    warmup = max(3, args.warmup) if args.impl != "reference" else max(0, args.warmup)
    rank = int(os.environ.get("RANK", "0"))
    if args.impl == "reference" and rank != 0:
        return
    path, data_kind = stream_path(args.sf)
    config = {"workload": f"lineitem SF{args.sf:g} read_arrow full scan, 16 columns, {BATCH_ROWS}-row batches, ONE stream "
                          "whose batches are sharded contiguously over the GPUs",
              "batch_rows": BATCH_ROWS, "file_bytes": os.path.getsize(path),
              "l2": "inputs larger than L2 (no flush needed)",
              "sharding": "contiguous record-batch ranges of one stream (parallel.batch_range)",
              "generator": ("DuckDB dbgen (reference's tpch extension) -> reference ArrowAppender -> IPC framing with "
                            "validity bitmaps (tools/dbgen_lineitem.py)") if data_kind == "dbgen"
              else "numpy-drawn lineitem written by pyarrow (oracle/_ref absent)"}
    if args.impl == "reference":
        reference_arm(args, path, data_kind, config, steps, warmup)
    else:
        b200_arm(args, path, data_kind, config, steps, warmup)


def _stdout_for_the_json_line_only():
    """The contract is ONE JSON line on stdout.  Native libraries write there too (NCCL prints its version line to fd 1
    when NCCL_DEBUG is set in the environment): send fd 1 to stderr for the whole run and keep the real stdout for
    print()."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)


if __name__ == "__main__":
    _stdout_for_the_json_line_only()
    main()
    sys.stdout.flush()
