"""Host-side mirror of the reference extension's read surface, on top of the C ABI (include/b2arrow.h).

    read_arrow(path | [paths], columns=...)      src/scanner/read_arrow.cpp:44-86
    scan_arrow_ipc([(ptr,size) | bytes-like])    src/scanner/scan_arrow_ipc.cpp:20-64
    ArrowIPCReader / ArrowScan                   IPCStreamReader / ArrowFileScan (src/ipc, src/file_scanner)

PyTorch is used only for device memory and streams.  Nothing here computes on the CPU: decoding always
goes through b2a_scan_decode and fails loudly without a CUDA device."""
from __future__ import annotations

import ctypes as C
import glob as _glob
from dataclasses import dataclass

import numpy as np

from . import _lib
from ._lib import B2A_NONE, VECTOR_SIZE, B2AError, check, lib


@dataclass
class Field:
    index: int
    name: str
    format: str
    type: str
    op: str
    out_width: int
    src_width: int
    children: list
    parent: int
    nullable: bool
    decimal_width: int
    decimal_scale: int
    param: int
    dict_id: int
    index_type: str


class ArrowIPCReader:
    """Framing + slicing of one Arrow IPC stream / file (host only)."""

    def __init__(self, handle, keepalive=None, owned=True):
        self._h = handle
        self._keep = keepalive
        self._owned = owned  # False: the handle belongs to a b2a_multifile
        n_top, n_flat = C.c_uint32(), C.c_uint32()
        check(lib.b2a_reader_num_fields(self._h, C.byref(n_top), C.byref(n_flat)))
        self.n_flat = n_flat.value
        self.fields: list[Field] = []
        for i in range(self.n_flat):
            fi = _lib.FieldInfo()
            check(lib.b2a_reader_field(self._h, i, C.byref(fi)))
            self.fields.append(Field(i, fi.name.decode("utf-8", "replace"), fi.format.decode("utf-8", "replace"), _lib.TYPE_NAMES[fi.type_id],
                                     _lib.OP_NAMES[fi.op], fi.out_width, fi.src_width, [], fi.parent,
                                     bool(fi.nullable), fi.decimal_width, fi.decimal_scale, fi.param, fi.dict_id,
                                     _lib.TYPE_NAMES[fi.index_type_id]))
        for f in self.fields:
            if f.parent >= 0:
                self.fields[f.parent].children.append(f.index)
        self.columns = [f for f in self.fields if f.parent < 0]
        self.projection = None

    @classmethod
    def open_file(cls, path: str) -> "ArrowIPCReader":
        h = C.c_void_p()
        check(lib.b2a_reader_open_file(path.encode(), C.byref(h)))
        return cls(h)

    @classmethod
    def open_buffers(cls, buffers) -> "ArrowIPCReader":
        """buffers: list of objects exposing the buffer protocol, or (ptr, size) tuples — the
        LIST(STRUCT(ptr POINTER, size UBIGINT)) argument of scan_arrow_ipc."""
        keep, arr = [], (_lib.IpcBuffer * len(buffers))()
        for i, b in enumerate(buffers):
            if isinstance(b, tuple):
                arr[i].ptr, arr[i].size = b
            else:
                a = np.frombuffer(b, dtype=np.uint8)
                keep.append(a)
                arr[i].ptr, arr[i].size = a.ctypes.data, a.size
        h = C.c_void_p()
        check(lib.b2a_reader_open_buffers(arr, len(buffers), C.byref(h)))
        return cls(h, keep)

    def close(self):
        if self._h:
            if self._owned:
                lib.b2a_reader_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_projection(self, names: list[str]):
        arr = (C.c_char_p * len(names))(*[n.encode() for n in names])
        check(lib.b2a_reader_set_projection(self._h, arr, len(names)))
        self.projection = list(names)

    @property
    def num_batches(self) -> int:
        nb, nr = C.c_uint64(), C.c_uint64()
        check(lib.b2a_reader_num_batches(self._h, C.byref(nb), C.byref(nr)))
        return nb.value

    @property
    def num_rows(self) -> int:
        nb, nr = C.c_uint64(), C.c_uint64()
        check(lib.b2a_reader_num_batches(self._h, C.byref(nb), C.byref(nr)))
        return nr.value

    def batch_info(self, i: int) -> _lib.BatchInfo:
        bi = _lib.BatchInfo()
        check(lib.b2a_reader_batch_info(self._h, i, C.byref(bi)))
        return bi

    @property
    def used_footer(self) -> bool:
        """the message index came from the Arrow file footer rather than from walking the message prefixes"""
        u = C.c_int32()
        check(lib.b2a_reader_used_footer(self._h, C.byref(u)))
        return bool(u.value)

    def image(self) -> np.ndarray:
        p, n = C.c_void_p(), C.c_uint64()
        check(lib.b2a_reader_image(self._h, C.byref(p), C.byref(n)))
        if n.value == 0:
            return np.zeros(0, np.uint8)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value,))

    def schema_metadata(self) -> dict:
        out, n, k, v = {}, C.c_uint32(), C.c_char_p(), C.c_char_p()
        check(lib.b2a_reader_schema_metadata(self._h, 0, C.byref(k), C.byref(v), C.byref(n)))
        for i in range(n.value):
            check(lib.b2a_reader_schema_metadata(self._h, i, C.byref(k), C.byref(v), C.byref(n)))
            out[k.value.decode()] = v.value.decode()
        return out

    def output_columns(self) -> list[Field]:
        if self.projection is None:
            return self.columns
        names = _dedup([c.name for c in self.columns])
        by = {n: c for n, c in zip(names, self.columns)}
        return [by[n] for n in self.projection]


def _dedup(names):
    """QueryResult::DeduplicateColumns (duckdb/src/main/query_result.cpp:71-93)."""
    seen, out = {}, []
    for name in names:
        low = name.lower()
        if low not in seen:
            seen[low] = 1
            out.append(name)
        else:
            new = f"{name}_{seen[low]}"
            while new.lower() in seen:
                seen[low] += 1
                new = f"{name}_{seen[low]}"
            seen[new.lower()] = 1
            out.append(new)
    return out


class ArrowScan:
    """One scan of a batch range of a reader on one GPU (b2a_scan)."""

    def __init__(self, reader: ArrowIPCReader, device: int = 0, batch_lo: int = 0, batch_hi: int | None = None,
                 host_alias: bool = False, device_alias: bool = False):
        """device_alias: B2A_SCAN_DEVICE_ALIAS — DirectConversion columns stay in the DEVICE image for device consumers
        (filter / sum_product): no copy into the arena.
        host_alias: B2A_SCAN_HOST_ALIAS — DirectConversion columns stay in the host image (zero copy, like the
        reference); such a scan is consumed through read_host() only."""
        self.reader = reader
        self.device = device
        self.host_alias = host_alias
        hi = reader.num_batches if batch_hi is None else batch_hi
        self.batch_lo, self.batch_hi = batch_lo, hi
        h = C.c_void_p()
        self.device_alias = device_alias
        flags = (_lib.SCAN_HOST_ALIAS if host_alias else 0) | (_lib.SCAN_DEVICE_ALIAS if device_alias else 0)
        check(lib.b2a_scan_create_ex(reader._h, device, batch_lo, hi, flags,
                                     C.byref(h)))
        self._h = h
        ub = C.c_uint64()
        check(lib.b2a_scan_upload_bytes(self._h, C.byref(ub)))
        self.upload_bytes = ub.value
        v = [C.c_uint64() for _ in range(6)]
        check(lib.b2a_scan_sizes(self._h, *[C.byref(x) for x in v]))
        (self.image_lo, self.image_bytes, self.out_bytes, self.algorithmic_bytes, self.n_tasks,
         self.n_tiles) = [x.value for x in v]
        nc = C.c_uint64()
        check(lib.b2a_scan_num_chunks(self._h, C.byref(nc)))
        self.n_chunks = nc.value
        so, sb, nj = C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(lib.b2a_scan_scratch(self._h, C.byref(so), C.byref(sb), C.byref(nj)))
        # compressed bodies: scratch region behind the image slice that receives the decompressed buffers
        self.scratch_offset, self.scratch_bytes, self.n_decompress_jobs = so.value, sb.value, nj.value
        self.d_out = None
        self.d_image = None

    def close(self):
        if getattr(self, "_h", None):
            lib.b2a_scan_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- device plumbing (torch: memory + streams only)
    def _torch(self):
        import torch
        if not torch.cuda.is_available():
            raise B2AError(5, "no CUDA device: the Arrow IPC scan path has no CPU fallback")
        return torch

    def _stream_ptr(self):
        torch = self._torch()
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def upload(self):
        """H2D of the stream image slice (internally allocated device buffer)."""
        check(lib.b2a_scan_upload(self._h, self._stream_ptr()))

    def set_device_image(self, d_image):
        """d_image: torch uint8 CUDA tensor holding image[image_lo : image_lo+image_bytes] (256-byte aligned)."""
        need = self.scratch_offset + self.scratch_bytes + 64 if self.scratch_bytes else self.image_bytes
        if d_image.numel() < need:
            raise B2AError(1, f"device image too small: {d_image.numel()} bytes, the scan needs {need} "
                              "(stream slice + decompression scratch)")
        self.d_image = d_image
        check(lib.b2a_scan_set_device_image(self._h, C.c_void_p(d_image.data_ptr())))

    @property
    def device_image_bytes(self) -> int:
        """bytes a caller-provided device image must have (stream slice, + scratch of compressed scans)"""
        return self.scratch_offset + self.scratch_bytes + 64 if self.scratch_bytes else self.image_bytes + 64

    def download_scratch(self):
        """host copy (pinned) of the decompressed buffers: what string_t pointers of a compressed scan refer to"""
        torch = self._torch()
        h = torch.empty(max(1, self.scratch_bytes), dtype=torch.uint8, pin_memory=True)
        check(lib.b2a_scan_download_scratch(self._h, C.c_void_p(h.data_ptr()), self._stream_ptr()))
        torch.cuda.current_stream(self.device).synchronize()
        return h

    def alloc_out(self):
        torch = self._torch()
        if self.d_out is None:
            self.d_out = torch.empty(max(self.out_bytes, 256), dtype=torch.uint8, device=f"cuda:{self.device}")
        return self.d_out

    def decode(self, d_out=None, string_ptr_base: int = 0):
        d_out = self.alloc_out() if d_out is None else d_out
        check(lib.b2a_scan_decode(self._h, C.c_void_p(d_out.data_ptr()), string_ptr_base, self._stream_ptr()))
        return d_out

    def finish(self):
        check(lib.b2a_scan_finish(self._h, self._stream_ptr()))

    def download(self, d_out=None, h_out=None) -> np.ndarray:
        torch = self._torch()
        d_out = self.d_out if d_out is None else d_out
        if h_out is None:
            h_out = torch.empty(self.out_bytes, dtype=torch.uint8, pin_memory=True)
        check(lib.b2a_scan_download(self._h, C.c_void_p(d_out.data_ptr()), C.c_void_p(h_out.data_ptr()),
                                    self._stream_ptr()))
        torch.cuda.current_stream(self.device).synchronize()
        return h_out.numpy()

    def read_host(self, h_out=None, n_slices: int = 0) -> np.ndarray:
        """b2a_read_arrow_host: host image -> (pipelined H2D, decode, D2H) -> host vectors in pinned memory."""
        torch = self._torch()
        if h_out is None:
            h_out = torch.empty(max(self.out_bytes, 256), dtype=torch.uint8, pin_memory=True)
        check(lib.b2a_read_arrow_host(self._h, C.c_void_p(h_out.data_ptr()), n_slices))
        self._h_out = h_out
        return h_out.numpy()

    def read_device(self, d_out=None, n_slices: int = 0):
        """b2a_read_arrow_device: host image -> pipelined H2D of the projected buffers + decode -> vectors in d_out
        (device); the current stream is ordered behind the last decode (no host sync, nothing copied back)."""
        d_out = self.alloc_out() if d_out is None else d_out
        check(lib.b2a_read_arrow_device(self._h, C.c_void_p(d_out.data_ptr()), n_slices, self._stream_ptr()))
        return d_out

    @property
    def gpu_launches(self) -> int:
        return int(lib.b2a_scan_gpu_launches(self._h))

    # -- pushed-down filters and the Q6-style consumer (new behaviour; the reference filters above the scan)
    def filter(self, predicates, d_out=None, out=None):
        """predicates: [(column name | projected index, op, int constant)], ANDed.  Returns (sel, count, total):
        torch tensors uint32[n_chunks*2048], uint32[n_chunks], uint64[1] on the device — one selection vector per
        2 048-row DuckDB vector."""
        torch = self._torch()
        d_out = self.d_out if d_out is None else d_out
        names = [f.name for f in self.reader.output_columns()]
        arr = (_lib.Predicate * len(predicates))()
        for i, (col, op, const) in enumerate(predicates):
            arr[i].column = names.index(col) if isinstance(col, str) else col
            arr[i].cmp = _lib.CMP[op]
            arr[i].constant = int(const)
        dev = f"cuda:{self.device}"
        if out is not None:  # (sel, count, total) of an earlier call: no allocation (CUDA-graph capture)
            sel, count, total = out
        else:
            sel = torch.empty(max(1, self.n_chunks) * VECTOR_SIZE, dtype=torch.int32, device=dev)
            count = torch.zeros(max(1, self.n_chunks), dtype=torch.int32, device=dev)
            total = torch.zeros(1, dtype=torch.int64, device=dev)
        check(lib.b2a_scan_filter(self._h, C.c_void_p(d_out.data_ptr()), arr, len(predicates),
                                  C.c_void_p(sel.data_ptr()), C.c_void_p(count.data_ptr()),
                                  C.c_void_p(total.data_ptr()), self._stream_ptr()))
        return sel, count, total

    def filter_sum_product(self, predicates, col_a, col_b, d_out=None, out=None):
        """b2a_scan_filter_sum_product: count and sum(a*b) over the rows passing the ANDed predicates in ONE launch (no
        selection vectors).  Returns (total int64[1], sum int64[2] = {lo, hi}) device tensors, written by work queued on the
        current stream; `out` = the pair of an earlier call (no allocation: CUDA-graph capture)."""
        torch = self._torch()
        d_out = self.d_out if d_out is None else d_out
        names = [f.name for f in self.reader.output_columns()]
        arr = (_lib.Predicate * len(predicates))()
        for i, (col, op, const) in enumerate(predicates):
            arr[i].column = names.index(col) if isinstance(col, str) else col
            arr[i].cmp = _lib.CMP[op]
            arr[i].constant = int(const)
        ia = names.index(col_a) if isinstance(col_a, str) else col_a
        ib = names.index(col_b) if isinstance(col_b, str) else col_b
        dev = f"cuda:{self.device}"
        if out is not None:
            total, res = out
        else:
            total = torch.zeros(1, dtype=torch.int64, device=dev)
            res = torch.zeros(2, dtype=torch.int64, device=dev)
        check(lib.b2a_scan_filter_sum_product(self._h, C.c_void_p(d_out.data_ptr()), arr, len(predicates), ia, ib,
                                              C.c_void_p(total.data_ptr()), C.c_void_p(res.data_ptr()),
                                              self._stream_ptr()))
        return total, res

    def sum_product(self, col_a, col_b, sel=None, count=None, d_out=None) -> int:
        """sum(a*b) over the selected rows (all rows when sel is None) as an exact Python int (HUGEINT)."""
        torch = self._torch()
        d_out = self.d_out if d_out is None else d_out
        names = [f.name for f in self.reader.output_columns()]
        ia = names.index(col_a) if isinstance(col_a, str) else col_a
        ib = names.index(col_b) if isinstance(col_b, str) else col_b
        res = torch.zeros(2, dtype=torch.int64, device=f"cuda:{self.device}")
        check(lib.b2a_scan_sum_product(self._h, C.c_void_p(d_out.data_ptr()), ia, ib,
                                       C.c_void_p(sel.data_ptr() if sel is not None else 0),
                                       C.c_void_p(count.data_ptr() if count is not None else 0),
                                       C.c_void_p(res.data_ptr()), self._stream_ptr()))
        lo, hi = [int(x) for x in res.cpu().tolist()]
        return (hi << 64) | (lo & 0xFFFFFFFFFFFFFFFF)

    def sum_product_async(self, col_a, col_b, sel=None, count=None, d_out=None, result=None):
        """sum_product without the read-back: returns the int64[2] device tensor {lo, hi} of the 128-bit sum, written by
        work queued on the current stream (graph-capturable once the column table is cached by a first call)."""
        torch = self._torch()
        d_out = self.d_out if d_out is None else d_out
        names = [f.name for f in self.reader.output_columns()]
        ia = names.index(col_a) if isinstance(col_a, str) else col_a
        ib = names.index(col_b) if isinstance(col_b, str) else col_b
        res = torch.zeros(2, dtype=torch.int64, device=f"cuda:{self.device}") if result is None else result
        check(lib.b2a_scan_sum_product(self._h, C.c_void_p(d_out.data_ptr()), ia, ib,
                                       C.c_void_p(sel.data_ptr() if sel is not None else 0),
                                       C.c_void_p(count.data_ptr() if count is not None else 0),
                                       C.c_void_p(res.data_ptr()), self._stream_ptr()))
        return res

    # -- layout
    def column(self, batch: int, flat: int) -> _lib.ColumnOut:
        co = _lib.ColumnOut()
        check(lib.b2a_scan_column(self._h, batch, flat, C.byref(co)))
        return co

    def dictionary(self, flat: int) -> _lib.ColumnOut:
        co, df = _lib.ColumnOut(), C.c_uint32()
        check(lib.b2a_scan_dictionary(self._h, flat, C.byref(co), C.byref(df)))
        return co


def read_arrow(path, columns: list[str] | None = None, device: int = 0, host_alias: bool = True,
               union_by_name: bool = False, filename=False, hive_partitioning: bool = False,
               hive_types_autocast: bool = True):
    """`FROM read_arrow('x.arrows')` / `read_arrow([...])` / `read_arrow('dir/*.arrow', union_by_name=true, ...)`.
    One plain path without multi-file options -> DecodedScan (the decoded vectors on the host in canonical layout);
    a list, a glob or any multi-file option -> MultiFileScan (src/file_scanner: global schema, per-file column mapping,
    casts on the device, constant columns)."""
    multi = (isinstance(path, (list, tuple)) or any(c in path for c in "*?[") or union_by_name or filename
             or hive_partitioning)
    if multi:
        return MultiFileScan(path, columns=columns, device=device, union_by_name=union_by_name, filename=filename,
                             hive_partitioning=hive_partitioning, hive_types_autocast=hive_types_autocast)
    r = ArrowIPCReader.open_file(path)
    if columns:
        r.set_projection(columns)
    return _scan_to_host(r, device, host_alias)


def scan_arrow_ipc(buffers, columns: list[str] | None = None, device: int = 0, host_alias: bool = True) -> "DecodedScan":
    r = ArrowIPCReader.open_buffers(buffers)
    if columns:
        r.set_projection(columns)
    return _scan_to_host(r, device, host_alias)


def _scan_to_host(reader: ArrowIPCReader, device: int, host_alias: bool = True) -> "DecodedScan":
    """The drop-in path: host stream in, host vectors out (b2a_read_arrow_host).  With host_alias the fixed-width
    columns alias the host image exactly like the reference's DirectConversion; without it every column goes through
    upload -> decode -> download (the device-resident layout, copied back whole)."""
    image = reader.image()
    compressed = any(reader.batch_info(b).compressed for b in range(reader.num_batches))
    if host_alias and not compressed:
        scan = ArrowScan(reader, device, host_alias=True)
        arena = scan.read_host()
        return DecodedScan(scan, arena, image, image.ctypes.data if image.size else 0)
    scan = ArrowScan(reader, device)
    scan.upload()
    if scan.scratch_bytes:
        # compressed bodies: the Arrow buffers the string_t pointers refer to exist only after decompression; the host
        # consumer gets a copy of the scratch region and pointers into it
        import torch
        h_scratch = torch.empty(scan.scratch_bytes, dtype=torch.uint8, pin_memory=True)
        base = h_scratch.data_ptr() - scan.scratch_offset
        scan.decode(string_ptr_base=base)
        scan.finish()
        check(lib.b2a_scan_download_scratch(scan._h, C.c_void_p(h_scratch.data_ptr()), scan._stream_ptr()))
        arena = scan.download()
        ds = DecodedScan(scan, arena, image, base - scan.image_lo)
        ds._h_scratch = h_scratch
        return ds
    # string_t pointers refer to the HOST copy of the stream, exactly where the reference's pointers point
    base = image.ctypes.data + scan.image_lo if image.size else 0
    scan.decode(string_ptr_base=base)
    scan.finish()
    arena = scan.download()
    return DecodedScan(scan, arena, image, base - scan.image_lo if image.size else 0)


class DecodedScan:
    """Host copy of a decoded scan: `arena` (the output layout of b2a_scan_column) + the stream image the
    string_t pointers refer to."""

    def __init__(self, scan: ArrowScan, arena: np.ndarray, image: np.ndarray, ptr_origin: int):
        self.scan, self.arena, self.image, self.ptr_origin = scan, arena, image, ptr_origin
        self.reader = scan.reader

    @property
    def num_rows(self):
        return sum(self.reader.batch_info(b).rows for b in range(self.scan.batch_lo, self.scan.batch_hi))

    def column(self, batch: int, flat: int) -> "ColumnView":
        return ColumnView(self, self.scan.column(batch, flat), self.reader.fields[flat], batch)


class ColumnView:
    """numpy views of one decoded field of one batch, in the canonical comparison form of SURVEY §8c."""

    def __init__(self, ds: DecodedScan, co: _lib.ColumnOut, field: Field, batch: int, is_dict_values=False):
        self.ds, self.co, self.field, self.batch = ds, co, field, batch
        self.rows = int(co.rows)
        self.is_dict_values = is_dict_values

    @property
    def is_dictionary(self):
        return self.field.dict_id >= 0 and not self.is_dict_values

    @property
    def width(self):
        return 4 if self.is_dictionary else self.field.out_width

    @property
    def valid(self) -> np.ndarray:
        """bool[rows] — first `rows` bits of the ValidityMask; all True when the mask was left untouched."""
        if self.field.type == "SQLNULL":
            return np.zeros(self.rows, bool)
        if self.co.validity_off == B2A_NONE:
            return np.ones(self.rows, bool)
        nbytes = (self.rows + 7) // 8
        raw = self.ds.arena[self.co.validity_off: self.co.validity_off + nbytes]
        return np.unpackbits(raw, bitorder="little")[: self.rows].astype(bool)

    def validity_words(self, extra_rows=0) -> np.ndarray:
        n = (self.rows + extra_rows + 63) // 64
        return self.ds.arena[self.co.validity_off: self.co.validity_off + 8 * n].view(np.uint64)

    @property
    def data(self) -> np.ndarray:
        """uint8[rows, width]"""
        w = self.width
        if w and self.co.data_off == B2A_NONE and self.co.image_off != B2A_NONE:  # host-alias: the Arrow buffer itself
            return self.ds.image[self.co.image_off: self.co.image_off + self.rows * w].reshape(self.rows, w)
        if w == 0 or self.co.data_off == B2A_NONE:
            return np.zeros((self.rows, 0), np.uint8)
        return self.ds.arena[self.co.data_off: self.co.data_off + self.rows * w].reshape(self.rows, w)

    @property
    def valid_count(self) -> int:
        if self.co.stats_off == B2A_NONE:
            return self.rows
        return int(self.ds.arena[self.co.stats_off: self.co.stats_off + 8].view(np.uint64)[0])

    def chunk_child_starts(self) -> np.ndarray:
        n = (self.rows + VECTOR_SIZE - 1) // VECTOR_SIZE + 1
        return self.ds.arena[self.co.aux_off: self.co.aux_off + 8 * n].view(np.uint64)

    def chunk_child_windows(self) -> np.ndarray:
        """list views: (child start row, child rows) of every 2 048-row chunk (b2a_column_out.aux_off)"""
        n = (self.rows + VECTOR_SIZE - 1) // VECTOR_SIZE
        return self.ds.arena[self.co.aux_off: self.co.aux_off + 16 * n].view(np.uint64).reshape(n, 2)

    def children(self) -> list["ColumnView"]:
        if self.is_dictionary:
            return []
        return [self.ds.column(self.batch, c) for c in self.field.children]

    def dictionary(self) -> "ColumnView":
        co = self.ds.scan.dictionary(self.field.index)
        return ColumnView(self.ds, co, self.field, self.batch, is_dict_values=True)

    def string_parts(self):
        """(len u32[rows], inline/prefix bytes uint8[rows,12], pointer offset from the column's char buffer i64[rows])"""
        d = self.data
        lens = d[:, 0:4].copy().view(np.uint32).ravel()
        ptr = d[:, 8:16].copy().view(np.uint64).ravel().astype(np.int64)
        rel = ptr - np.int64(self.ds.ptr_origin + int(self.co.chars_off))
        return lens, d[:, 4:16], rel


# ==================================================================================================== multi-file
@dataclass
class GlobalColumn:
    index: int
    name: str
    type: str      # LogicalType::ToString() of the global column
    type_id: str   # TYPE_NAMES entry
    kind: int      # MF_COL_DATA | MF_COL_FILENAME | MF_COL_HIVE
    decimal: tuple


@dataclass
class ConstantColumn:
    """A constant vector of a multi-file scan: filename / hive partition value (text + global type), or NULL."""
    rows: int
    value: str | None
    column: GlobalColumn


class CastColumnView(ColumnView):
    """A column that went through a multi-file cast: data (and VARCHAR heap) live in the file's FINAL arena, the
    validity mask is the source vector's (scan arena)."""

    def __init__(self, ds, co, field, batch, final: np.ndarray, final_origin: int):
        super().__init__(ds, co, field, batch)
        self.final, self.final_origin = final, final_origin

    @property
    def data(self) -> np.ndarray:
        w = self.field.out_width
        return self.final[self.co.data_off: self.co.data_off + self.rows * w].reshape(self.rows, w)

    def string_parts(self):
        d = self.data
        lens = d[:, 0:4].copy().view(np.uint32).ravel()
        ptr = d[:, 8:16].copy().view(np.uint64).ravel().astype(np.int64)
        return lens, d[:, 4:16], ptr - np.int64(self.final_origin + int(self.co.chars_off))


class MultiFileFile:
    """One file of a multi-file scan, decoded (and cast) on the device and copied back to the host."""

    def __init__(self, mfs: "MultiFileScan", index: int, path: str, reader: ArrowIPCReader):
        self.mfs, self.index, self.path, self.reader = mfs, index, path, reader
        self.ds = None
        self._fin = None
        self.final = None
        self.gpu_launches = 0

    def map(self, column: int) -> _lib.MfMapInfo:
        mi = _lib.MfMapInfo()
        check(lib.b2a_multifile_map(self.mfs._h, self.index, column, C.byref(mi)))
        return mi

    def scan(self, batch_lo: int = 0, batch_hi: int | None = None):
        """upload -> decode -> casts -> download (device-resident layout copied back; string_t pointers are host
        addresses of the file's stream image / of the cast heap)."""
        import torch
        mfs = self.mfs
        image = self.reader.image()
        scan = ArrowScan(self.reader, mfs.device, batch_lo, batch_hi)
        scan.upload()
        base = image.ctypes.data + scan.image_lo if image.size else 0
        if scan.scratch_bytes:  # compressed bodies: string_t pointers refer to the host copy of the scratch region
            self._h_scratch = torch.empty(scan.scratch_bytes, dtype=torch.uint8, pin_memory=True)
            base = self._h_scratch.data_ptr() - scan.scratch_offset
        scan.decode(string_ptr_base=base)
        fin = C.c_void_p()
        check(lib.b2a_multifile_finalize_create(mfs._h, self.index, scan._h, C.byref(fin)))
        self._fin = fin
        fb, nc = C.c_uint64(), C.c_uint64()
        check(lib.b2a_multifile_finalize_sizes(fin, C.byref(fb), C.byref(nc)))
        self.n_casts = nc.value
        h_final = torch.empty(max(fb.value, 1), dtype=torch.uint8, pin_memory=True)
        if nc.value:
            d_final = torch.empty(fb.value, dtype=torch.uint8, device=f"cuda:{mfs.device}")
            check(lib.b2a_multifile_finalize_run(fin, C.c_void_p(scan.d_out.data_ptr()), C.c_void_p(d_final.data_ptr()),
                                                 base, h_final.data_ptr(), scan._stream_ptr()))
            check(lib.b2a_multifile_finalize_finish(fin, scan._stream_ptr()))
            h_final.copy_(d_final)
        scan.finish()
        if scan.scratch_bytes:
            check(lib.b2a_scan_download_scratch(scan._h, C.c_void_p(self._h_scratch.data_ptr()), scan._stream_ptr()))
        arena = scan.download()
        self._h_final = h_final
        self.final = h_final.numpy()
        self.ds = DecodedScan(scan, arena, image, base - scan.image_lo if image.size else 0)
        self.gpu_launches = scan.gpu_launches + int(lib.b2a_multifile_finalize_gpu_launches(fin))
        return self

    def close(self):
        if self._fin:
            lib.b2a_multifile_finalize_destroy(self._fin)
            self._fin = None
        if self.ds is not None:
            self.ds.scan.close()

    def column(self, batch: int, column: int):
        """Vector of GLOBAL column `column` for record batch `batch` of this file: ColumnView (the file's own vector),
        CastColumnView, or ConstantColumn."""
        mi = self.map(column)
        gc = self.mfs.all_columns[column]
        if mi.kind in (_lib.MF_MAP_NULL, _lib.MF_MAP_CONSTANT):
            rows = self.reader.batch_info(batch).rows
            value = None if mi.constant_is_null else mi.constant.decode()
            return ConstantColumn(int(rows), value, gc)
        if mi.kind == _lib.MF_MAP_LOCAL:
            return self.ds.column(batch, mi.local_flat)
        co = _lib.ColumnOut()
        check(lib.b2a_multifile_finalize_column(self._fin, batch, column, C.byref(co)))
        src = self.reader.fields[mi.local_flat]
        if gc.type_id == "DECIMAL":  # DuckDB's physical type for the width
            width = 2 if gc.decimal[0] <= 4 else 4 if gc.decimal[0] <= 9 else 8 if gc.decimal[0] <= 18 else 16
        else:
            width = {"VARCHAR": 16, "DOUBLE": 8, "FLOAT": 4, "BIGINT": 8, "UBIGINT": 8, "INTEGER": 4, "UINTEGER": 4,
                     "SMALLINT": 2, "USMALLINT": 2, "TINYINT": 1, "UTINYINT": 1, "TIMESTAMP": 8, "HUGEINT": 16, "DATE": 4}[gc.type_id]
        f = Field(src.index, gc.name, "", gc.type_id, "CAST", width, src.out_width, [], -1, True, 0, 0, 0, -1, "INVALID")
        return CastColumnView(self.ds, co, f, batch, self.final, self._h_final.data_ptr())


class MultiFileScan:
    """read_arrow over several files (src/file_scanner + DuckDB's MultiFileReader in the reference)."""

    def __init__(self, paths, columns=None, device: int = 0, union_by_name=False, filename=False,
                 hive_partitioning=False, hive_types_autocast=True):
        self.device = device
        plist = list(paths) if isinstance(paths, (list, tuple)) else [paths]
        arr = (C.c_char_p * len(plist))(*[p.encode() for p in plist])
        opt = _lib.MultiFileOptions()
        opt.union_by_name = int(bool(union_by_name))
        opt.filename = int(bool(filename))
        self._fn_col = filename.encode() if isinstance(filename, str) else None
        opt.filename_column = self._fn_col
        opt.hive_partitioning = int(bool(hive_partitioning))
        opt.hive_types_autocast = int(bool(hive_types_autocast))
        h = C.c_void_p()
        check(lib.b2a_multifile_open(arr, len(plist), C.byref(opt), C.byref(h)))
        self._h = h
        if columns:
            names = (C.c_char_p * len(columns))(*[n.encode() for n in columns])
            check(lib.b2a_multifile_set_projection(self._h, names, len(columns)))
        nf, nc, npj = C.c_uint32(), C.c_uint32(), C.c_uint32()
        check(lib.b2a_multifile_counts(self._h, C.byref(nf), C.byref(nc), C.byref(npj)))
        self.all_columns = []
        for i in range(nc.value):
            ci = _lib.MfColumnInfo()
            check(lib.b2a_multifile_column(self._h, i, C.byref(ci)))
            self.all_columns.append(GlobalColumn(i, ci.name.decode(), ci.type_name.decode(), _lib.TYPE_NAMES[ci.type_id],
                                                 ci.kind, (ci.decimal_width, ci.decimal_scale)))
        self.columns = []
        for i in range(npj.value):
            c = C.c_uint32()
            check(lib.b2a_multifile_projected(self._h, i, C.byref(c)))
            self.columns.append(self.all_columns[c.value])
        self.files = []
        for i in range(nf.value):
            path, rh = C.c_char_p(), C.c_void_p()
            check(lib.b2a_multifile_file(self._h, i, C.byref(path), C.byref(rh)))
            self.files.append(MultiFileFile(self, i, path.value.decode(), ArrowIPCReader(rh, owned=False)))

    @property
    def names(self):
        return [c.name for c in self.columns]

    @property
    def types(self):
        return [c.type for c in self.columns]

    @property
    def num_rows(self):
        return sum(f.reader.num_rows for f in self.files)

    def scan_all(self):
        for f in self.files:
            f.scan()
        return self

    def close(self):
        for f in self.files:
            f.close()
        if getattr(self, "_h", None):
            lib.b2a_multifile_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ======================================================================================================== writer
class DeviceVector:
    """A flat DuckDB vector (or the end-to-end concatenation of the vectors of a column) in device memory —
    the Python face of b2a_vector.  `data` / `validity` are torch CUDA tensors (uint8 bytes / int64 mask words);
    VARCHAR data is string_t[rows] whose pointers are device addresses, LIST / MAP data is list_entry_t[rows]."""

    def __init__(self, type: str, rows: int, data=None, validity=None, children=(), name: str = "",
                 decimal=(0, 0), param: int = 0, timezone: str | None = None, keep=()):
        self.type, self.rows, self.data, self.validity = type, int(rows), data, validity
        self.children, self.name, self.decimal, self.param, self.timezone = list(children), name, decimal, param, timezone
        self.keep = list(keep)  # tensors the string_t pointers refer to (string heaps)

    def _fill(self, cv: _lib.Vector, keep: list):
        cv.type_id = _lib.T[self.type]
        cv.decimal_width, cv.decimal_scale = self.decimal
        cv.param = self.param
        name = self.name.encode()
        keep.append(name)
        cv.name = name
        if self.timezone is not None:
            tz = self.timezone.encode()
            keep.append(tz)
            cv.timezone = tz
        cv.data = self.data.data_ptr() if self.data is not None and self.data.numel() else None
        cv.validity = self.validity.data_ptr() if self.validity is not None else None
        cv.rows = self.rows
        cv.n_children = len(self.children)
        if self.children:
            arr = (_lib.Vector * len(self.children))()
            keep.append(arr)
            for i, c in enumerate(self.children):
                c._fill(arr[i], keep)
            cv.children = arr


class ArrowWriter:
    """b2a_writer: COPY ... (FORMAT ARROWS) of a device-resident table (src/writer/write_arrow_stream.cpp)."""

    def __init__(self, columns: list[DeviceVector], rows: int, row_group_size: int = 122_880,
                 kv_metadata: dict | None = None, device: int = 0, arrow_large_buffer_size: bool = False,
                 arrow_output_list_view: bool = False, produce_arrow_string_view: bool = False,
                 arrow_lossless_conversion: bool = False):
        self.device, self.columns = device, columns
        self._keep: list = []
        arr = (_lib.Vector * len(columns))()
        for i, c in enumerate(columns):
            c._fill(arr[i], self._keep)
        opt = _lib.WriteOptions()
        opt.row_group_size = row_group_size
        opt.flags = ((_lib.WRITE_LARGE_OFFSETS if arrow_large_buffer_size else 0)
                     | (_lib.WRITE_LIST_VIEW if arrow_output_list_view else 0)
                     | (_lib.WRITE_STRING_VIEW if produce_arrow_string_view else 0)
                     | (_lib.WRITE_LOSSLESS if arrow_lossless_conversion else 0))
        kv = kv_metadata or {}
        if kv:
            ks = (C.c_char_p * len(kv))(*[str(k).encode() for k in kv])
            vs = (C.c_char_p * len(kv))(*[str(v).encode() for v in kv.values()])
            self._keep += [ks, vs]
            opt.n_kv_metadata, opt.kv_keys, opt.kv_values = len(kv), ks, vs
        h = C.c_void_p()
        check(lib.b2a_writer_create(arr, len(columns), rows, C.byref(opt), device, C.byref(h)))
        self._h = h
        self.stream_bytes = self.n_batches = self.body_bytes = self.algorithmic_bytes = None

    def close(self):
        if getattr(self, "_h", None):
            lib.b2a_writer_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream_ptr(self):
        import torch
        if not torch.cuda.is_available():
            raise B2AError(5, "no CUDA device: the Arrow IPC writer path has no CPU fallback")
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def measure(self):
        check(lib.b2a_writer_measure(self._h, self._stream_ptr()))
        v = [C.c_uint64() for _ in range(4)]
        check(lib.b2a_writer_sizes(self._h, *[C.byref(x) for x in v]))
        self.stream_bytes, self.n_batches, self.body_bytes, self.algorithmic_bytes = [x.value for x in v]

    def encode(self, d_stream):
        check(lib.b2a_writer_encode(self._h, C.c_void_p(d_stream.data_ptr()), self._stream_ptr()))

    def finish(self):
        check(lib.b2a_writer_finish(self._h, self._stream_ptr()))

    def headers(self, h_stream: np.ndarray):
        check(lib.b2a_writer_headers(self._h, C.c_void_p(h_stream.ctypes.data)))

    def schema_message(self) -> bytes:
        """the encapsulated Schema message (host only; works without a device)"""
        ptr, size = C.c_void_p(), C.c_uint64()
        check(lib.b2a_writer_schema_message(self._h, C.byref(ptr), C.byref(size)))
        return C.string_at(ptr, size.value)

    @property
    def preamble_bytes(self) -> int:
        """schema message + DictionaryBatch messages (one per ENUM field): the stream before its first record batch"""
        n = C.c_uint64()
        check(lib.b2a_writer_preamble(self._h, C.byref(n)))
        return int(n.value)

    def batch(self, i: int) -> _lib.BatchInfo:
        bi = _lib.BatchInfo()
        check(lib.b2a_writer_batch(self._h, i, C.byref(bi)))
        return bi

    @property
    def gpu_launches(self) -> int:
        return int(lib.b2a_writer_gpu_launches(self._h))

    def write(self) -> np.ndarray:
        """measure -> encode -> finish -> D2H -> headers; returns the complete IPC stream (host, pinned)."""
        import torch
        if self.stream_bytes is None:
            self.measure()
        d_stream = torch.empty(self.stream_bytes, dtype=torch.uint8, device=f"cuda:{self.device}")
        self.encode(d_stream)
        self.finish()
        h = torch.empty(self.stream_bytes, dtype=torch.uint8, pin_memory=True)
        h.copy_(d_stream)
        torch.cuda.current_stream(self.device).synchronize()
        out = h.numpy()
        self.headers(out)
        self._h_stream = h
        return out


def copy_to_arrows(columns: list[DeviceVector], path: str | None = None, row_group_size: int = 122_880,
                   chunk_size: int | None = None, kv_metadata: dict | None = None, device: int = 0,
                   arrow_large_buffer_size: bool = False, arrow_output_list_view: bool = False,
                   produce_arrow_string_view: bool = False, arrow_lossless_conversion: bool = False) -> np.ndarray:
    """`COPY tbl TO 'x.arrows' (FORMAT ARROWS, ROW_GROUP_SIZE n, KV_METADATA {...})`
    (src/writer/write_arrow_stream.cpp:54-125: `chunk_size` is an alias of `row_group_size`).
    `arrow_large_buffer_size` / `arrow_output_list_view` / `produce_arrow_string_view` / `arrow_lossless_conversion` are
    the session settings of those names (`SET ... = true`),
    which the reference's writer sees through `context.GetClientProperties()` (src/writer/arrow_stream_writer.cpp:11)."""
    rows = columns[0].rows if columns else 0
    w = ArrowWriter(columns, rows, chunk_size or row_group_size, kv_metadata, device, arrow_large_buffer_size,
                    arrow_output_list_view, produce_arrow_string_view, arrow_lossless_conversion)
    stream = w.write()
    if path is not None:
        with open(path, "wb") as f:
            f.write(memoryview(stream))
    stream_copy = stream.copy()
    w.close()
    return stream_copy


def to_arrow_ipc(columns: list[DeviceVector], device: int = 0,
                 arrow_large_buffer_size: bool = False, arrow_output_list_view: bool = False,
                 produce_arrow_string_view: bool = False, arrow_lossless_conversion: bool = False) -> list[tuple[bytes, bool]]:
    """`FROM to_arrow_ipc((FROM t))`: rows of (ipc BLOB, header BOOL) — first the schema message, then one
    record-batch message per 120 x 2 048 rows (src/writer/to_arrow_ipc.cpp:21,116-155, to_arrow_ipc.hpp:28)."""
    rows = columns[0].rows if columns else 0
    w = ArrowWriter(columns, rows, 120 * VECTOR_SIZE, None, device, arrow_large_buffer_size, arrow_output_list_view,
                    produce_arrow_string_view, arrow_lossless_conversion)
    stream = w.write()
    out = []
    out.append((bytes(stream[:w.preamble_bytes]), True))  # schema (+ the dictionaries of ENUM columns)
    for i in range(w.n_batches):
        bi = w.batch(i)
        out.append((bytes(stream[bi.meta_offset: bi.body_offset + bi.body_length]), False))
    w.close()
    return out
