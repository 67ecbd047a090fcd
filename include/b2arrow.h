/*
 * b2arrow.h — thin C ABI between the DuckDB-facing host code (src/ipc, src/scanner,
 * src/file_scanner, src/writer) and the sm_100a kernels.
 *
 * Every entry point is `extern "C"`, takes plain pointers / sizes / POD structs, returns an int
 * status (b2a_status) and never throws across the boundary.  b2a_last_error() returns the
 * thread-local message of the last failing call.
 *
 * Reference interfaces replaced (paths relative to the pdet/duckdb-arrow checkout):
 *   - b2a_reader_*      : IPCStreamReader / IPCFileStreamReader / IPCBufferStreamReader
 *                         (src/ipc/stream_reader/base_stream_reader.cpp:52-212,
 *                          ipc_file_stream_reader.cpp:47-138, ipc_buffer_stream_reader.cpp:12-76)
 *                         plus the nanoarrow calls they make (ArrowIpcDecoderDecodeHeader /
 *                         DecodeSchema / DecodeArrayFromShared, base_stream_reader.cpp:112-140).
 *   - b2a_scan_*        : ArrowTableFunction::ArrowScanFunction + ArrowToDuckDB
 *                         (duckdb/src/function/table/arrow.cpp:172-201,
 *                          duckdb/src/function/table/arrow_conversion.cpp:1421-1484) — one call
 *                         converts whole record batches instead of one 2 048-row chunk.
 *   - b2a_scan_filter   : new (the reference has filter_pushdown=false, src/scanner/read_arrow.cpp:46-48);
 *                         semantics = DuckDB ConstantFilter over the scanned vectors.
 *   - b2a_writer_*      : ColumnDataCollectionSerializer::Serialize + ArrowAppender + nanoarrow encoder
 *                         (src/writer/column_data_collection_serializer.cpp:80-120,
 *                          duckdb/src/common/arrow/arrow_appender.cpp:37-131) and ArrowStreamWriter
 *                         (src/writer/arrow_stream_writer.cpp:55-82).
 */
#ifndef B2ARROW_H
#define B2ARROW_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2A_VECTOR_SIZE 2048u /* STANDARD_VECTOR_SIZE, duckdb/src/include/duckdb/common/vector_size.hpp:16-20 */
#define B2A_NONE UINT64_MAX   /* "no such buffer" in every *_off field */

typedef enum b2a_status {
	B2A_OK = 0,
	B2A_EINVAL = 1,     /* InvalidInputException */
	B2A_EIO = 2,        /* IOException (framing, truncated message, bad continuation token) */
	B2A_ENOTIMPL = 3,   /* NotImplementedException (decimal256, dense union, ...) */
	B2A_ECONVERSION = 4, /* ConversionException (checked multiply overflow, >4GB string, index > u32) */
	B2A_ECUDA = 5,      /* CUDA runtime error, or no device: there is no CPU fallback */
	B2A_EINTERNAL = 6,  /* InternalException (projection of unknown / duplicate column) */
	B2A_EOOM = 7
} b2a_status;

/* DuckDB logical types the path can produce (names follow LogicalTypeId). */
typedef enum b2a_type_id {
	B2A_T_INVALID = 0,
	B2A_T_SQLNULL,
	B2A_T_BOOLEAN,
	B2A_T_TINYINT,
	B2A_T_SMALLINT,
	B2A_T_INTEGER,
	B2A_T_BIGINT,
	B2A_T_UTINYINT,
	B2A_T_USMALLINT,
	B2A_T_UINTEGER,
	B2A_T_UBIGINT,
	B2A_T_HUGEINT,
	B2A_T_UHUGEINT,
	B2A_T_FLOAT,
	B2A_T_DOUBLE,
	B2A_T_DECIMAL,
	B2A_T_VARCHAR,
	B2A_T_BLOB,
	B2A_T_DATE,
	B2A_T_TIME,
	B2A_T_TIMESTAMP,
	B2A_T_TIMESTAMP_SEC,
	B2A_T_TIMESTAMP_MS,
	B2A_T_TIMESTAMP_NS,
	B2A_T_TIMESTAMP_TZ,
	B2A_T_INTERVAL,
	B2A_T_UUID,
	B2A_T_LIST,
	B2A_T_STRUCT,
	B2A_T_MAP,
	B2A_T_ARRAY,
	B2A_T_UNION, /* sparse union (arrow_duck_schema.cpp:235-259): a STRUCT-like node whose OWN data is the UTINYINT tag
	               per row (DuckDB's UNION = STRUCT{tag, members...}); children = the members */
	/* Arrow extension types registered by default (duckdb/src/common/arrow/arrow_type_extension.cpp:366-390) */
	B2A_T_TIME_TZ, /* arrow.opaque {DuckDB, time_tz} over w:8: dtime_tz_t bits, copied */
	B2A_T_BIT,     /* arrow.opaque {DuckDB, bit} over z / Z: string_t like BLOB */
	B2A_T_VARINT,  /* arrow.opaque {DuckDB, varint} over z / Z */
	B2A_T_JSON,    /* arrow.json over u / U / vu: VARCHAR with the JSON alias */
	B2A_T_ENUM     /* writer only (the scan returns dictionary-encoded columns as VARCHAR, like the reference): data = the
	                  codes, uint8 / uint16 / uint32 by dictionary size (EnumTypeInfo::DictType: <= 255, <= 65 535, else);
	                  children[0] = the VARCHAR vector of the dictionary values in insert order (EnumType::GetValuesInsertOrder).
	                  Written as a dictionary-encoded utf8 field "C" / "S" / "I" + one DictionaryBatch per stream
	                  (arrow_converter.cpp:362-386, appender/enum_data.hpp:24-87) */
} b2a_type_id;

/* Conversion op of one field = which kernel body runs for its value buffer(s). */
typedef enum b2a_op {
	B2A_OP_NONE = 0,        /* struct / fixed-size-list parent / null: no value buffer */
	B2A_OP_COPY,            /* DirectConversion, arrow_conversion.cpp:454-461 (+ endian swap) */
	B2A_OP_BOOL,            /* :869-891 */
	B2A_OP_STRING32,        /* SetVectorString<uint32_t>, :409-423,926-933 */
	B2A_OP_STRING64,        /* SetVectorString<uint64_t>, :919-925 */
	B2A_OP_STRING_VIEW,     /* SetVectorStringView, :425-452 */
	B2A_OP_BLOB_FIXED,      /* ArrowToDuckDBBlob fixed size, :347-360 */
	B2A_OP_LIST32,          /* ConvertArrowListOffsetsTemplated<uint32_t>, :139-165 */
	B2A_OP_LIST64,          /* ... <uint64_t> */
	B2A_OP_MUL_I32,         /* TimeConversion<int32_t> (checked multiply by param), :463-478 */
	B2A_OP_MUL_I64,         /* TimeConversion<int64_t>/TimestampTZConversion, :463-511 */
	B2A_OP_DIV_I64,         /* ns -> us truncating division by param, :986-994,1018-1026 */
	B2A_OP_DATE64,          /* ms -> days, :951-961 */
	B2A_OP_INTERVAL_MUL,    /* IntervalConversionUs (unchecked validity!), :513-525 */
	B2A_OP_INTERVAL_DIV,    /* duration ns, :1052-1062 */
	B2A_OP_INTERVAL_MONTHS, /* :527-537 */
	B2A_OP_INTERVAL_MDN,    /* month-day-nano, :539-549 */
	B2A_OP_DECIMAL,         /* ConvertDecimal narrowing/widening, :768-844 */
	B2A_OP_UUID,            /* UUIDConversion, :480-495 */
	B2A_OP_DICT_INDEX,      /* SetSelectionVector*, :1227-1348 */
	B2A_OP_VALIDITY,        /* GetValidityMask, :63-108 (internal: emitted per nullable node) */
	B2A_OP_REE,             /* ColumnArrowToDuckDBRunEndEncoded + FlattenRunEnds, :551-767: the field has the type of its
	                           values child; run ends are searched per row and the decoded values gathered */
	B2A_OP_UNION,           /* sparse UNION, :1173-1221: tag bytes, the union's mask and the members' masks from the
	                           type_ids buffer and the members' own validity */
	B2A_OP_LISTVIEW32,      /* ConvertArrowListViewOffsetsTemplated, :166-203 (+vl): offsets + sizes buffers */
	B2A_OP_LISTVIEW64,      /* +vL */
	B2A_OP_BOOL8            /* arrow.bool8 over int8: ArrowBool8::ArrowToDuck, arrow_type_extension.cpp:344-350 (v != 0) */
} b2a_op;

/* One node of the schema tree, flattened in pre-order (the order of RecordBatch.nodes). */
typedef struct b2a_field_info {
	char name[128];
	char format[48];      /* Arrow C data interface format string ("l", "d:15,2,64", "u", "tdD", "+l", ...) */
	int32_t type_id;      /* b2a_type_id: the DuckDB LogicalType, per arrow_duck_schema.cpp:59-297 */
	int32_t op;           /* b2a_op for the node's value buffer */
	int32_t out_width;    /* bytes per row of the DuckDB vector (0: no flat data) */
	int32_t src_width;    /* bytes per value in the Arrow buffer (0: bit-packed / none) */
	int32_t n_children;
	int32_t first_child;  /* flat index of first child, -1 if none (children are at first_child, next_sibling, ...) */
	int32_t next_sibling; /* flat index, -1 if last */
	int32_t parent;       /* flat index, -1 for top-level columns */
	int32_t nullable;
	int32_t n_buffers;    /* IPC buffers this node consumes */
	int32_t decimal_width;
	int32_t decimal_scale;
	int64_t param;        /* fixed size / list size / multiplier / divisor */
	int64_t dict_id;      /* dictionary id, -1 if the field is not dictionary encoded */
	int32_t index_type_id; /* b2a_type_id of the dictionary indices (when dict_id >= 0) */
	int32_t index_width;
} b2a_field_info;

typedef struct b2a_ipc_buffer { /* scan_arrow_ipc argument: LIST(STRUCT(ptr POINTER, size UBIGINT)) */
	const void *ptr;
	uint64_t size;
} b2a_ipc_buffer;

typedef struct b2a_batch_info {
	uint64_t rows;
	uint64_t body_offset;  /* offset of the message body inside the stream image */
	uint64_t body_length;
	uint64_t meta_offset;  /* offset of the 8-byte message prefix */
	uint64_t buffer_bytes; /* sum of Buffer.length over projected columns (algorithmic input bytes) */
	uint32_t compressed;   /* BodyCompression present */
	uint32_t pad;
} b2a_batch_info;

/* Where one field of one record batch landed in the output arena.
 * Layout of a decoded column = the reference's per-2 048-row vectors laid end to end:
 * chunk k of the batch is rows [k*2048, (k+1)*2048) of every slab. */
typedef struct b2a_column_out {
	uint64_t rows;         /* FieldNode.length */
	uint64_t null_count;   /* FieldNode.null_count */
	uint64_t data_off;     /* rows * out_width bytes, B2A_NONE if the type has no flat data */
	uint64_t validity_off; /* ceil(rows/64) uint64 words; B2A_NONE = mask untouched (all valid) */
	uint64_t aux_off;      /* LIST/MAP: uint64[ceil(rows/2048)+1] child start row of each chunk (+ end);
	                          list views (B2A_OP_LISTVIEW*): uint64[ceil(rows/2048)][2] = {child start row, child rows} per
	                          chunk (the reference converts sum(sizes) child rows from the chunk's lowest offset) */
	uint64_t chars_off;    /* strings: stream-image offset of the character buffer (pointer base) */
	uint64_t chars_len;
	uint64_t stats_off;    /* uint64: valid-row count computed on device (popc of the mask) */
	uint64_t image_off;    /* B2A_SCAN_HOST_ALIAS: the vector aliases the stream image at this offset (rows * out_width
	                          bytes, the reference's DirectConversion); data_off is then B2A_NONE.  Else B2A_NONE. */
} b2a_column_out;

/* Pushed-down predicate: ConstantFilter{comparison, constant} on a top-level integer-like column. */
typedef enum b2a_cmp { B2A_CMP_EQ = 0, B2A_CMP_NE, B2A_CMP_LT, B2A_CMP_LE, B2A_CMP_GT, B2A_CMP_GE } b2a_cmp;
typedef struct b2a_predicate {
	uint32_t column; /* index into the projected top-level columns */
	uint32_t cmp;    /* b2a_cmp */
	int64_t constant;
} b2a_predicate;

typedef struct b2a_reader b2a_reader;
typedef struct b2a_scan b2a_scan;

const char *b2a_version(void);
const char *b2a_last_error(void);
int b2a_device_count(int *n);

/* ---------------- src/ipc: framing + flatbuffer slicing (host) ---------------- */
int b2a_reader_open_file(const char *path, b2a_reader **out);
int b2a_reader_open_buffers(const b2a_ipc_buffer *bufs, uint32_t n, b2a_reader **out);
void b2a_reader_close(b2a_reader *r);
/* n_top: projected top-level columns; n_flat: nodes of the (projected) flattened tree */
int b2a_reader_num_fields(b2a_reader *r, uint32_t *n_top, uint32_t *n_flat);
int b2a_reader_field(b2a_reader *r, uint32_t flat_index, b2a_field_info *out);
int b2a_reader_set_projection(b2a_reader *r, const char *const *names, uint32_t n);
int b2a_reader_num_batches(b2a_reader *r, uint64_t *n_batches, uint64_t *n_rows);
int b2a_reader_batch_info(b2a_reader *r, uint64_t batch, b2a_batch_info *out);
/* the host bytes of the stream (mmap of the file, or the caller's single buffer) */
int b2a_reader_image(b2a_reader *r, const void **data, uint64_t *size);
/* 1 when the message index of an Arrow FILE came from its footer (Block{offset, metaDataLength, bodyLength}) instead of
 * the sequential walk over the message prefixes the reference does (ipc_file_stream_reader.cpp:96-132); same messages
 * either way.  B2A_NO_FOOTER=1 in the environment forces the walk. */
int b2a_reader_used_footer(b2a_reader *r, int *used);
int b2a_reader_schema_metadata(b2a_reader *r, uint32_t i, const char **key, const char **value, uint32_t *n);

/* ---------------- src/scanner + kernels: record batches -> DuckDB vectors ---------------- */
int b2a_scan_create(b2a_reader *r, int device, uint64_t batch_lo, uint64_t batch_hi, b2a_scan **out);
/* flags: B2A_SCAN_HOST_ALIAS = the consumer of the vectors lives on the HOST next to the stream image, so columns the
 * reference converts with DirectConversion (arrow_conversion.cpp:454-461: FlatVector::SetData on the Arrow buffer, no
 * copy) are NOT moved through the GPU either: b2a_column_out.image_off points into the host image.  Only buffers that
 * need real conversion work (strings, bools, validity, lists, scalar transforms, swapped endianness) are uploaded,
 * converted and downloaded.  Such a scan can be used with b2a_read_arrow_host only. */
#define B2A_SCAN_HOST_ALIAS 0x1u
/* B2A_SCAN_DEVICE_ALIAS = the same zero-copy rule for consumers that live on the DEVICE (b2a_scan_filter,
 * b2a_scan_sum_product, a caller's own kernels): DirectConversion columns are uploaded with the image but not copied into
 * the arena — b2a_column_out.image_off says where in the DEVICE image (d_image + image_off - image_lo) the vector is,
 * data_off is B2A_NONE; validity masks and every column that needs conversion are decoded into the arena as usual.
 * A Q6-shaped scan (four fixed-width columns) then runs no decode launch at all.  For b2a_scan_decode /
 * b2a_read_arrow_device; not for b2a_read_arrow_host (a host consumer cannot follow a device address), not with
 * B2A_SCAN_HOST_ALIAS, not for compressed bodies. */
#define B2A_SCAN_DEVICE_ALIAS 0x2u
int b2a_scan_create_ex(b2a_reader *r, int device, uint64_t batch_lo, uint64_t batch_hi, uint32_t flags, b2a_scan **out);
void b2a_scan_destroy(b2a_scan *s);
int b2a_scan_sizes(b2a_scan *s, uint64_t *image_lo, uint64_t *image_bytes, uint64_t *out_bytes,
                   uint64_t *algorithmic_bytes, uint64_t *n_tasks, uint64_t *n_tiles);
/* bytes b2a_read_arrow_host moves host -> device per call (the merged image ranges the tasks read) */
int b2a_scan_upload_bytes(b2a_scan *s, uint64_t *bytes);
/* Host image -> device (one cudaMemcpyAsync per call; `host_image` should be pinned for full PCIe rate). */
int b2a_scan_upload(b2a_scan *s, void *cuda_stream);
/* Or: the [image_lo, image_lo+image_bytes) slice of the stream is already resident at d_image. */
int b2a_scan_set_device_image(b2a_scan *s, const void *d_image);
/* Launch the decode kernels.  d_out = caller-owned device arena of out_bytes (256-byte aligned).
 * string_ptr_base: address that stream-image offset `image_lo` maps to in the address space of whoever
 * consumes the string_t pointers (0 => the device image itself). */
int b2a_scan_decode(b2a_scan *s, void *d_out, uint64_t string_ptr_base, void *cuda_stream);
/* Block until `cuda_stream` drains and turn the device error word into a status. */
int b2a_scan_finish(b2a_scan *s, void *cuda_stream);
int b2a_scan_download(b2a_scan *s, const void *d_out, void *h_out, void *cuda_stream);
int b2a_scan_column(b2a_scan *s, uint64_t batch, uint32_t flat_field, b2a_column_out *out);
/* Dictionary of a dictionary-encoded field (decoded once per scan, shared by every batch). */
int b2a_scan_dictionary(b2a_scan *s, uint32_t flat_field, b2a_column_out *out, uint32_t *dict_flat_field);
uint64_t b2a_scan_gpu_launches(b2a_scan *s);
/* Compressed bodies (RecordBatch.compression = ZSTD; the reference decompresses buffer by buffer with ZSTD_decompress,
 * src/ipc/stream_reader/base_stream_reader.cpp:11-50; LZ4_FRAME, which the reference has no decompressor for, is
 * decoded as well): every projected buffer is expanded on the device into a scratch
 * region that lies `scratch_offset` bytes behind image offset image_lo, before the decode launch of the same call.
 * A caller-provided device image (b2a_scan_set_device_image) must therefore be scratch_offset + scratch_bytes + 64 long;
 * its first image_bytes hold the stream slice.  string_t pointers of a compressed scan point into the scratch region:
 * a host consumer copies it with b2a_scan_download_scratch and passes string_ptr_base = h_scratch - scratch_offset. */
int b2a_scan_scratch(b2a_scan *s, uint64_t *scratch_offset, uint64_t *scratch_bytes, uint64_t *n_jobs);
int b2a_scan_download_scratch(b2a_scan *s, void *h_scratch, void *cuda_stream);

/* Predicates -> selection vectors: for every 2 048-row chunk of every batch, sel_t[<=2048] + count.
 * d_sel: uint32[total_chunks * 2048]; d_count: uint32[total_chunks]; d_total: uint64 (sum of counts). */
int b2a_scan_filter(b2a_scan *s, const void *d_out, const b2a_predicate *preds, uint32_t n_preds, void *d_sel,
                    void *d_count, void *d_total, void *cuda_stream);
int b2a_scan_num_chunks(b2a_scan *s, uint64_t *n_chunks);
/* b2a_scan_filter without any device allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, same kernel, same results; the kernel's per-vector tables (16 bytes per vector and
 * predicate + 4 bytes per vector), which b2a_scan_filter keeps in cudaMalloc'ed memory cached in the scan, go into a
 * fixed 4 MiB `static __device__` array that every GPU loading the library holds until exit.  Limit: a scan whose
 * tables exceed 4 MiB (more than 49 932 vectors = 102 M rows under 5 predicates, 31 775 under 8) fails with
 * B2A_EINVAL and a message that says so.  The call issues no cudaMalloc / cudaFree / allocating library routine.  All
 * calls on a device share the array and each overwrites the part it needs (nothing has to start at zero): calls that
 * overlap on the GPU — other streams, other scans, other host threads — race; serialise them on one stream. */
int b2a_scan_filter_no_alloc(b2a_scan *s, const void *d_out, const b2a_predicate *preds, uint32_t n_preds, void *d_sel,
                             void *d_count, void *d_total, void *cuda_stream);
/* Q6-style consumer: sum(a*b) over selected rows of two int64 DECIMAL columns (HUGEINT accumulator, lo/hi). */
int b2a_scan_sum_product(b2a_scan *s, const void *d_out, uint32_t col_a, uint32_t col_b, const void *d_sel,
                         const void *d_count, void *d_result_i128, void *cuda_stream);

/* The two above fused for a consumer that needs no selection vectors (the Q6 shape, benchmark/lineitem.py:41): *d_total =
 * rows that pass the predicates (uint64), d_result_i128 = sum(a*b) over them (rows where a or b is NULL add nothing, like
 * DuckDB's sum).  One launch; nothing the reference has (it filters and aggregates above the scan). */
int b2a_scan_filter_sum_product(b2a_scan *s, const void *d_out, const b2a_predicate *preds, uint32_t n_preds,
                                uint32_t col_a, uint32_t col_b, void *d_total, void *d_result_i128, void *cuda_stream);

/* One call, host buffers in, host vectors out (upload, decode, download pipelined over `n_slices`). */
int b2a_read_arrow_host(b2a_scan *s, void *h_out_pinned, uint32_t n_slices);
/* The same for streams with compressed bodies (BodyCompression ZSTD / LZ4_FRAME): the decompressed buffers exist only
 * on the device, so the character buffers string_t point into come back as well — into `h_scratch_pinned`
 * (b2a_scan_scratch().scratch_bytes) — and the pointers refer to that copy.  Uncompressed scans ignore the argument. */
int b2a_read_arrow_host_scratch(b2a_scan *s, void *h_out_pinned, void *h_scratch_pinned, uint32_t n_slices);
/* Host buffers in, DEVICE vectors out, for consumers that run on the GPU (b2a_scan_filter, b2a_scan_sum_product): only
 * the buffers of the PROJECTED columns cross PCIe (the reference's projection pushdown, src/scanner/read_arrow.cpp:44-50,
 * base_stream_reader.cpp:146-212), slice by slice behind the decode launches; nothing is copied back.  Vectors land in
 * d_out (device, b2a_scan_sizes.out_bytes, 256-byte aligned); string_t point into the device image.  `cuda_stream` first
 * orders the pipeline behind the caller's earlier work on d_out and is then made to wait for the last decode: work
 * queued on it afterwards sees every vector.  Not for B2A_SCAN_HOST_ALIAS scans. */
int b2a_read_arrow_device(b2a_scan *s, void *d_out, uint32_t n_slices, void *cuda_stream);

/* ---------------- src/file_scanner: multi-file read_arrow (lists, globs, union_by_name, filename, hive) ----------------
 * Replaces what the reference gets from DuckDB's MultiFileReader framework through ArrowMultiFileInfo / ArrowFileScan
 * (src/file_scanner/arrow_multi_file_info.cpp:57-72,103-121, src/file_scanner/arrow_file_scan.cpp:11-26,75-80):
 *   bind     MultiFileReader::BindReader / BindUnionReader / BindOptions
 *            (duckdb/src/common/multi_file/multi_file_reader.cpp:220-286,528-580,653-672)
 *   mapping  MultiFileColumnMapper, by name (duckdb/src/common/multi_file/multi_file_column_mapper.cpp:134-192,494-585)
 *   finalize MultiFileReader::FinalizeChunk (multi_file_reader.cpp:435-458): column reference | cast | constant
 * Binding runs on the host; the cast expressions run on the device over whole decoded record batches
 * (b2a_multifile_finalize_*); constants (filename, hive partition values, union_by_name NULL columns) stay constants,
 * like DuckDB's CONSTANT_VECTORs. */
typedef struct b2a_multifile_options {
	uint32_t union_by_name;       /* read_arrow(..., union_by_name = true) */
	uint32_t filename;            /* filename = true: adds the VARCHAR column `filename_column` */
	uint32_t hive_partitioning;   /* hive_partitioning = true: key=value path components become columns */
	uint32_t hive_types_autocast; /* default 1: partition values that all parse as BIGINT / DATE get that type */
	const char *filename_column;  /* NULL: "filename" */
} b2a_multifile_options;

typedef enum b2a_mf_column_kind { B2A_MF_COL_DATA = 0, B2A_MF_COL_FILENAME = 1, B2A_MF_COL_HIVE = 2 } b2a_mf_column_kind;
typedef enum b2a_mf_map_kind {
	B2A_MF_MAP_LOCAL = 0,   /* the file's own vector (b2a_scan_column of local_flat) */
	B2A_MF_MAP_CAST = 1,    /* the file's vector cast to the global type (b2a_multifile_finalize_column) */
	B2A_MF_MAP_NULL = 2,    /* constant NULL (column missing from this file under union_by_name) */
	B2A_MF_MAP_CONSTANT = 3 /* constant value given as text (filename / hive partition value) */
} b2a_mf_map_kind;

typedef struct b2a_mf_column_info {
	char name[128];
	char type_name[256];  /* LogicalType::ToString() of the global column type */
	int32_t type_id;      /* b2a_type_id */
	int32_t decimal_width;
	int32_t decimal_scale;
	int32_t kind;         /* b2a_mf_column_kind */
} b2a_mf_column_info;

typedef struct b2a_mf_map_info {
	int32_t kind;         /* b2a_mf_map_kind */
	int32_t local_flat;   /* flat field index inside the file (LOCAL / CAST), else -1 */
	int32_t from_type_id; /* CAST: the file's column type */
	int32_t to_type_id;   /* the global column type */
	int32_t constant_is_null;
	int32_t pad;
	const char *constant; /* CONSTANT: text of the value (owned by the multi-file handle) */
} b2a_mf_map_info;

typedef struct b2a_multifile b2a_multifile;
typedef struct b2a_mf_finalize b2a_mf_finalize;

/* paths: file names and/or glob patterns (expanded in sorted order, like FileSystem::Glob) */
int b2a_multifile_open(const char *const *paths, uint32_t n_paths, const b2a_multifile_options *opt,
                       b2a_multifile **out);
void b2a_multifile_close(b2a_multifile *m);
int b2a_multifile_counts(b2a_multifile *m, uint32_t *n_files, uint32_t *n_columns, uint32_t *n_projected);
int b2a_multifile_column(b2a_multifile *m, uint32_t column, b2a_mf_column_info *out);
/* projection pushdown over GLOBAL column names; each file's reader is narrowed to the local columns it must deliver */
int b2a_multifile_set_projection(b2a_multifile *m, const char *const *names, uint32_t n);
int b2a_multifile_projected(b2a_multifile *m, uint32_t i, uint32_t *column);
/* the reader of file `file` is owned by the multi-file handle: scan it with b2a_scan_create, do not close it */
int b2a_multifile_file(b2a_multifile *m, uint32_t file, const char **path, b2a_reader **reader);
int b2a_multifile_map(b2a_multifile *m, uint32_t file, uint32_t column, b2a_mf_map_info *out);
/* the cast expressions of file `file` over a decoded scan of its reader */
int b2a_multifile_finalize_create(b2a_multifile *m, uint32_t file, b2a_scan *scan, b2a_mf_finalize **out);
void b2a_multifile_finalize_destroy(b2a_mf_finalize *fin);
int b2a_multifile_finalize_sizes(b2a_mf_finalize *fin, uint64_t *final_bytes, uint64_t *n_casts);
/* d_out: the decoded arena of the scan; d_final: device arena of final_bytes (256-byte aligned) for the cast results.
 * src_string_ptr_base: the string_ptr_base the scan was decoded with (0 = device image addresses);
 * dst_string_ptr_base: address final-arena offset 0 maps to for the consumer of VARCHAR results (0 = d_final). */
int b2a_multifile_finalize_run(b2a_mf_finalize *fin, const void *d_out, void *d_final, uint64_t src_string_ptr_base,
                               uint64_t dst_string_ptr_base, void *cuda_stream);
/* waits for the stream and maps the device error word to ConversionException / NotImplementedException */
int b2a_multifile_finalize_finish(b2a_mf_finalize *fin, void *cuda_stream);
/* location of a CAST column of one batch: data_off / chars_off are offsets into the FINAL arena, validity_off and
 * stats_off into the SCAN arena (the mask is shared with the source vector) */
int b2a_multifile_finalize_column(b2a_mf_finalize *fin, uint64_t batch, uint32_t column, b2a_column_out *out);
uint64_t b2a_multifile_finalize_gpu_launches(b2a_mf_finalize *fin);

/* ---------------- src/writer + kernels: DuckDB vectors -> Arrow IPC stream ----------------
 * Replaces ColumnDataCollectionSerializer::Serialize (src/writer/column_data_collection_serializer.cpp:80-120),
 * ArrowAppender (duckdb/src/common/arrow/arrow_appender.cpp:37-131 + appender/<type>_data.hpp), the nanoarrow
 * encoder it calls (ArrowIpcEncoderEncodeSimpleRecordBatch) and ArrowStreamWriter (src/writer/arrow_stream_writer.cpp
 * :55-82).  The input is a whole table of flat DuckDB vectors resident in DEVICE memory; it is cut into record batches
 * of `row_group_size` rows (COPY option, default 122 880: src/writer/write_arrow_stream.cpp:33,70-76). */
typedef struct b2a_vector {
	int32_t type_id;          /* b2a_type_id of the DuckDB LogicalType */
	int32_t decimal_width;    /* DECIMAL(width, scale) */
	int32_t decimal_scale;
	int32_t n_children;       /* STRUCT: fields; LIST / ARRAY: 1; MAP: 1 (the STRUCT(key, value) entry vector);
	                             UNION: the members (flat types), data = the UTINYINT tags, validity = the union's */
	int64_t param;            /* ARRAY: list size */
	const char *name;         /* column name / struct field name (NULL: "") */
	const char *timezone;     /* TIMESTAMP_TZ: ClientProperties.time_zone (NULL: "UTC") */
	const void *data;         /* device: values | string_t[] | list_entry_t[] | bool bytes; NULL for STRUCT/ARRAY/SQLNULL */
	const uint64_t *validity; /* device: ValidityMask words (bit = 1 valid); NULL = all valid */
	uint64_t rows;            /* vector length (LIST child: ListVector::GetListSize) */
	const struct b2a_vector *children;
} b2a_vector;

typedef struct b2a_write_options {
	uint64_t row_group_size;      /* 0 = 122 880 */
	uint32_t n_kv_metadata;       /* COPY option kv_metadata (write_arrow_stream.cpp:87-105) */
	const char *const *kv_keys;
	const char *const *kv_values;
	uint32_t flags;               /* B2A_WRITE_* (ClientProperties of the session, arrow_stream_writer.cpp:11) */
} b2a_write_options;

/* SET arrow_large_buffer_size = true (ClientProperties.arrow_offset_size = LARGE): VARCHAR / UUID -> "U", BLOB / BIT
 * -> "Z", LIST -> "+L" with int64 offsets (arrow_appender.cpp:184-188,237-251,288-292; arrow_converter.cpp:179-196,
 * 262-279,289-293).  MAP keeps int32 offsets: Arrow has no large map (arrow_appender.cpp:296-299). */
#define B2A_WRITE_LARGE_OFFSETS 0x1u
/* SET arrow_output_list_view = true (ClientProperties.arrow_use_list_view): LIST -> "+vl" ("+vL" with LARGE_OFFSETS):
 * ArrowListViewData, an offsets and a sizes buffer of `rows` entries each (appender/list_view_data.hpp:59-103,
 * arrow_appender.cpp:281-286, arrow_converter.cpp:282-287).  MAP is not affected. */
#define B2A_WRITE_LIST_VIEW 0x2u
/* SET produce_arrow_string_view = true: VARCHAR -> "vu" (ArrowVarcharToStringViewData, appender/varchar_data.hpp:129-202;
 * arrow_appender.cpp:234-235, arrow_converter.cpp:189-190): 16-byte views + ONE variadic data buffer holding the strings
 * longer than 12 bytes end to end (RecordBatch.variadicBufferCounts = [1] per such field).  BLOB / BIT keep offsets. */
#define B2A_WRITE_STRING_VIEW 0x4u
/* SET arrow_lossless_conversion = true (ClientProperties.arrow_lossless_conversion; arrow_converter.cpp:118-128,161-279,
 * arrow_appender.cpp:161-191, arrow_type_extension.cpp:366-390): types Arrow has no 1:1 type for keep their bits and get
 * an extension type instead of a lossy stand-in — UUID -> arrow.uuid over w:16 (big-endian bytes), HUGEINT / UHUGEINT ->
 * arrow.opaque{DuckDB, hugeint | uhugeint} over w:16, TIME_TZ -> arrow.opaque{DuckDB, time_tz} over w:8, BIT / VARINT ->
 * arrow.opaque{DuckDB, bit | varint} over z / Z, BOOLEAN -> arrow.bool8 over c (one byte per row), JSON -> arrow.json. */
#define B2A_WRITE_LOSSLESS 0x8u

typedef struct b2a_writer b2a_writer;

int b2a_writer_create(const b2a_vector *columns, uint32_t n_columns, uint64_t rows, const b2a_write_options *opt,
                      int device, b2a_writer **out);
void b2a_writer_destroy(b2a_writer *w);
/* Pass 1: data-dependent sizes (string bytes, list child counts) are reduced on the device, read back, and the
 * stream is laid out (message offsets, per-buffer body offsets).  Synchronises `cuda_stream`. */
int b2a_writer_measure(b2a_writer *w, void *cuda_stream);
/* algorithmic_bytes = what the encode launch must read and write: the bytes of the input vectors (validity words, values,
 * string_t / list_entry_t records), the characters of the strings longer than 12 bytes (a non-inlined string_t holds a
 * pointer into the heap; summed by pass 1) and every byte of the produced bodies. */
int b2a_writer_sizes(b2a_writer *w, uint64_t *stream_bytes, uint64_t *n_batches, uint64_t *body_bytes,
                     uint64_t *algorithmic_bytes);
/* Pass 2: every buffer of every record batch is written at its final stream offset inside d_stream
 * (device, stream_bytes, 256-byte aligned).  Message headers are NOT written by the device. */
int b2a_writer_encode(b2a_writer *w, void *d_stream, void *cuda_stream);
/* Waits for the encode, reads back null counts and the device error word. */
int b2a_writer_finish(b2a_writer *w, void *cuda_stream);
/* Host side of the stream (after b2a_writer_finish): Schema message, one RecordBatch message header per batch and
 * the end-of-stream marker, each written at its stream offset inside h_stream (which already holds, or will receive,
 * the D2H copy of the bodies). */
int b2a_writer_headers(b2a_writer *w, void *h_stream);
/* rows / meta_offset / body_offset / body_length of record batch i of the produced stream */
int b2a_writer_batch(b2a_writer *w, uint64_t i, b2a_batch_info *out);
uint64_t b2a_writer_gpu_launches(b2a_writer *w);
/* The encapsulated Schema message of the stream (ArrowStreamWriter::WriteSchema, src/writer/arrow_stream_writer.cpp:55-60;
 * types = ArrowConverter::ToArrowSchema, duckdb/src/common/arrow/arrow_converter.cpp:115-432).  Host only: available
 * right after b2a_writer_create, also without a device.  The pointer stays valid until b2a_writer_destroy. */
int b2a_writer_schema_message(b2a_writer *w, const void **data, uint64_t *size);
/* Bytes of the stream before its first record batch message: the Schema message followed by one DictionaryBatch message
 * (header + body) per ENUM field, in dictionary-id order.  A caller that cuts record-batch messages out of several
 * streams (COPY writes one row group per call, to_arrow_ipc one blob per message) sends [0, preamble) of the first
 * stream once.  Valid after b2a_writer_measure; the dictionary bodies are there after b2a_writer_encode + the D2H copy. */
int b2a_writer_preamble(b2a_writer *w, uint64_t *bytes);

/* ---------------- host-facing writer: HOST vectors in, HOST stream bytes out ----------------
 * What COPY ... (FORMAT ARROWS) and to_arrow_ipc hold is host memory: the flat vectors of one row group
 * (ColumnDataCollectionSerializer::Serialize(ColumnDataCollection&) concatenates the collection into ONE DataChunk and
 * runs ArrowConverter::ToArrowArray over it, src/writer/column_data_collection_serializer.cpp:80-115).  `host_columns`
 * is that chunk as a b2a_vector tree with HOST pointers (data, validity; string_t records may point into DuckDB's
 * string heaps).  The call uploads the tree, runs the b2a_writer kernels (measure + encode), and returns the complete
 * stream — Schema message, one RecordBatch message per `row_group_size` rows, end-of-stream marker — in pinned host
 * memory.  `writer` stays valid until release: b2a_writer_batch(writer, i) gives the byte range of message i, so a
 * COPY sink writes the schema once and then only the RecordBatch messages of every row group. */
typedef struct b2a_host_stream {
	void *data;            /* pinned host memory, `size` bytes: the stream */
	uint64_t size;
	uint64_t n_batches;
	b2a_writer *writer;    /* owned by the stream */
	uint64_t gpu_launches; /* kernels launched for this stream */
	void *priv;
} b2a_host_stream;
int b2a_write_arrow_host(const b2a_vector *host_columns, uint32_t n_columns, uint64_t rows, const b2a_write_options *opt,
                         int device, b2a_host_stream *out);
void b2a_host_stream_release(b2a_host_stream *s);

/* Pinned (page-locked) host memory for the arenas b2a_read_arrow_host fills: a DuckDB-side caller must not link the
 * CUDA runtime itself. */
int b2a_pinned_alloc(void **out, uint64_t bytes);
void b2a_pinned_free(void *p);

#ifdef __cplusplus
}
#endif
#endif /* B2ARROW_H */
