/*
 * oracle.h — CPU restatement of the reference's Arrow -> DuckDB vector conversion.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is linked into, imported by or called from the product
 * (duckdb_arrow_b200/, include/); only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference leg may use it, as the checker or the timed CPU baseline.
 *
 * Follows duckdb/src/function/table/arrow_conversion.cpp (ColumnArrowToDuckDB and helpers) and
 * duckdb/src/function/table/arrow.cpp:172-201 (the 2 048-row chunk loop), function by function;
 * each function in oracle.c cites the lines it restates.  Input is the Arrow C data interface
 * (buffers produced by pyarrow's IPC reader in the tests, i.e. NOT by the product's own slicer).
 *
 * Parity pinning: the oracle's output is checked against the golden contents (*.json.gz) of the
 * arrow-testing integration streams, the expected values of the reference's SQL tests
 * (test/sql/read_arrow.test, read_arrow_file.test) and pyarrow's decoding of the same files
 * (tests/test_oracle_golden.py).  The reference's own C++ could not be compiled here without
 * running its cmake build, so no oracle/_ref exists.
 */
#ifndef B2A_ORACLE_H
#define B2A_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define O_VECTOR_SIZE 2048

/* logical type ids: same numbering as b2a_type_id in include/b2arrow.h (kept in sync by tests/test_abi.py) */
enum {
	O_INVALID = 0, O_SQLNULL, O_BOOLEAN, O_TINYINT, O_SMALLINT, O_INTEGER, O_BIGINT, O_UTINYINT, O_USMALLINT,
	O_UINTEGER, O_UBIGINT, O_HUGEINT, O_UHUGEINT, O_FLOAT, O_DOUBLE, O_DECIMAL, O_VARCHAR, O_BLOB, O_DATE, O_TIME,
	O_TIMESTAMP, O_TIMESTAMP_SEC, O_TIMESTAMP_MS, O_TIMESTAMP_NS, O_TIMESTAMP_TZ, O_INTERVAL, O_UUID, O_LIST,
	O_STRUCT, O_MAP, O_ARRAY, O_UNION, O_TIME_TZ, O_BIT, O_VARINT, O_JSON,
	O_ENUM /* writer side only (oracle_write.py): the scan never produces it */
};

/* ArrowVariableSizeType (duckdb/src/include/duckdb/function/table/arrow/enum/arrow_variable_size_type.hpp) */
enum { O_SIZE_NORMAL = 0, O_SIZE_SUPER = 1, O_SIZE_FIXED = 2, O_SIZE_VIEW = 3 };
/* ArrowDateTimeType */
enum { O_DT_MILLISECONDS = 0, O_DT_MICROSECONDS, O_DT_NANOSECONDS, O_DT_SECONDS, O_DT_DAYS, O_DT_MONTHS, O_DT_MONTH_DAY_NANO };

typedef struct OType {
	int32_t id;          /* DuckDB logical type of the VALUES (for dictionary columns: of the dictionary values) */
	int32_t width;       /* bytes per row of the DuckDB vector */
	int32_t size_type;   /* strings / blobs / lists */
	int32_t datetime;    /* O_DT_* */
	int32_t bit_width;   /* decimal: 32 / 64 / 128 */
	int32_t index_type;  /* dictionary: logical type of the indices (O_TINYINT ... O_UBIGINT), else 0 */
	int64_t fixed_size;  /* fixed-size binary / fixed-size list */
	int32_t n_children;
	int32_t has_dictionary;
	struct OType **children;
	/* O_UNION (sparse): width 1 = the UTINYINT tag child of DuckDB's STRUCT{tag, members...}; children = the members */
	int32_t run_end_encoded; /* 1: children = {run_ends type, values type}; id / width are those of the values */
	int32_t bool8;           /* O_BOOLEAN only: arrow.bool8 — storage converted as TINYINT, then ArrowBool8::ArrowToDuck */
	int32_t list_view;       /* O_LIST only: ArrowListInfo::IsView() — buffers = validity, offsets, sizes */
} OType;

/* the subset of struct ArrowArray the converter reads */
typedef struct OArray {
	int64_t length;
	int64_t null_count;
	int64_t offset;
	int64_t n_buffers;
	int64_t n_children;
	const void **buffers;
	struct OArray **children;
	struct OArray *dictionary;
} OArray;

/* Converted column of one record batch: the per-chunk DuckDB vectors laid end to end. */
typedef struct OSlab {
	uint64_t rows;
	uint64_t cap;
	int32_t width;
	int32_t type_id;
	uint8_t *data;   /* rows * width */
	uint8_t *valid;  /* 1 byte per row (the bits of the ValidityMask) */
	int32_t n_children;
	int32_t all_null; /* SQLNULL: constant NULL vector */
	struct OSlab **children;
	uint32_t *sel;      /* dictionary vectors: selection per row */
	struct OSlab *dict; /* dictionary values (+1 NULL slot when the column can contain NULL) */
} OSlab;

/* Runs the reference's scan loop over one column of one record batch: for chunk_offset = 0, 2048, ...
 * converts min(2048, length - chunk_offset) rows exactly like ArrowToDuckDB and appends the chunk's vector
 * to the slab.  Returns 0, or an error code (1 conversion, 2 not implemented, 3 invalid input) with message. */
int oracle_scan_column(const OArray *array, const OType *type, int64_t parent_offset, OSlab **out, char *errbuf,
                       int errlen);
void oracle_slab_free(OSlab *s);

/* Full scan of a struct array (= record batch) — used by the timed CPU baseline: converts every column chunk by
 * chunk and discards the vectors; returns a checksum so the work cannot be optimised away. */
uint64_t oracle_scan_batch_checksum(const OArray *batch, OType *const *types, int n_cols);

/* a1 + a2 + the scan: DecodeBody through the 4 KiB bounce buffer, nanoarrow FULL validation, scan (see oracle.c) */
uint8_t *oracle_decode_body(const uint8_t *file_image, uint64_t body_len);
int oracle_validate_full(const OArray *array, const OType *type);
uint64_t oracle_read_arrow_batch_checksum(const uint8_t *body, uint64_t body_len, const OArray *batch,
                                          OType *const *types, int n_cols);

#ifdef __cplusplus
}
#endif
#endif
