// Device-visible descriptor of one encode task of the writer = one Arrow buffer (or offsets+data pair) of one
// field of one record batch.  The write planner (src/writer/write_plan.cpp) emits them; b2a_encode_kernel walks
// the whole table in one launch, like the decode side.
#pragma once
#include <stdint.h>

/* ops: each restates one ArrowAppender routine (duckdb/src/include/duckdb/common/arrow/appender/...) */
enum b2a_enc_op {
	B2A_ENC_NONE = 0,
	B2A_ENC_VALIDITY,     /* AppendValidity, append_data.hpp:127-149 (+ null_count) */
	B2A_ENC_COPY,         /* ArrowScalarBaseData<T>::Append, scalar_data.hpp:90-119 */
	B2A_ENC_WIDEN16,      /* ArrowScalarData<int32_t,int16_t>: DECIMAL(<=4) int16 -> decimal32 */
	B2A_ENC_INTERVAL,     /* ArrowIntervalConverter: interval_t -> month-day-nano, NULLs skipped */
	B2A_ENC_BOOL,         /* ArrowBoolData::Append, bool_data.cpp:12-39 */
	B2A_ENC_STRING32,     /* ArrowVarcharData<string_t, ArrowVarcharConverter, int32_t>, varchar_data.hpp:53-111 */
	B2A_ENC_STRING64,     /* ... int64_t offsets (arrow_large_buffer_size) */
	B2A_ENC_LIST32,       /* ArrowListData<int32_t>::AppendOffsets, list_data.hpp:56-98 */
	B2A_ENC_LIST64,
	B2A_ENC_EXPAND_ARRAY, /* child selection of a fixed-size list under a selection (fixed_size_list_data.cpp:16-28) */
	B2A_ENC_FILL_VALID,   /* all-valid bitmap: ResizeValidity(.., 0xFF), append_data.hpp:116-119 */
	/* ops below run in the full instantiation of b2a_encode_kernel only (rare types; keeps the common one lean) */
	B2A_ENC_TIMETZ,       /* ArrowScalarData<int64_t, dtime_tz_t, ArrowTimeTzConverter>: bits >> 24, NULLs skipped */
	B2A_ENC_UUID32,       /* ArrowVarcharData<hugeint_t, ArrowUUIDConverter, int32_t>: 36 characters per valid row */
	B2A_ENC_UNION_TAG,    /* ArrowUnionData::Append, union_data.cpp:24-52: type id = tag of a valid row, else 0 */
	B2A_ENC_UUID64,       /* ArrowVarcharData<hugeint_t, ArrowUUIDConverter> (int64 offsets, arrow_large_buffer_size) */
	B2A_ENC_LISTVIEW32,   /* ArrowListViewData<int32_t>::AppendListMetadata, list_view_data.hpp:59-103: offsets -> dst,
	                         sizes -> dst2, `rows` entries each; a NULL row = {running offset, 0} */
	B2A_ENC_LISTVIEW64,
	B2A_ENC_UUID_BLOB,    /* ArrowScalarData<hugeint_t, hugeint_t, ArrowUUIDBlobConverter>, scalar_data.hpp:70-87: the UUID's 16
	                         bytes big-endian (upper ^ sign bit first), NULL rows skipped */
	B2A_ENC_STRINGVIEW    /* ArrowVarcharToStringViewData::Append, varchar_data.hpp:136-187: arrow_string_view_t[rows] -> dst,
	                         the characters of the strings longer than 12 bytes -> dst2 (buffer index 0, int32 offset) */
};

#define B2A_ENC_OP_IS_RARE(op) ((op) >= B2A_ENC_TIMETZ)

#define B2A_ENC_F_LAST_PAD 0x1u /* unused */

#define B2A_ENC_STR_TILE 1024u /* rows per tile of the scan-based ops (256 threads x 4 rows) */

/* writer device error codes (atomicMax), continuing the numbering of task.h */
#define B2A_DEV_ENC_OFFSET_I32 16u /* "Arrow Appender: The maximum total string size for regular string buffers is ..." */
#define B2A_DEV_ENC_LIST_I32 17u   /* "Arrow Appender: The maximum combined list offset for regular list buffers is ..." */

struct __attribute__((aligned(16))) b2a_enc_task {
	uint64_t src;      /* device address: input vector data */
	uint64_t validity; /* device address: ValidityMask words of the input vector, 0 = all valid */
	uint64_t sel;      /* device address: uint32 selection (input row of output row i), 0 = identity */
	uint64_t row0;     /* identity mapping: input row of output row 0 */
	uint64_t rows;     /* output rows */
	uint64_t dst;      /* stream offset of the main buffer (values / offsets / bits); EXPAND: device address of child sel */
	uint64_t dst2;     /* stream offset of the character buffer */
	uint64_t scratch;  /* device address: uint64 per tile of this task — tile sums, then (after the scan) tile bases */
	uint64_t counter;  /* device address of a uint64: null_count (VALIDITY, atomicAdd) / total (scan kernel) */
	int64_t param;     /* EXPAND_ARRAY: list size */
	uint32_t tile0;
	uint32_t tile_rows;
	uint16_t op;
	uint8_t src_width;
	uint8_t dst_width;
	uint32_t flags;
	uint64_t dst_len; /* unpadded byte length of the main buffer (zero padding up to 8 is written by the last tile) */
};
