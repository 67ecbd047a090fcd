// Launch interface of the multi-file cast kernel (host side).  The casts are the per-column expressions
// MultiFileReader::FinalizeChunk evaluates when a file's column type differs from the global column type
// (duckdb/src/common/multi_file/multi_file_reader.cpp:435-458, multi_file_column_mapper.cpp:494-585).
#pragma once
#include <stdint.h>

#define B2A_CAST_DEV_OK 0u
#define B2A_CAST_DEV_RANGE 1u      /* ConversionException: value out of range for the destination type */
#define B2A_CAST_DEV_PARSE 2u      /* ConversionException: could not convert string to the destination type */
#define B2A_CAST_DEV_PRECISION 3u  /* NotImplemented: VARCHAR -> DOUBLE where the reference reads past the end of the string */

struct __attribute__((aligned(16))) b2a_cast_task {
	uint64_t src;      /* scan arena offset: source vector data */
	uint64_t validity; /* scan arena offset: ValidityMask words, UINT64_MAX = all valid */
	uint64_t dst;      /* final arena offset: destination vector data */
	uint64_t heap;     /* final arena offset: string heap of this task (heap_stride bytes per row) */
	uint64_t rows;
	uint32_t tile0;    /* first global tile (2 048 rows each) */
	uint16_t from, to; /* b2a_type_id */
	uint8_t src_width, dst_width;
	uint8_t heap_stride;                  /* VARCHAR results: B2A_CAST_HEAP_STRIDE or _WIDE */
	uint8_t decimal_width, decimal_scale; /* DECIMAL sources */
	uint8_t dst_decimal_width, dst_decimal_scale; /* DECIMAL results */
	uint8_t pad[1];
};

#define B2A_CAST_HEAP_STRIDE 32u      /* integers (<= 20 chars), FLOAT / DOUBLE (<= 24), DATE (<= 18), TIME (<= 15) */
#define B2A_CAST_HEAP_STRIDE_WIDE 48u /* TIMESTAMP (<= 33), HUGEINT / DECIMAL of more than 18 digits (<= 40) */

namespace b2a {

struct CastLaunch {
	const b2a_cast_task *d_tasks = nullptr;
	const uint32_t *d_tile_task = nullptr;
	uint32_t n_tiles = 0;
	const uint8_t *scan_arena = nullptr; /* decoded vectors (device) */
	uint8_t *final_arena = nullptr;      /* cast results (device) */
	/* VARCHAR sources: string_t pointers in the scan arena are in the consumer's address space; adding this delta
	 * turns them into device addresses of the stream image */
	int64_t src_ptr_delta = 0;
	/* VARCHAR results: address of final-arena offset 0 in the consumer's address space */
	uint64_t dst_ptr_origin = 0;
	uint32_t *d_err = nullptr;
	int sm_count = 148;
};

int LaunchCast(const CastLaunch &l, void *stream);

} // namespace b2a
