// Device-visible descriptor of one conversion task = one IPC buffer (or buffer pair) of one field of
// one record batch.  The host slicer (src/ipc) emits them, the planner (src/scanner) assigns output
// offsets and tile ranges, and ONE launch of b2a_decode_kernel walks the whole table.
#pragma once
#include <stdint.h>

#define B2A_MAX_MASKS 4
#define B2A_TASK_F_BSWAP 0x1u      /* stream is big-endian: swap values/offsets (K8) */
#define B2A_TASK_F_SIGNED 0x2u     /* DICT_INDEX / DECIMAL: source is signed */
#define B2A_TASK_F_ADD_NULL 0x4u   /* VALIDITY of a dictionary: extra NULL slot at index `rows` */
#define B2A_TASK_F_BLOB 0x8u
#define B2A_TASK_F_PHASE2 0x10u    /* reads vectors another task decoded (REE gather, UNION): runs in the second launch */

/* device error codes (atomicMax into the scan's error word) */
#define B2A_DEV_OK 0u
#define B2A_DEV_BAD_OFFSETS 1u     /* offsets not monotone / beyond the data buffer (nanoarrow FULL validation) */
#define B2A_DEV_MUL_OVERFLOW 2u    /* "Could not convert ... to Microsecond" */
#define B2A_DEV_STRING_4GB 3u      /* "DuckDB does not support Strings over 4GB" */
#define B2A_DEV_INDEX_U32 4u       /* "DuckDB only supports indices that fit on an uint32" */
#define B2A_DEV_UNION_TAG 5u       /* "Arrow union tag out of range: %d" */
#define B2A_DEV_DICT_INDEX 6u      /* dictionary index outside [0, dictionary length) */
#define B2A_DEV_VIEW_RANGE 7u      /* string view: buffer index / offset + length outside its variadic data buffer */
#define B2A_DEV_LIST_CHILD 8u      /* list: last offset beyond the child array */

struct b2a_mask_ref {
	uint64_t src; /* image offset of an Arrow validity bitmap */
	uint64_t div; /* row r of this task reads bit r/div (fixed-size-list broadcast) */
};

struct __attribute__((aligned(16))) b2a_task {
	uint64_t src;      /* image offset: values / offsets / views */
	uint64_t src2;     /* image offset: character data (strings) */
	uint64_t src2_len; /* bytes of character data */
	uint64_t dst;      /* output arena offset */
	uint64_t aux;      /* output arena offset: list chunk starts; validity: popcount cell */
	uint64_t rows;
	int64_t param;
	uint32_t tile0;     /* first global tile of this task */
	uint32_t tile_rows; /* rows per tile (multiple of 2048) */
	uint16_t op;        /* b2a_op */
	uint8_t src_width;
	uint8_t dst_width;
	uint32_t flags;
	uint32_t n_masks;   /* effective validity = AND of masks[0..n_masks) */
	uint32_t variadic0; /* STRING_VIEW: first entry of this task's variadic buffer table */
	struct b2a_mask_ref masks[B2A_MAX_MASKS];
};
