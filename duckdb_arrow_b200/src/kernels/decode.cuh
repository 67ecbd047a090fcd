// Launch interface of the decode / filter kernels (host side of src/kernels).
#pragma once
#include "b2arrow.h"
#include "task.h"

#include <stdint.h>

#define B2A_NONE_OFF 0xFFFFFFFFFFFFFFFFull

namespace b2a {

struct DecodeLaunch {
	const b2a_task *d_tasks = nullptr;
	const uint32_t *d_tile_task = nullptr; // tile -> task index
	uint32_t n_tiles = 0;   // tiles to process in this launch ...
	uint32_t tile_begin = 0; // ... starting at this global tile
	const uint8_t *image_origin = nullptr; // device address of stream-image offset 0 (may lie before the allocation)
	uint8_t *out = nullptr;                // output arena
	uint64_t string_ptr_origin = 0;        // address of stream-image offset 0 in the consumer's address space
	uint32_t *d_err = nullptr;
	const uint64_t *d_variadic = nullptr;  // image offsets of view data buffers
	int sm_count = 148;
	bool lean = false; // every task is COPY / VALIDITY / BOOL / STRING32 / STRING64: the lean kernel instantiation
};

// returns cudaError_t as int
int LaunchDecode(const DecodeLaunch &l, void *stream);
// tiles of tasks flagged B2A_TASK_F_PHASE2 (they read vectors the decode launch produced): b2a_ree_kernel
int LaunchPhase2(const DecodeLaunch &l, void *stream);

struct FilterLaunch {
	const uint8_t *out = nullptr;       // decoded arena
	const uint64_t *d_chunk_col = nullptr; // per chunk, per predicate: arena offset of the chunk's column data
	const uint32_t *d_chunk_rows = nullptr;
	uint64_t n_chunks = 0;
	uint32_t n_preds = 0;
	b2a_predicate preds[8];
	uint8_t widths[8];  // bytes per value of each predicate's column (4 or 8)
	uint32_t *d_sel = nullptr;
	uint32_t *d_count = nullptr;
	unsigned long long *d_total = nullptr;
	int sm_count = 148;
};
int LaunchFilter(const FilterLaunch &l, void *stream);
// filter + sum(a * b) in one kernel: l.d_chunk_col holds n_preds + 2 {data, validity} pairs per chunk (the predicates',
// then a's and b's); l.d_sel / l.d_count are not used; l.d_total += passing rows, d_result (lo, hi) += the 128-bit sum
int LaunchFilterSum(const FilterLaunch &l, unsigned long long *d_result, void *stream);

// The no-allocation variant (b2a_scan_filter_no_alloc): the per-chunk tables live in a file-scope __device__ array of
// filter.cu instead of cudaMalloc'ed memory.  l.d_chunk_col / l.d_chunk_rows are ignored; the HOST tables are copied into
// the array on `stream` (h_chunk_col: 2 * n_preds uint64 per chunk, h_chunk_rows: one uint32 per chunk) before the launch.
// Returns 0, a cudaError_t, or -1 when the tables do not fit FilterNoAllocScratchBytes().
uint64_t FilterNoAllocScratchBytes();
uint64_t FilterNoAllocBytesNeeded(uint64_t n_chunks, uint32_t n_preds);
int LaunchFilterNoAlloc(const FilterLaunch &l, const uint64_t *h_chunk_col, const uint32_t *h_chunk_rows, void *stream);

struct SumProductLaunch {
	const uint8_t *out = nullptr;
	const uint64_t *d_chunk_col = nullptr; // per chunk: {a data, a validity, b data, b validity} arena offsets
	const uint32_t *d_chunk_rows = nullptr;
	const uint32_t *d_sel = nullptr;
	const uint32_t *d_count = nullptr;
	uint64_t n_chunks = 0;
	unsigned long long *d_result = nullptr; // lo, hi of the int128 sum
	int sm_count = 148;
};
int LaunchSumProduct(const SumProductLaunch &l, void *stream);

} // namespace b2a
