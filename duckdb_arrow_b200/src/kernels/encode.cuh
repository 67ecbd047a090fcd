// Launch interface of the encode kernels (host side of src/kernels, writer direction).
#pragma once
#include "b2arrow.h"
#include "enc_task.h"

#include <stdint.h>

namespace b2a {

struct EncodeLaunch {
	const b2a_enc_task *d_tasks = nullptr;
	const uint32_t *d_tile_task = nullptr;
	uint32_t n_tasks = 0;
	uint32_t n_tiles = 0;
	uint8_t *stream_out = nullptr;  // device image of the IPC stream being produced
	uint32_t *d_err = nullptr;
	uint64_t *d_heap_bytes = nullptr; // measure: += bytes of the strings longer than 12 (what pass 2 reads from the heap)
	int sm_count = 148;
	bool full = false; // some task needs an op of the full instantiation (B2A_ENC_OP_IS_RARE)
};

// each returns cudaError_t as int
int LaunchEncMeasure(const EncodeLaunch &l, void *stream); // tile sums + per-task scan
int LaunchEncExpand(const EncodeLaunch &l, void *stream);  // child selection vectors
int LaunchEncode(const EncodeLaunch &l, void *stream);     // pass 2
// member masks of a UNION vector (write_plan.hpp: WVec::union_member); out_words = uint64[ceil(rows / 64)]
int LaunchEncUnionMask(const void *tags, const void *union_valid, const void *member_valid, uint32_t member,
                       uint64_t rows, void *out_words, int sm_count, void *stream);

} // namespace b2a
