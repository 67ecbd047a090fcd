// Internal definitions behind the opaque handles of include/b2arrow.h, shared by the capi*.cu translation units.
#pragma once
#include "b2arrow.h"
#include "ipc/ipc_reader.hpp"
#include "kernels/scatter.cuh"
#include "scanner/host_stage.hpp"
#include "scanner/scan_plan.hpp"

#include <string>
#include <vector>

struct b2a_reader {
	b2a::Reader reader;
	std::vector<std::string> meta_keys, meta_vals;
};

struct b2a_scan {
	b2a_reader *r = nullptr;
	b2a::ScanPlan plan;
	int device = 0;
	int sm_count = 148;
	b2a_task *d_tasks = nullptr;
	uint32_t *d_tile_task = nullptr;
	uint64_t *d_variadic = nullptr;
	uint32_t *d_err = nullptr;
	uint8_t *d_image_owned = nullptr;   // allocated by b2a_scan_upload
	const uint8_t *d_image = nullptr;   // device address of image offset plan.image_lo
	uint64_t launches = 0;
	// filter support
	uint64_t *d_chunk_col = nullptr;   // filter: per chunk, per predicate {data, validity}
	uint64_t *d_chunk_col2 = nullptr;  // sum_product: per chunk {a, va, b, vb}
	uint64_t *d_chunk_col3 = nullptr;  // filter_sum_product: per chunk, per predicate {data, validity}, then {a, va, b, vb}
	uint32_t *d_chunk_rows = nullptr;
	std::vector<uint32_t> chunk_col_key, chunk_col2_key, chunk_col3_key;
	// compressed bodies: decompress jobs, per-warp decoder workspaces, job counter + error word
	b2a_decomp_job *d_jobs = nullptr;
	uint32_t *d_job_order = nullptr; // all jobs, largest output first (whole-scan launches: shortens the tail)
	void *d_workspaces = nullptr;
	uint32_t *d_decomp_state = nullptr; // [0] job counter, [1] zstd error code
	uint32_t n_workers = 0;
	// host path over a pageable image (mmap'ed file): staged uploads, see scanner/host_stage.hpp
	int stage_mode = 0; // 0 undecided, 1 direct cudaMemcpyAsync per range (pinned image), 2 staged
	int stage_io = 0;   // staged: 0 memcpy from the image (mmap / caller memory), 1 pread(fd), 2 pread(O_DIRECT fd)
	int stage_fd = -1;
	b2a::StagePlan stage_plan;
	b2a::ScatterPiece *d_stage_pieces = nullptr;
};

namespace b2a {
// one thread-local error slot behind b2a_last_error() (capi.cu)
int SetLastError(int code, const std::string &msg);
// QueryResult::DeduplicateColumns (duckdb/src/main/query_result.cpp:71-93), ipc_reader.cpp
void DeduplicateColumns(std::vector<std::string> &names);
} // namespace b2a
