// Scan planner (host): record-batch range x projected fields -> conversion tasks, tile table and the
// layout of the output arena.  This is the part of ArrowScanFunction / ArrowToDuckDB
// (duckdb/src/function/table/arrow.cpp:172-201, arrow_conversion.cpp:1421-1484) that decides WHICH
// conversion runs on WHICH buffer; the conversions themselves are the kernels in src/kernels.
#pragma once
#include "../ipc/ipc_reader.hpp"
#include "../kernels/decomp.cuh"
#include "../kernels/task.h"

#include <map>
#include <vector>

namespace b2a {

struct ScanPlan {
	uint64_t batch_lo = 0, batch_hi = 0;
	uint32_t n_flat = 0;
	// slice of the stream image the tasks read: [image_lo, image_hi), image_lo is 256-byte aligned
	uint64_t image_lo = 0, image_hi = 0;
	uint64_t out_bytes = 0;
	uint64_t stats_base = 0, stats_bytes = 0; // region of the arena that must be zeroed before a decode
	uint64_t algorithmic_in = 0, algorithmic_out = 0;
	uint64_t total_rows = 0;
	uint64_t n_chunks = 0; // sum over batches of ceil(rows/2048)

	std::vector<b2a_task> tasks;
	std::vector<uint32_t> tile_task;
	std::vector<uint64_t> variadic;
	std::vector<b2a_column_out> cols;            // [(batch - batch_lo) * n_flat + flat]
	std::map<int, b2a_column_out> dict_cols;     // flat field -> decoded dictionary values
	std::vector<int> top_columns;                // projected top-level columns (indices into schema.fields)
	std::vector<uint64_t> batch_rows;            // rows of each batch in range
	std::vector<uint64_t> batch_chunk0;          // first global chunk index of each batch
	// per batch (+1 sentinel): first tile, first output byte, and the image range its tasks read
	std::vector<uint32_t> batch_tile0;
	std::vector<uint64_t> batch_out0;
	std::vector<uint64_t> batch_image_lo, batch_image_hi;

	// host-alias mode (B2A_SCAN_HOST_ALIAS): DirectConversion columns alias the host image, no task is emitted
	bool alias_direct = false;
	bool alias_on_device = false; // B2A_SCAN_DEVICE_ALIAS: the aliased buffers are read on the device, so they are uploaded
	// image ranges the tasks of each batch read (merged when close), + the ranges of the dictionary batches;
	// the host path uploads exactly these
	typedef std::vector<std::pair<uint64_t, uint64_t>> RangeList;
	std::vector<RangeList> batch_ranges;
	RangeList tail_ranges;
	uint64_t upload_bytes = 0;

	// ---- compressed bodies (RecordBatch.compression): every projected buffer of a compressed batch is one job of the
	// decompress launch; its output lives in the scratch region [scratch_base, scratch_base + scratch_bytes) of the
	// image offset space (directly behind the uploaded slice), and the tasks read it from there
	std::vector<b2a_decomp_job> jobs;
	std::vector<uint32_t> batch_job0; // first job of each batch (+1 sentinel); dictionary jobs follow
	uint64_t scratch_base = 0, scratch_bytes = 0;
	bool compressed = false;
	// tiles [phase2_tile0, n_tiles) belong to tasks that read vectors decoded by the tiles before them (REE)
	uint32_t phase2_tile0 = 0;
	// no task outside {COPY, VALIDITY, BOOL, STRING32, STRING64}: the lean decode kernel can run the scan
	bool lean = false;

	void Build(const Reader &reader, uint64_t lo, uint64_t hi);

	const b2a_column_out &Column(uint64_t batch, uint32_t flat) const {
		return cols[(batch - batch_lo) * n_flat + flat];
	}

private:
	struct Chain {
		b2a_mask_ref m[B2A_MAX_MASKS];
		uint32_t n = 0;
		// list rebase chain (see chunk_base_row in kernels/decode.cu): ancestors between the top-level column and
		// this field that change the row indexing
		std::vector<std::pair<uint64_t, uint64_t>> levels; // (kind, value)
		uint64_t top_rows = 0;
	};
	uint64_t Alloc(uint64_t bytes);
	void Touch(uint64_t off, uint64_t len);
	void AddTask(b2a_task t, uint64_t in_bytes_per_row_x16, uint64_t out_bytes_per_row);
	// returns number of nodes consumed
	void EmitField(const Reader &reader, const Field &f, const RecordBatch &rb, size_t &node_idx, size_t &buf_idx,
	               size_t &var_idx, const Chain &parent_chain, bool list_child, bool selected, b2a_column_out *slots,
	               bool as_dictionary_values, bool ignore_dictionary);
	void SkipField(const Field &f, const RecordBatch &rb, size_t &node_idx, size_t &buf_idx, size_t &var_idx,
	               bool ignore_dictionary);
	bool big_endian = false;
	uint32_t next_tile = 0;
	RangeList *cur_ranges = nullptr;
	// candidate jobs: one per non-empty buffer of a compressed batch; becomes a job when a task first touches it
	struct DecompCand {
		uint64_t src_off, src_len, dst_len;
		uint32_t codec;
		int64_t job;
	};
	std::vector<DecompCand> cands;
	std::vector<b2a_task> phase2_pending;
	RecordBatch VirtualBatch(const Reader &reader, const RecordBatch &rb);
	uint64_t Resolve(uint64_t off) const;
};

} // namespace b2a
