//===----------------------------------------------------------------------===//
// b2a_scanner.hpp — the B200 scan of ONE Arrow IPC stream (file or caller buffers) behind DuckDB's vector model.
//
// Takes the place of what the reference stacks up per stream: IPCStreamReader (src/ipc/stream_reader/*) +
// ArrowIPCStreamFactory / IpcArrayStream (src/ipc/stream_factory.cpp, array_stream.cpp) + ArrowScanLocalState +
// ArrowTableFunction::ArrowScanFunction / ArrowToDuckDB (duckdb/src/function/table/arrow.cpp:172-201,
// arrow_conversion.cpp:1421-1484).  Framing, slicing and every per-value conversion happen inside libb2arrow.so
// (include/b2arrow.h); what is left here is pointing DuckDB vectors at the decoded slabs, 2 048 rows at a time —
// the zero-copy idiom of the reference's DirectConversion (arrow_conversion.cpp:454-461), applied to every type.
//
// The stream is decoded in slices of whole record batches; while DuckDB consumes slice i a worker thread runs
// b2a_read_arrow_host for slice i + 1 (PCIe both ways + kernels overlap the executor).
//===----------------------------------------------------------------------===//
#pragma once

#include "b2arrow.h"

#include "duckdb/common/types/data_chunk.hpp"
#include "duckdb/common/types/vector.hpp"
#include "duckdb/common/unordered_map.hpp"

#include <condition_variable>
#include <mutex>
#include <thread>

namespace duckdb {
namespace ext_nanoarrow {

//! b2a_status -> the DuckDB exception class the reference would have thrown (message = b2a_last_error())
[[noreturn]] void B2AThrow(int status);
//! device used by this process (B2A_DEVICE, default 0)
int B2ADevice();

//! Pinned host arena holding the decoded vectors of one slice; attached to every vector that points into it — the
//! keep-alive the reference uses for Arrow buffers (ArrowAuxiliaryData, arrow_conversion.cpp:862-864).  Arenas go back
//! to a process-wide pool: page-locking memory costs milliseconds per 100 MB.
struct B2ASlab {
	B2ASlab(data_ptr_t data_p, idx_t capacity_p) : data(data_p), capacity(capacity_p) {
	}
	~B2ASlab();
	data_ptr_t data;
	idx_t capacity;
	static shared_ptr<B2ASlab> Acquire(idx_t bytes);
};

struct B2ASlabAuxData : public VectorAuxiliaryData {
	explicit B2ASlabAuxData(shared_ptr<B2ASlab> slab_p)
	    : VectorAuxiliaryData(VectorAuxiliaryDataType::ARROW_AUXILIARY), slab(std::move(slab_p)) {
	}
	shared_ptr<B2ASlab> slab;
};

class B2AScanner {
public:
	//! takes ownership of `reader` (b2a_reader_open_file / b2a_reader_open_buffers)
	B2AScanner(b2a_reader *reader, int device);
	~B2AScanner();

	const vector<string> &Names() const {
		return names;
	}
	const vector<LogicalType> &Types() const {
		return types;
	}
	//! projection pushdown (IPCStreamReader::SetColumnProjection, base_stream_reader.cpp:146-212): `columns` = top-level
	//! column index behind every column of the chunks Next() will fill; starts decoding the first slice
	void Begin(const vector<idx_t> &columns);
	//! ArrowScanFunction: at most STANDARD_VECTOR_SIZE rows per call, 0 rows = end of stream
	void Next(DataChunk &chunk);
	//! 0..100
	double Progress() const;
	uint64_t GpuLaunches() const {
		return gpu_launches;
	}

private:
	struct Slice {
		b2a_scan *scan = nullptr;
		shared_ptr<B2ASlab> slab;
		uint64_t batch_lo = 0, batch_hi = 0;
		unordered_map<uint32_t, unique_ptr<Vector>> dictionaries; // flat field -> decoded dictionary values
		int status = 0;
		string error;
		~Slice();
	};

	LogicalType TypeOf(uint32_t flat) const;
	void PointVector(Slice &slice, Vector &vec, uint32_t flat, uint64_t batch, idx_t offset, idx_t count);
	unique_ptr<Slice> Load(uint64_t lo, uint64_t hi);
	void Request(uint64_t slice);
	unique_ptr<Slice> Take();
	void WorkerLoop();

	int device;
	b2a_reader *reader = nullptr;
	vector<b2a_field_info> fields;
	vector<uint32_t> top_level; // flat index of every top-level column
	vector<string> names;
	vector<LogicalType> types;
	vector<uint32_t> projected_flat;
	const_data_ptr_t image = nullptr;
	vector<uint64_t> slice_bounds; // batch index where slice k starts (+ end)
	uint64_t n_batches = 0, total_rows = 0, rows_done = 0;
	idx_t n_slices = 0, requested = 0, taken = 0;
	unique_ptr<Slice> current;
	uint64_t batch = 0;
	idx_t chunk_offset = 0;
	uint64_t gpu_launches = 0;

	// prefetch worker (one thread per scanner: the C ABI keeps its streams / staging per thread and device)
	std::thread worker;
	std::mutex mu;
	std::condition_variable cv;
	bool stop = false, has_request = false, has_result = false;
	uint64_t req_lo = 0, req_hi = 0;
	unique_ptr<Slice> result;
};

} // namespace ext_nanoarrow
} // namespace duckdb
