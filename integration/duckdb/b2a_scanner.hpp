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
// The stream is decoded in slices of whole record batches by one producer thread (b2a_read_arrow_host per slice: PCIe both
// ways + kernels overlap the executor), a few slices ahead of the consumers.  A consumer is a Cursor: one DuckDB scan thread
// holding one record batch of a decoded slice and walking its 2 048-row vectors — the reference gives ONE file one scan
// thread (arrow_file_scan.cpp:35-41); here every thread of the pipeline takes the next record batch of the same file, in
// order (slices are large so that PCIe runs at its rate, the unit handed to a thread is small so that the last one is).
//===----------------------------------------------------------------------===//
#pragma once

#include "b2arrow.h"

#include "duckdb/common/types/data_chunk.hpp"
#include "duckdb/common/types/vector.hpp"
#include "duckdb/common/unordered_map.hpp"

#include <atomic>
#include <condition_variable>
#include <deque>
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

//! Finished b2a_scan handles of one scanner, destroyed together when the last slice and the scanner are gone:
//! b2a_scan_destroy frees device memory, cudaFree waits for the device — in the middle of a scan that would stall the
//! producer's pipeline once per slice
struct B2AScanBin {
	~B2AScanBin();
	void Add(b2a_scan *scan);
	std::mutex mu;
	vector<b2a_scan *> scans;
};

class B2AScanner {
	struct Slice;

public:
	//! One consumer's position: the slice it holds and where in it the next vector starts
	struct Cursor {
		Cursor();
		~Cursor();
		Cursor(Cursor &&) noexcept;
		Cursor &operator=(Cursor &&) noexcept;
		shared_ptr<Slice> slice;
		uint64_t batch = 0, batch_end = 0; // record batches [batch, batch_end) of the slice are this cursor's
		idx_t chunk_offset = 0;
	};

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
	//! column index behind every column of the chunks that will be filled; starts the producer.  Once per scanner.
	void Begin(const vector<idx_t> &columns);
	bool Begun() const {
		return begun;
	}
	//! ArrowScanFunction for a single consumer: at most STANDARD_VECTOR_SIZE rows per call, 0 rows = end of stream
	void Next(DataChunk &chunk);
	//! Several consumers: TakeSlice gives `cursor` the next record batch of the stream (false: none left; blocks while the
	//! producer is still decoding its slice; thread-safe), NextFrom fills `chunk` from the cursor's rows (0 rows: exhausted)
	bool TakeSlice(Cursor &cursor);
	void NextFrom(Cursor &cursor, DataChunk &chunk);
	//! 0..100
	double Progress() const;
	uint64_t GpuLaunches() const {
		return gpu_launches;
	}

private:
	struct Slice {
		b2a_scan *scan = nullptr;
		shared_ptr<B2AScanBin> bin;
		shared_ptr<B2ASlab> slab;
		uint64_t batch_lo = 0, batch_hi = 0;
		std::mutex dict_mu; // a slice is shared by the cursors that walk its record batches
		unordered_map<uint32_t, unique_ptr<Vector>> dictionaries; // flat field -> decoded dictionary values
		uint64_t next_batch = 0; // first record batch not yet handed to a cursor (under B2AScanner::mu)
		int status = 0;
		string error;
		~Slice();
	};

	LogicalType TypeOf(uint32_t flat) const;
	void PointVector(Slice &slice, Vector &vec, uint32_t flat, uint64_t batch, idx_t offset, idx_t count);
	unique_ptr<Slice> Load(uint64_t lo, uint64_t hi);
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
	uint64_t n_batches = 0, total_rows = 0;
	idx_t n_slices = 0;
	bool begun = false;
	std::atomic<uint64_t> rows_done {0};
	std::atomic<uint64_t> gpu_launches {0};
	Cursor own; // the single consumer of Next()
	shared_ptr<B2AScanBin> bin;

	// producers: two threads take turns at the slices (each call of b2a_read_arrow_host leases its own streams / staging
	// ring / arena), so that the pipeline of slice k + 1 fills while the one of slice k drains; slices are handed out in
	// order, at most `depth` started ones ahead of the consumers
	vector<std::thread> workers;
	std::mutex mu;
	std::condition_variable cv;
	bool stop = false;
	idx_t claimed = 0;  // next slice a producer takes
	idx_t produced = 0; // slices [0, produced) are in `ready` or already handed out (n_slices after a failure)
	idx_t handed_out = 0, depth = 3;
	std::deque<shared_ptr<Slice>> ready;                 // finished, in order
	unordered_map<idx_t, shared_ptr<Slice>> out_of_order; // finished before their predecessor
};

} // namespace ext_nanoarrow
} // namespace duckdb
