// Host staging of the upload ranges: the byte ranges a slice of record batches needs are packed by a few
// host threads from the stream image (an mmap'ed file, or any pageable buffer) into a ring of PINNED buffers, each slice
// then crosses PCIe as ONE large copy and is scattered to its place in the device image by b2a_scatter_kernel.
//
// Replaces the reference's 4 KiB BufferedFileReader loop (src/ipc/stream_reader/ipc_file_stream_reader.cpp:80-138 copies
// every message body through one small buffer on one thread) for images cudaMemcpyAsync cannot read asynchronously.
// This is I/O staging, not conversion: no byte is interpreted on the host.
#pragma once
#include "scanner/scan_plan.hpp"

#include <atomic>
#include <cstdint>
#include <thread>
#include <vector>

namespace b2a {

struct StagePiece {
	uint64_t src;   // offset inside the host image
	uint64_t stage; // offset inside the slice's staging buffer (congruent to `dst` modulo 16)
	uint64_t dst;   // offset inside the device image
	uint64_t bytes;
};

struct StagePlan {
	static constexpr uint64_t kPieceBytes = 1u << 20;
	std::vector<StagePiece> pieces;     // slice by slice
	std::vector<uint32_t> slice_piece0; // n_slices + 1
	std::vector<uint64_t> slice_bytes;  // staging bytes of each slice
	uint64_t max_slice_bytes = 0;
	uint32_t n_slices = 0;

	// align = 16: pieces keep their phase inside a 16-byte vector (all the scatter kernel needs).  align = 4096
	// (O_DIRECT reads): a piece's staging offset is congruent to its FILE offset modulo 4096 and every range is padded
	// to whole blocks, so that the enclosing aligned block range can be read straight into the pinned buffer.
	void Build(const ScanPlan &plan, uint32_t n_slices, uint64_t align = 16);
	uint64_t align = 16;
};

// Worker threads that pack pieces in order; the caller gates them slice by slice (a staging buffer may be refilled only
// after its previous content has crossed the bus) and waits for a slice to be complete before sending it.
class StageWorkers {
public:
	// io: 0 = memcpy from the (mmap'ed / caller) image, 1 = pread(fd) into the pinned ring, 2 = pread(fd opened with
	// O_DIRECT) of the enclosing 4 KiB blocks (plan built with align = 4096)
	StageWorkers(const StagePlan &plan, const uint8_t *image, uint8_t *const *ring, uint32_t ring_size, uint32_t n_threads,
	             int io = 0, int fd = -1, uint64_t file_size = 0);
	bool Failed() const {
		return failed.load(std::memory_order_acquire);
	}
	~StageWorkers();
	void Allow(uint32_t slice);      // slices <= `slice` may be packed
	void WaitPacked(uint32_t slice); // blocks until every piece of `slice` is in its staging buffer

private:
	void Run();
	const StagePlan &plan;
	const uint8_t *image;
	uint8_t *const *ring;
	uint32_t ring_size;
	std::atomic<uint32_t> next_piece {0};
	std::atomic<int64_t> allowed {-1};
	std::vector<std::atomic<uint32_t>> remaining;
	std::vector<std::thread> threads;
	int io = 0, fd = -1;
	uint64_t file_size = 0;
	std::atomic<bool> failed {false};
};

} // namespace b2a
