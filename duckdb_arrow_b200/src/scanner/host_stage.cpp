#include "scanner/host_stage.hpp"

#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <unistd.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

namespace b2a {

// Copy into the pinned ring with non-temporal stores: the bytes are read next by the copy engine, not by a core, and a
// plain store to a line that is not in the cache first READS it (memcpy of pieces <= 1 MiB stays below glibc's own
// non-temporal threshold): a third of the staging threads' memory traffic.  Source and destination of a piece share their
// phase inside a 16-byte vector (StagePlan), so both become aligned together.
static void StreamCopy(uint8_t *dst, const uint8_t *src, uint64_t n) {
#if defined(__SSE2__)
	static const bool plain = std::getenv("B2A_STAGE_PLAIN_MEMCPY") != nullptr;
	if (!plain && n >= 256 && ((reinterpret_cast<uintptr_t>(dst) ^ reinterpret_cast<uintptr_t>(src)) & 15) == 0) {
		const uint64_t head = (16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15;
		std::memcpy(dst, src, head);
		dst += head;
		src += head;
		n -= head;
		const uint64_t nv = n >> 4;
		const __m128i *s = reinterpret_cast<const __m128i *>(src);
		__m128i *d = reinterpret_cast<__m128i *>(dst);
		uint64_t i = 0;
		for (; i + 4 <= nv; i += 4) {
			const __m128i a = _mm_load_si128(s + i), b = _mm_load_si128(s + i + 1), c = _mm_load_si128(s + i + 2),
			              e = _mm_load_si128(s + i + 3);
			_mm_stream_si128(d + i, a);
			_mm_stream_si128(d + i + 1, b);
			_mm_stream_si128(d + i + 2, c);
			_mm_stream_si128(d + i + 3, e);
		}
		for (; i < nv; i++) {
			_mm_stream_si128(d + i, _mm_load_si128(s + i));
		}
		_mm_sfence();
		std::memcpy(dst + (nv << 4), src + (nv << 4), n & 15);
		return;
	}
#endif
	std::memcpy(dst, src, n);
}


void StagePlan::Build(const ScanPlan &plan, uint32_t n_slices_p, uint64_t align_p) {
	n_slices = n_slices_p;
	align = align_p;
	pieces.clear();
	slice_piece0.assign(1, 0);
	slice_bytes.clear();
	max_slice_bytes = 0;
	const uint64_t n_batches = plan.batch_hi - plan.batch_lo;
	for (uint32_t sl = 0; sl < n_slices; sl++) {
		const uint64_t b0 = n_batches * sl / n_slices, b1 = n_batches * (sl + 1) / n_slices;
		uint64_t cursor = 0;
		for (uint64_t b = b0; b < b1; b++) {
			for (auto &r : plan.batch_ranges[b]) {
				if (r.second <= r.first) {
					continue;
				}
				// keep the range's phase inside a 16-byte vector so that the device-side scatter copies aligned vectors
				// (image_lo is a multiple of 256, so a staging offset congruent to the FILE offset modulo `align` is also
				// congruent to the device offset modulo 16)
				cursor = ((cursor + align - 1) & ~(align - 1)) + (r.first & (align - 1));
				for (uint64_t o = r.first; o < r.second; o += kPieceBytes) {
					const uint64_t n = std::min<uint64_t>(kPieceBytes, r.second - o);
					pieces.push_back(StagePiece {o, cursor, o - plan.image_lo, n});
					cursor += n;
				}
				if (align > 16) { // block reads overshoot the range's end: keep the next range out of that block
					cursor = (cursor + align - 1) & ~(align - 1);
				}
			}
		}
		slice_piece0.push_back(static_cast<uint32_t>(pieces.size()));
		slice_bytes.push_back(cursor);
		max_slice_bytes = std::max(max_slice_bytes, cursor);
	}
}

StageWorkers::StageWorkers(const StagePlan &plan_p, const uint8_t *image_p, uint8_t *const *ring_p, uint32_t ring_size_p,
                           uint32_t n_threads, int io_p, int fd_p, uint64_t file_size_p)
    : plan(plan_p), image(image_p), ring(ring_p), ring_size(ring_size_p), remaining(plan_p.n_slices), io(io_p), fd(fd_p),
      file_size(file_size_p) {
	for (uint32_t sl = 0; sl < plan.n_slices; sl++) {
		remaining[sl].store(plan.slice_piece0[sl + 1] - plan.slice_piece0[sl], std::memory_order_relaxed);
	}
	for (uint32_t t = 0; t < n_threads; t++) {
		threads.emplace_back([this] { Run(); });
	}
}

StageWorkers::~StageWorkers() {
	allowed.store(static_cast<int64_t>(plan.n_slices), std::memory_order_release); // let stragglers run to the end
	for (auto &t : threads) {
		t.join();
	}
}

void StageWorkers::Allow(uint32_t slice) {
	int64_t cur = allowed.load(std::memory_order_relaxed);
	while (cur < static_cast<int64_t>(slice) &&
	       !allowed.compare_exchange_weak(cur, static_cast<int64_t>(slice), std::memory_order_release)) {
	}
}

void StageWorkers::WaitPacked(uint32_t slice) {
	while (remaining[slice].load(std::memory_order_acquire) != 0) {
		std::this_thread::yield();
	}
}

void StageWorkers::Run() {
	const uint32_t total = static_cast<uint32_t>(plan.pieces.size());
	uint32_t slice = 0;
	while (true) {
		const uint32_t q = next_piece.fetch_add(1, std::memory_order_relaxed);
		if (q >= total) {
			return;
		}
		while (plan.slice_piece0[slice + 1] <= q) {
			slice++;
		}
		while (allowed.load(std::memory_order_acquire) < static_cast<int64_t>(slice)) {
			std::this_thread::yield();
		}
		const StagePiece &pc = plan.pieces[q];
		uint8_t *dst = ring[slice % ring_size] + pc.stage;
		if (io == 0 || fd < 0) {
			StreamCopy(dst, image + pc.src, pc.bytes);
		} else {
			// the reference reads through BufferedFileReader, 4 KiB per read() (ipc_file_stream_reader.cpp:80-138); here
			// every piece (<= 1 MiB) is one pread straight into pinned memory, many in flight
			uint64_t off = pc.src, left = pc.bytes;
			if (io == 2) { // O_DIRECT: whole blocks around the piece (the plan keeps dst - (off % 4096) block aligned)
				const uint64_t a0 = off & ~uint64_t(4095);
				dst -= off - a0;
				left = ((pc.src + pc.bytes + 4095) & ~uint64_t(4095)) - a0;
				off = a0;
			}
			while (left > 0) {
				const ssize_t got = pread(fd, dst, left, static_cast<off_t>(off));
				if (got < 0 && errno == EINTR) {
					continue;
				}
				if (got <= 0) {
					if (io == 2 && got == 0 && off >= file_size) {
						break; // the last block of the file is short
					}
					failed.store(true, std::memory_order_release);
					break;
				}
				dst += got;
				off += static_cast<uint64_t>(got);
				left -= std::min<uint64_t>(left, static_cast<uint64_t>(got));
			}
		}
		remaining[slice].fetch_sub(1, std::memory_order_release);
	}
}

} // namespace b2a
