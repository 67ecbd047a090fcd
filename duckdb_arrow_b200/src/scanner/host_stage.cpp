#include "scanner/host_stage.hpp"

#include <cstring>

namespace b2a {

void StagePlan::Build(const ScanPlan &plan, uint32_t n_slices_p) {
	n_slices = n_slices_p;
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
				const uint64_t dst0 = r.first - plan.image_lo;
				cursor = ((cursor + 15) & ~uint64_t(15)) + (dst0 & 15);
				for (uint64_t o = r.first; o < r.second; o += kPieceBytes) {
					const uint64_t n = std::min<uint64_t>(kPieceBytes, r.second - o);
					pieces.push_back(StagePiece {o, cursor, o - plan.image_lo, n});
					cursor += n;
				}
			}
		}
		slice_piece0.push_back(static_cast<uint32_t>(pieces.size()));
		slice_bytes.push_back(cursor);
		max_slice_bytes = std::max(max_slice_bytes, cursor);
	}
}

StageWorkers::StageWorkers(const StagePlan &plan_p, const uint8_t *image_p, uint8_t *const *ring_p, uint32_t ring_size_p,
                           uint32_t n_threads)
    : plan(plan_p), image(image_p), ring(ring_p), ring_size(ring_size_p), remaining(plan_p.n_slices) {
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
		std::memcpy(ring[slice % ring_size] + pc.stage, image + pc.src, pc.bytes);
		remaining[slice].fetch_sub(1, std::memory_order_release);
	}
}

} // namespace b2a
