#include "scatter.cuh"

#include <cuda_runtime.h>

namespace b2a {

static constexpr uint32_t kScatterThreads = 256;

__global__ void __launch_bounds__(kScatterThreads)
    b2a_scatter_kernel(const ScatterPiece *__restrict__ pieces, uint32_t n_pieces, const uint8_t *__restrict__ stage,
                       uint8_t *__restrict__ image) {
	for (uint32_t q = blockIdx.x; q < n_pieces; q += gridDim.x) {
		const ScatterPiece pc = pieces[q];
		const uint8_t *src = stage + pc.stage;
		uint8_t *dst = image + pc.dst;
		uint64_t n = pc.bytes;
		// src and dst share their phase inside a 16-byte vector (StagePlan::Build)
		const uint64_t head = min(n, static_cast<uint64_t>((16u - (reinterpret_cast<uintptr_t>(dst) & 15u)) & 15u));
		if (threadIdx.x < head) {
			dst[threadIdx.x] = src[threadIdx.x];
		}
		src += head;
		dst += head;
		n -= head;
		const uint64_t vecs = n >> 4;
		const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
		uint4 *d4 = reinterpret_cast<uint4 *>(dst);
		constexpr uint32_t kU = 4;
		uint64_t i = threadIdx.x;
		for (; i + (kU - 1) * kScatterThreads < vecs; i += kU * kScatterThreads) {
			uint4 v[kU];
#pragma unroll
			for (uint32_t k = 0; k < kU; k++) {
				v[k] = __ldcs(s4 + i + k * kScatterThreads);
			}
#pragma unroll
			for (uint32_t k = 0; k < kU; k++) {
				d4[i + k * kScatterThreads] = v[k];
			}
		}
		for (; i < vecs; i += kScatterThreads) {
			d4[i] = __ldcs(s4 + i);
		}
		const uint64_t tail = n - (vecs << 4);
		if (threadIdx.x < tail) {
			dst[(vecs << 4) + threadIdx.x] = src[(vecs << 4) + threadIdx.x];
		}
	}
}

int LaunchScatter(const ScatterPiece *d_pieces, uint32_t n_pieces, const uint8_t *d_stage, uint8_t *d_image, int sm_count,
                  void *stream) {
	if (n_pieces == 0) {
		return 0;
	}
	uint32_t grid = static_cast<uint32_t>(sm_count) * 4;
	if (grid > n_pieces) {
		grid = n_pieces;
	}
	b2a_scatter_kernel<<<grid, kScatterThreads, 0, static_cast<cudaStream_t>(stream)>>>(d_pieces, n_pieces, d_stage, d_image);
	return static_cast<int>(cudaGetLastError());
}

} // namespace b2a
