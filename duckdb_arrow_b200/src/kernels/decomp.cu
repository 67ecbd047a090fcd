// b2a_decompress_kernel: one warp per compressed IPC buffer at a time (jobs are handed out through an atomic counter,
// largest-first is not needed: buffers of one batch are similar).  RAW jobs are a warp-wide copy; ZSTD jobs run the
// frame decoder of zstd_decode.cuh warp-wide (see its header) with the warp's workspace (entropy tables + literal buffer) in global memory;
// LZ4 jobs run lz4_decode.cuh the same way (no workspace).
#include "decomp.cuh"

#include "lz4_decode.cuh"
#include "zstd_decode.cuh"

#include <cuda_runtime.h>

namespace b2a {

static constexpr int kWarpsPerCta = 4;
// per warp in shared memory: the entropy tables
static constexpr uint32_t kSmemPerWarp = (sizeof(zstd::Tables) + 15) & ~15u;
static constexpr uint32_t kSmemPerCta = kSmemPerWarp * kWarpsPerCta;
static_assert(kSmemPerCta <= 227 * 1024, "decompress kernel: shared memory per CTA");

uint64_t DecompWorkspaceBytes() {
	return (sizeof(zstd::Workspace) + 255) & ~uint64_t(255);
}

uint32_t DecompWorkers(int sm_count) {
	const uint32_t ctas_per_sm = (227u * 1024u) / (kSmemPerCta + 1024u);
	return static_cast<uint32_t>(sm_count) * (ctas_per_sm ? ctas_per_sm : 1) * kWarpsPerCta;
}

__global__ void __launch_bounds__(kWarpsPerCta * 32) b2a_decompress_kernel(DecompLaunch l, uint64_t ws_bytes) {
	const uint32_t lane = threadIdx.x & 31;
	const uint32_t worker = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
	if (worker >= l.n_workers) {
		return;
	}
	zstd::Workspace *w =
	    reinterpret_cast<zstd::Workspace *>(static_cast<uint8_t *>(l.d_workspaces) + static_cast<uint64_t>(worker) * ws_bytes);
	extern __shared__ __align__(16) uint8_t smem[];
	uint8_t *my = smem + (threadIdx.x >> 5) * kSmemPerWarp;
	const zstd::Ctx cx = {static_cast<int>(lane), 32, reinterpret_cast<zstd::Tables *>(my)};
	while (true) {
		uint32_t j = 0;
		if (lane == 0) {
			j = l.job_begin + atomicAdd(l.d_counter, 1u);
		}
		j = __shfl_sync(0xFFFFFFFFu, j, 0);
		if (j >= l.job_end) {
			break;
		}
		const b2a_decomp_job job = l.d_jobs[l.d_order ? l.d_order[j] : j];
		const uint8_t *src = l.image_origin + job.src;
		uint8_t *dst = l.image_origin + job.dst;
		if (job.codec == B2A_CODEC_RAW) {
			// both sides are 8-byte aligned (IPC buffers are, scratch slots are 64-byte aligned)
			const uint64_t n8 = job.dst_len >> 3;
			const uint64_t *s8 = reinterpret_cast<const uint64_t *>(src);
			uint64_t *d8 = reinterpret_cast<uint64_t *>(dst);
			for (uint64_t i = lane; i < n8; i += 32) {
				d8[i] = s8[i];
			}
			for (uint64_t i = (n8 << 3) + lane; i < job.dst_len; i += 32) {
				dst[i] = src[i];
			}
		} else if (job.codec == B2A_CODEC_LZ4) {
			const lz4::Ctx lx = {static_cast<int>(lane), 32};
			const int rc = lz4::decode_frame(lx, src, job.src_len, dst, job.dst_len);
			if (rc != lz4::OK && lane == 0) {
				atomicMax(l.d_err, B2A_DECOMP_LZ4_ERR + static_cast<uint32_t>(rc));
			}
		} else {
			// every lane runs the decoder (uniform control flow, lane-parallel copies and Huffman streams)
			const int rc = zstd::decode_frame(cx, *w, src, job.src_len, dst, job.dst_len);
			if (rc != zstd::OK && lane == 0) {
				atomicMax(l.d_err, static_cast<uint32_t>(rc));
			}
		}
		__syncwarp();
	}
}

int LaunchDecompress(const DecompLaunch &l, void *stream) {
	if (l.job_end <= l.job_begin) {
		return 0;
	}
	const uint32_t ctas = (l.n_workers + kWarpsPerCta - 1) / kWarpsPerCta;
	// the attribute is per device: a process may launch on several GPUs
	static bool attr_set[64] = {};
	int dev = 0;
	if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
		dev = 0;
	}
	if (!attr_set[dev]) {
		cudaError_t e = cudaFuncSetAttribute(b2a_decompress_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
		                                     static_cast<int>(kSmemPerCta));
		if (e != cudaSuccess) {
			return static_cast<int>(e);
		}
		attr_set[dev] = true;
	}
	b2a_decompress_kernel<<<ctas, kWarpsPerCta * 32, kSmemPerCta, static_cast<cudaStream_t>(stream)>>>(
	    l, DecompWorkspaceBytes());
	return static_cast<int>(cudaGetLastError());
}

} // namespace b2a
