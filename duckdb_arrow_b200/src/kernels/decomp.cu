// b2a_decompress_kernel: one warp per compressed IPC buffer at a time (jobs are handed out through an atomic counter,
// largest-first is not needed: buffers of one batch are similar).  RAW jobs are a warp-wide copy; ZSTD jobs run the
// frame decoder of zstd_decode.cuh on lane 0 with the warp's workspace (entropy tables + literal buffer) in global memory.
#include "decomp.cuh"

#include "zstd_decode.cuh"

#include <cuda_runtime.h>

namespace b2a {

static constexpr int kWarpsPerCta = 4;

uint64_t DecompWorkspaceBytes() {
	return (sizeof(zstd::Workspace) + 255) & ~uint64_t(255);
}

uint32_t DecompWorkers(int sm_count) {
	return static_cast<uint32_t>(sm_count) * 4 * kWarpsPerCta;
}

__global__ void __launch_bounds__(kWarpsPerCta * 32) b2a_decompress_kernel(DecompLaunch l, uint64_t ws_bytes) {
	const uint32_t lane = threadIdx.x & 31;
	const uint32_t worker = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
	if (worker >= l.n_workers) {
		return;
	}
	zstd::Workspace *w =
	    reinterpret_cast<zstd::Workspace *>(static_cast<uint8_t *>(l.d_workspaces) + static_cast<uint64_t>(worker) * ws_bytes);
	while (true) {
		uint32_t j = 0;
		if (lane == 0) {
			j = l.job_begin + atomicAdd(l.d_counter, 1u);
		}
		j = __shfl_sync(0xFFFFFFFFu, j, 0);
		if (j >= l.job_end) {
			break;
		}
		const b2a_decomp_job job = l.d_jobs[j];
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
		} else if (lane == 0) {
			const int rc = zstd::decode_frame(*w, src, job.src_len, dst, job.dst_len);
			if (rc != zstd::OK) {
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
	b2a_decompress_kernel<<<ctas, kWarpsPerCta * 32, 0, static_cast<cudaStream_t>(stream)>>>(l, DecompWorkspaceBytes());
	return static_cast<int>(cudaGetLastError());
}

} // namespace b2a
