// Launch interface of the IPC body decompression kernel.  A compressed record batch (RecordBatch.compression, format/
// Message.fbs BodyCompression) stores every buffer as `int64 uncompressed_length` + one codec frame, or
// uncompressed_length = -1 followed by the raw bytes.  The reference decompresses buffer by buffer on the CPU
// (src/ipc/stream_reader/base_stream_reader.cpp:11-50: ZSTD only; LZ4_FRAME is decoded here as well); here every projected buffer becomes one job and one
// launch expands all of them into the scratch region behind the device image, where the decode kernel reads them.
#pragma once
#include <stdint.h>

#define B2A_CODEC_RAW 0u  /* uncompressed_length == -1: plain copy */
#define B2A_CODEC_ZSTD 1u
#define B2A_CODEC_LZ4 2u  /* LZ4_FRAME: not in the reference (it installs ZSTD only), SURVEY §8f N-4 */

/* d_err values: 1..5 = zstd::Error, B2A_DECOMP_LZ4_ERR + 1..5 = lz4::Error */
#define B2A_DECOMP_LZ4_ERR 16u

#define B2A_DECOMP_OK 0u

struct __attribute__((aligned(16))) b2a_decomp_job {
	uint64_t src;     /* image offset of the frame (after the 8-byte length prefix) */
	uint64_t src_len;
	uint64_t dst;     /* image offset inside the scratch region */
	uint64_t dst_len; /* uncompressed length */
	uint32_t codec;
	uint32_t pad[3];
};

namespace b2a {

struct DecompLaunch {
	const b2a_decomp_job *d_jobs = nullptr;
	uint32_t job_begin = 0, job_end = 0;
	const uint32_t *d_order = nullptr; // optional: the k-th job handed out is d_order[k] (largest first); covers all jobs
	uint8_t *image_origin = nullptr; // device address of stream-image offset 0
	void *d_workspaces = nullptr;    // DecompWorkspaceBytes() per worker warp
	uint32_t n_workers = 0;
	uint32_t *d_counter = nullptr;   // zeroed before the launch
	uint32_t *d_err = nullptr;       // decoder error code (atomicMax, see B2A_DECOMP_LZ4_ERR), 0 = ok
	int sm_count = 148;
};

uint64_t DecompWorkspaceBytes();
uint32_t DecompWorkers(int sm_count);
int LaunchDecompress(const DecompLaunch &l, void *stream);

} // namespace b2a
