// C ABI (include/b2arrow.h): exception firewall + device memory plumbing around the planner and kernels.
// There is NO CPU fallback: every b2a_scan_* entry point that needs the device fails with B2A_ECUDA
// when no CUDA device is usable.
#include "capi_internal.hpp"
#include "kernels/decode.cuh"
#include "kernels/decomp.cuh"

#include <cuda_runtime.h>

#include <mutex>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <thread>

using namespace b2a;

static thread_local std::string g_last_error;

static int Fail(int code, const std::string &msg) {
	g_last_error = msg;
	return code;
}

namespace b2a {
// shared with capi_writer.cu: one thread-local error slot behind b2a_last_error()
int SetLastError(int code, const std::string &msg) {
	return Fail(code, msg);
}
} // namespace b2a

#define B2A_TRY try {
#define B2A_CATCH                                                                                                      \
	}                                                                                                                  \
	catch (const Error &e) {                                                                                           \
		return Fail(e.code, e.msg);                                                                                    \
	}                                                                                                                  \
	catch (const std::bad_alloc &) {                                                                                   \
		return Fail(B2A_EOOM, "out of host memory");                                                                   \
	}                                                                                                                  \
	catch (const std::exception &e) {                                                                                  \
		return Fail(B2A_EINTERNAL, e.what());                                                                          \
	}

#define CUDA_OK(expr)                                                                                                  \
	do {                                                                                                               \
		cudaError_t _e = (expr);                                                                                       \
		if (_e != cudaSuccess) {                                                                                       \
			return Fail(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " #expr);              \
		}                                                                                                              \
	} while (0)

extern "C" {

const char *b2a_version(void) {
	return "b2arrow 0.1.0 (sm_100a)";
}

const char *b2a_last_error(void) {
	return g_last_error.c_str();
}

int b2a_device_count(int *n) {
	int c = 0;
	cudaError_t e = cudaGetDeviceCount(&c);
	if (e != cudaSuccess) {
		*n = 0;
		cudaGetLastError();
		return Fail(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(e));
	}
	*n = c;
	return B2A_OK;
}

// ------------------------------------------------------------------ reader
int b2a_reader_open_file(const char *path, b2a_reader **out) {
	B2A_TRY
	std::unique_ptr<b2a_reader> r(new b2a_reader());
	r->reader.OpenFile(path);
	*out = r.release();
	return B2A_OK;
	B2A_CATCH
}

int b2a_reader_open_buffers(const b2a_ipc_buffer *bufs, uint32_t n, b2a_reader **out) {
	B2A_TRY
	std::unique_ptr<b2a_reader> r(new b2a_reader());
	r->reader.OpenBuffers(bufs, n);
	*out = r.release();
	return B2A_OK;
	B2A_CATCH
}

void b2a_reader_close(b2a_reader *r) {
	delete r;
}

int b2a_reader_num_fields(b2a_reader *r, uint32_t *n_top, uint32_t *n_flat) {
	B2A_TRY
	*n_top = static_cast<uint32_t>(r->reader.OutputColumns().size());
	*n_flat = static_cast<uint32_t>(r->reader.BaseSchema().flat.size());
	return B2A_OK;
	B2A_CATCH
}

static void CopyStr(char *dst, size_t cap, const std::string &s) {
	size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
	std::memcpy(dst, s.data(), n);
	dst[n] = 0;
}

int b2a_reader_field(b2a_reader *r, uint32_t flat_index, b2a_field_info *out) {
	B2A_TRY
	const Schema &s = r->reader.BaseSchema();
	if (flat_index >= s.flat.size()) {
		return Fail(B2A_EINVAL, "field index out of range");
	}
	const Field &f = *s.flat[flat_index];
	std::memset(out, 0, sizeof(*out));
	CopyStr(out->name, sizeof(out->name), f.name);
	CopyStr(out->format, sizeof(out->format), f.format);
	out->type_id = f.type_id;
	out->op = f.op;
	out->out_width = f.out_width;
	out->src_width = f.src_width;
	out->n_children = static_cast<int32_t>(f.children.size());
	out->first_child = f.children.empty() ? -1 : f.children[0]->flat_index;
	out->parent = f.parent_flat;
	out->next_sibling = -1;
	const std::vector<std::unique_ptr<Field>> *sibs = nullptr;
	if (f.parent_flat >= 0) {
		sibs = &s.flat[static_cast<size_t>(f.parent_flat)]->children;
	} else {
		sibs = &s.fields;
	}
	for (size_t i = 0; i + 1 < sibs->size(); i++) {
		if ((*sibs)[i].get() == &f) {
			out->next_sibling = (*sibs)[i + 1]->flat_index;
		}
	}
	out->nullable = f.nullable;
	out->n_buffers = f.n_buffers;
	out->decimal_width = f.decimal_width;
	out->decimal_scale = f.decimal_scale;
	out->param = f.param;
	out->dict_id = f.dict_id;
	out->index_type_id = f.index_type_id;
	out->index_width = f.index_width;
	return B2A_OK;
	B2A_CATCH
}

int b2a_reader_set_projection(b2a_reader *r, const char *const *names, uint32_t n) {
	B2A_TRY
	std::vector<std::string> v;
	for (uint32_t i = 0; i < n; i++) {
		v.emplace_back(names[i]);
	}
	r->reader.SetProjection(v);
	return B2A_OK;
	B2A_CATCH
}

int b2a_reader_num_batches(b2a_reader *r, uint64_t *n_batches, uint64_t *n_rows) {
	*n_batches = r->reader.Batches().size();
	*n_rows = r->reader.TotalRows();
	return B2A_OK;
}

int b2a_reader_batch_info(b2a_reader *r, uint64_t batch, b2a_batch_info *out) {
	B2A_TRY
	const auto &bs = r->reader.Batches();
	if (batch >= bs.size()) {
		return Fail(B2A_EINVAL, "batch index out of range");
	}
	const RecordBatch &b = bs[batch];
	std::memset(out, 0, sizeof(*out));
	out->rows = b.rows;
	out->body_offset = b.body_offset;
	out->body_length = b.body_length;
	out->meta_offset = b.meta_offset;
	out->compressed = b.compression >= 0;
	uint64_t sum = 0;
	for (auto &s : b.buffers) {
		sum += s.length;
	}
	out->buffer_bytes = sum;
	return B2A_OK;
	B2A_CATCH
}

int b2a_reader_image(b2a_reader *r, const void **data, uint64_t *size) {
	*data = r->reader.Image();
	*size = r->reader.ImageSize();
	return B2A_OK;
}

int b2a_reader_used_footer(b2a_reader *r, int *used) {
	if (!r || !used) {
		return Fail(B2A_EINVAL, "null argument");
	}
	*used = r->reader.UsedFooter() ? 1 : 0;
	return B2A_OK;
}

int b2a_reader_schema_metadata(b2a_reader *r, uint32_t i, const char **key, const char **value, uint32_t *n) {
	const auto &m = r->reader.BaseSchema().metadata;
	*n = static_cast<uint32_t>(m.size());
	if (i < m.size()) {
		*key = m[i].first.c_str();
		*value = m[i].second.c_str();
	}
	return B2A_OK;
}

// ------------------------------------------------------------------ scan
// Device memory owned by a scan comes from the device's stream-ordered pool (release threshold raised once per device) on
// a private stream that is synchronised right after the allocation, so the memory is usable from any stream when the
// helper returns.  A read_arrow over a file is many scans (the DuckDB extension decodes slice by slice): cudaMalloc /
// cudaFree per table cost more than a slice's PCIe time (6 ms of 12 per 256 MiB slice, B2A_SCANNER_TRACE), and every
// cudaFree waited for the device.  cudaGetDeviceProperties, 1-2 ms per call, went the same way (SmCount).
static std::mutex g_dev_mu;
static cudaStream_t g_alloc_stream[64] = {};
static int g_sm_count[64] = {};

static cudaError_t AllocStream(cudaStream_t *out) {
	int dev = 0;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) {
		return e;
	}
	if (dev < 0 || dev >= 64) {
		return cudaErrorInvalidDevice;
	}
	std::lock_guard<std::mutex> lock(g_dev_mu);
	if (!g_alloc_stream[dev]) {
		cudaMemPool_t pool;
		if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
			uint64_t threshold = UINT64_MAX;
			cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
		}
		e = cudaStreamCreateWithFlags(&g_alloc_stream[dev], cudaStreamNonBlocking);
		if (e != cudaSuccess) {
			g_alloc_stream[dev] = nullptr;
			return e;
		}
	}
	*out = g_alloc_stream[dev];
	return cudaSuccess;
}

#define ScanAlloc(pp, bytes) ScanAllocRaw(reinterpret_cast<void **>(pp), bytes)
static cudaError_t ScanAllocRaw(void **p, size_t bytes) {
	cudaStream_t st;
	cudaError_t e = AllocStream(&st);
	if (e != cudaSuccess) {
		return e;
	}
	void *q = nullptr;
	e = cudaMallocAsync(&q, bytes ? bytes : 16, st);
	if (e != cudaSuccess) {
		return e;
	}
	*p = q;
	return cudaStreamSynchronize(st);
}

static void ScanFree(void *p) {
	cudaStream_t st;
	if (p && AllocStream(&st) == cudaSuccess) {
		cudaFreeAsync(p, st);
	}
}

static cudaError_t SmCount(int device, int *out) {
	if (device < 0 || device >= 64) {
		return cudaErrorInvalidDevice;
	}
	{
		std::lock_guard<std::mutex> lock(g_dev_mu);
		if (g_sm_count[device]) {
			*out = g_sm_count[device];
			return cudaSuccess;
		}
	}
	int n = 0;
	cudaError_t e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
	if (e == cudaSuccess) {
		std::lock_guard<std::mutex> lock(g_dev_mu);
		g_sm_count[device] = n;
		*out = n;
	}
	return e;
}

static void FreeScanDevice(b2a_scan *s) {
	if (s->device >= 0) {
		cudaSetDevice(s->device);
		if (s->launches || s->d_image_owned) {
			cudaDeviceSynchronize(); // what the cudaFree of each buffer used to imply: nothing of this scan is in flight
		}
		ScanFree(s->d_tasks);
		ScanFree(s->d_tile_task);
		ScanFree(s->d_variadic);
		ScanFree(s->d_err);
		ScanFree(s->d_image_owned);
		ScanFree(s->d_chunk_col);
		ScanFree(s->d_chunk_col2);
		ScanFree(s->d_chunk_col3);
		ScanFree(s->d_chunk_rows);
		ScanFree(s->d_jobs);
		ScanFree(s->d_job_order);
		ScanFree(s->d_stage_pieces);
		ScanFree(s->d_workspaces);
		ScanFree(s->d_decomp_state);
		(void)cudaGetLastError();
	}
}

// total device bytes behind image offset image_lo: the uploaded slice (+ the scratch region of compressed scans)
static uint64_t DeviceImageBytes(const ScanPlan &p) {
	return p.compressed ? (p.scratch_base - p.image_lo) + p.scratch_bytes : (p.image_hi - p.image_lo);
}

// expands the compressed buffers of jobs [j0, j1) into the scratch region (no-op for uncompressed scans)
static int LaunchJobs(b2a_scan *s, uint32_t j0, uint32_t j1, cudaStream_t st) {
	if (j1 <= j0) {
		return B2A_OK;
	}
	cudaError_t ce = cudaMemsetAsync(s->d_decomp_state, 0, sizeof(uint32_t), st);
	if (ce != cudaSuccess) {
		g_last_error = std::string("CUDA error: ") + cudaGetErrorString(ce);
		return B2A_ECUDA;
	}
	DecompLaunch l;
	l.d_jobs = s->d_jobs;
	l.job_begin = j0;
	l.job_end = j1;
	l.d_order = (j0 == 0 && j1 == s->plan.jobs.size()) ? s->d_job_order : nullptr;
	l.image_origin = const_cast<uint8_t *>(s->d_image) - s->plan.image_lo;
	l.d_workspaces = s->d_workspaces;
	l.n_workers = s->n_workers;
	l.d_counter = s->d_decomp_state;
	l.d_err = s->d_decomp_state + 1;
	l.sm_count = s->sm_count;
	int e = LaunchDecompress(l, st);
	if (e != 0) {
		g_last_error = std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)) +
		               " launching b2a_decompress_kernel";
		return B2A_ECUDA;
	}
	s->launches++;
	return B2A_OK;
}

// device < 0: plan only (layout inspection on a machine without a GPU); every compute call then fails.
int b2a_scan_create(b2a_reader *r, int device, uint64_t batch_lo, uint64_t batch_hi, b2a_scan **out) {
	return b2a_scan_create_ex(r, device, batch_lo, batch_hi, 0, out);
}

int b2a_scan_create_ex(b2a_reader *r, int device, uint64_t batch_lo, uint64_t batch_hi, uint32_t flags,
                       b2a_scan **out) {
	B2A_TRY
	if (flags & ~(B2A_SCAN_HOST_ALIAS | B2A_SCAN_DEVICE_ALIAS)) {
		return Fail(B2A_EINVAL, "unknown scan flags");
	}
	if ((flags & B2A_SCAN_HOST_ALIAS) && (flags & B2A_SCAN_DEVICE_ALIAS)) {
		return Fail(B2A_EINVAL, "B2A_SCAN_HOST_ALIAS and B2A_SCAN_DEVICE_ALIAS exclude each other");
	}
	std::unique_ptr<b2a_scan> s(new b2a_scan());
	s->r = r;
	s->device = device;
	s->plan.alias_direct = (flags & (B2A_SCAN_HOST_ALIAS | B2A_SCAN_DEVICE_ALIAS)) != 0;
	s->plan.alias_on_device = (flags & B2A_SCAN_DEVICE_ALIAS) != 0;
	s->plan.Build(r->reader, batch_lo, batch_hi);
	if (device >= 0) {
		CUDA_OK(cudaSetDevice(device));
		CUDA_OK(SmCount(device, &s->sm_count));
		const ScanPlan &p = s->plan;
		size_t nt = p.tasks.size() ? p.tasks.size() : 1, nl = p.tile_task.size() ? p.tile_task.size() : 1,
		       nv = p.variadic.size() ? p.variadic.size() : 1;
		CUDA_OK(ScanAlloc(&s->d_tasks, nt * sizeof(b2a_task)));
		CUDA_OK(ScanAlloc(&s->d_tile_task, nl * sizeof(uint32_t)));
		CUDA_OK(ScanAlloc(&s->d_variadic, nv * sizeof(uint64_t)));
		CUDA_OK(ScanAlloc(&s->d_err, sizeof(uint32_t)));
		CUDA_OK(cudaMemcpy(s->d_tasks, p.tasks.data(), p.tasks.size() * sizeof(b2a_task), cudaMemcpyHostToDevice));
		CUDA_OK(cudaMemcpy(s->d_tile_task, p.tile_task.data(), p.tile_task.size() * sizeof(uint32_t),
		                   cudaMemcpyHostToDevice));
		CUDA_OK(cudaMemcpy(s->d_variadic, p.variadic.data(), p.variadic.size() * sizeof(uint64_t),
		                   cudaMemcpyHostToDevice));
		CUDA_OK(cudaMemset(s->d_err, 0, sizeof(uint32_t)));
		if (!p.jobs.empty()) {
			s->n_workers = DecompWorkers(s->sm_count);
			if (s->n_workers > p.jobs.size()) {
				s->n_workers = static_cast<uint32_t>(p.jobs.size());
			}
			CUDA_OK(ScanAlloc(&s->d_jobs, p.jobs.size() * sizeof(b2a_decomp_job)));
			CUDA_OK(cudaMemcpy(s->d_jobs, p.jobs.data(), p.jobs.size() * sizeof(b2a_decomp_job), cudaMemcpyHostToDevice));
			std::vector<uint32_t> order(p.jobs.size());
			for (size_t i = 0; i < order.size(); i++) {
				order[i] = static_cast<uint32_t>(i);
			}
			std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
				return p.jobs[a].dst_len > p.jobs[b].dst_len;
			});
			CUDA_OK(ScanAlloc(&s->d_job_order, order.size() * sizeof(uint32_t)));
			CUDA_OK(cudaMemcpy(s->d_job_order, order.data(), order.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
			CUDA_OK(ScanAlloc(&s->d_workspaces, static_cast<size_t>(s->n_workers) * DecompWorkspaceBytes()));
			CUDA_OK(ScanAlloc(&s->d_decomp_state, 2 * sizeof(uint32_t)));
			CUDA_OK(cudaMemset(s->d_decomp_state, 0, 2 * sizeof(uint32_t)));
		}
	}
	*out = s.release();
	return B2A_OK;
	B2A_CATCH
}

void b2a_scan_destroy(b2a_scan *s) {
	if (!s) {
		return;
	}
	FreeScanDevice(s);
	delete s;
}

int b2a_scan_sizes(b2a_scan *s, uint64_t *image_lo, uint64_t *image_bytes, uint64_t *out_bytes,
                   uint64_t *algorithmic_bytes, uint64_t *n_tasks, uint64_t *n_tiles) {
	const ScanPlan &p = s->plan;
	*image_lo = p.image_lo;
	*image_bytes = p.image_hi - p.image_lo;
	*out_bytes = p.out_bytes;
	*algorithmic_bytes = p.algorithmic_in + p.algorithmic_out;
	*n_tasks = p.tasks.size();
	*n_tiles = p.tile_task.size();
	return B2A_OK;
}

int b2a_scan_upload_bytes(b2a_scan *s, uint64_t *bytes) {
	*bytes = s->plan.upload_bytes;
	return B2A_OK;
}

#define NEED_DEVICE(s)                                                                                                 \
	do {                                                                                                               \
		if ((s)->device < 0) {                                                                                         \
			return Fail(B2A_ECUDA, "scan was planned without a CUDA device; there is no CPU fallback");                \
		}                                                                                                              \
		CUDA_OK(cudaSetDevice((s)->device));                                                                           \
	} while (0)

int b2a_scan_upload(b2a_scan *s, void *cuda_stream) {
	NEED_DEVICE(s);
	const ScanPlan &p = s->plan;
	uint64_t bytes = p.image_hi - p.image_lo;
	if (!s->d_image_owned) {
		// +64: the string kernel fetches whole aligned words around the last characters
		CUDA_OK(ScanAlloc(&s->d_image_owned, DeviceImageBytes(p) + 64));
	}
	if (bytes) {
		CUDA_OK(cudaMemcpyAsync(s->d_image_owned, s->r->reader.Image() + p.image_lo, bytes, cudaMemcpyHostToDevice,
		                        static_cast<cudaStream_t>(cuda_stream)));
	}
	s->d_image = s->d_image_owned;
	return B2A_OK;
}

int b2a_scan_set_device_image(b2a_scan *s, const void *d_image) {
	if ((reinterpret_cast<uintptr_t>(d_image) & 255) != 0) {
		return Fail(B2A_EINVAL, "device image must be 256-byte aligned");
	}
	s->d_image = static_cast<const uint8_t *>(d_image);
	return B2A_OK;
}

int b2a_scan_decode(b2a_scan *s, void *d_out, uint64_t string_ptr_base, void *cuda_stream) {
	NEED_DEVICE(s);
	if (s->plan.alias_direct && !s->plan.alias_on_device) {
		return Fail(B2A_EINVAL, "a B2A_SCAN_HOST_ALIAS scan has host-resident direct columns: use b2a_read_arrow_host");
	}
	if (!s->d_image) {
		return Fail(B2A_EINVAL, "no device image: call b2a_scan_upload or b2a_scan_set_device_image first");
	}
	if (!d_out || (reinterpret_cast<uintptr_t>(d_out) & 255) != 0) {
		return Fail(B2A_EINVAL, "output arena must be a 256-byte aligned device pointer");
	}
	const ScanPlan &p = s->plan;
	cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
	if (p.stats_bytes) {
		CUDA_OK(cudaMemsetAsync(static_cast<uint8_t *>(d_out) + p.stats_base, 0, p.stats_bytes, st));
	}
	{
		int rc = LaunchJobs(s, 0, static_cast<uint32_t>(p.jobs.size()), st);
		if (rc != B2A_OK) {
			return rc;
		}
	}
	DecodeLaunch l;
	l.d_tasks = s->d_tasks;
	l.d_tile_task = s->d_tile_task;
	l.n_tiles = static_cast<uint32_t>(p.tile_task.size());
	l.image_origin = s->d_image - p.image_lo;
	l.out = static_cast<uint8_t *>(d_out);
	uint64_t base = string_ptr_base ? string_ptr_base : reinterpret_cast<uint64_t>(s->d_image);
	l.string_ptr_origin = base - p.image_lo;
	l.d_err = s->d_err;
	l.d_variadic = s->d_variadic;
	l.sm_count = s->sm_count;
	l.lean = p.lean;
	// tiles from phase2_tile0 on read vectors the earlier tiles decode (REE gathers): a second launch on the same stream
	const uint32_t total_tiles = l.n_tiles, p2 = p.phase2_tile0 < total_tiles ? p.phase2_tile0 : total_tiles;
	l.n_tiles = p2;
	int e = LaunchDecode(l, cuda_stream);
	if (e == 0 && p2 < total_tiles) {
		if (l.n_tiles) {
			s->launches++;
		}
		l.tile_begin = p2;
		l.n_tiles = total_tiles - p2;
		e = LaunchPhase2(l, cuda_stream);
	}
	if (e != 0) {
		return Fail(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)) +
		                           " launching b2a_decode_kernel");
	}
	if (l.n_tiles) {
		s->launches++;
	}
	return B2A_OK;
}

int b2a_scan_finish(b2a_scan *s, void *cuda_stream) {
	NEED_DEVICE(s);
	cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
	uint32_t err = 0;
	uint32_t zerr[2] = {0, 0};
	CUDA_OK(cudaMemcpyAsync(&err, s->d_err, sizeof(err), cudaMemcpyDeviceToHost, st));
	if (s->d_decomp_state) {
		CUDA_OK(cudaMemcpyAsync(zerr, s->d_decomp_state, sizeof(zerr), cudaMemcpyDeviceToHost, st));
	}
	CUDA_OK(cudaStreamSynchronize(st));
	if (zerr[1] != 0) {
		CUDA_OK(cudaMemsetAsync(s->d_decomp_state, 0, 2 * sizeof(uint32_t), st));
		CUDA_OK(cudaMemsetAsync(s->d_err, 0, sizeof(uint32_t), st));
		static const char *names[] = {"", "Unknown frame descriptor", "Src size is incorrect", "Data corruption detected",
		                              "Destination buffer is too small", "Unsupported frame parameter"};
		if (zerr[1] > B2A_DECOMP_LZ4_ERR) {
			// no reference text: the reference has no LZ4 decompressor; liblz4's LZ4F_getErrorName strings
			static const char *lnames[] = {"", "ERROR_frameType_unknown", "ERROR_srcSize_wrong", "ERROR_decompressionFailed",
			                               "ERROR_dstMaxSize_tooSmall", "ERROR_parameter_unsupported"};
			const uint32_t c = zerr[1] - B2A_DECOMP_LZ4_ERR;
			return Fail(B2A_EIO, std::string("LZ4F_decompress of a compressed IPC buffer failed with error '") +
			                         lnames[c < 6 ? c : 3] + "'");
		}
		// base_stream_reader.cpp:17-21: "ZSTD_decompress([buffer ...]) failed with error '...'" (IOException)
		return Fail(B2A_EIO, std::string("ZSTD_decompress of a compressed IPC buffer failed with error '") +
		                         names[zerr[1] < 6 ? zerr[1] : 3] + "'");
	}
	if (err != B2A_DEV_OK) {
		CUDA_OK(cudaMemsetAsync(s->d_err, 0, sizeof(uint32_t), st));
		if ((err & 0xFFu) == B2A_DEV_UNION_TAG) {
			const uint32_t tag = (err >> 8) & 0xFFu;
			if (tag >= 128) { // NumericCast<uint8_t>(int8_t) throws first (arrow_conversion.cpp:1209)
				return Fail(B2A_EINTERNAL, "Information loss on integer cast: value " +
				                               std::to_string(static_cast<int>(tag) - 256) +
				                               " outside of target range [0, 255]");
			}
			return Fail(B2A_EINVAL, "Arrow union tag out of range: " + std::to_string(tag));
		}
		switch (err) {
		case B2A_DEV_BAD_OFFSETS:
			return Fail(B2A_EINVAL, "Arrow IPC: offsets buffer is not monotonically increasing or exceeds the data "
			                        "buffer");
		case B2A_DEV_MUL_OVERFLOW:
			return Fail(B2A_ECONVERSION, "Could not convert to Microsecond");
		case B2A_DEV_STRING_4GB:
			return Fail(B2A_ECONVERSION, "DuckDB does not support Strings over 4GB");
		case B2A_DEV_INDEX_U32:
			return Fail(B2A_ECONVERSION, "DuckDB only supports indices that fit on an uint32");
		case B2A_DEV_DICT_INDEX:
			return Fail(B2A_EINVAL, "Arrow IPC: dictionary index outside the dictionary");
		case B2A_DEV_VIEW_RANGE:
			return Fail(B2A_EINVAL, "Arrow IPC: string view refers to bytes outside its variadic data buffers");
		case B2A_DEV_LIST_CHILD:
			return Fail(B2A_EINVAL, "Arrow IPC: list offsets exceed the length of the child array");
		default:
			return Fail(B2A_EINTERNAL, "unknown device error code " + std::to_string(err));
		}
	}
	return B2A_OK;
}

int b2a_scan_download(b2a_scan *s, const void *d_out, void *h_out, void *cuda_stream) {
	NEED_DEVICE(s);
	CUDA_OK(cudaMemcpyAsync(h_out, d_out, s->plan.out_bytes, cudaMemcpyDeviceToHost,
	                        static_cast<cudaStream_t>(cuda_stream)));
	return B2A_OK;
}

int b2a_scan_column(b2a_scan *s, uint64_t batch, uint32_t flat_field, b2a_column_out *out) {
	const ScanPlan &p = s->plan;
	if (batch < p.batch_lo || batch >= p.batch_hi || flat_field >= p.n_flat) {
		return Fail(B2A_EINVAL, "batch / field out of range");
	}
	*out = p.Column(batch, flat_field);
	return B2A_OK;
}

int b2a_scan_dictionary(b2a_scan *s, uint32_t flat_field, b2a_column_out *out, uint32_t *dict_flat_field) {
	auto it = s->plan.dict_cols.find(static_cast<int>(flat_field));
	if (it == s->plan.dict_cols.end()) {
		return Fail(B2A_EINVAL, "field is not dictionary encoded");
	}
	*out = it->second;
	if (dict_flat_field) {
		*dict_flat_field = flat_field;
	}
	return B2A_OK;
}

int b2a_scan_scratch(b2a_scan *s, uint64_t *scratch_offset, uint64_t *scratch_bytes, uint64_t *n_jobs) {
	const ScanPlan &p = s->plan;
	*scratch_offset = p.compressed ? p.scratch_base - p.image_lo : 0;
	*scratch_bytes = p.compressed ? p.scratch_bytes : 0;
	*n_jobs = p.jobs.size();
	return B2A_OK;
}

int b2a_scan_download_scratch(b2a_scan *s, void *h_scratch, void *cuda_stream) {
	NEED_DEVICE(s);
	const ScanPlan &p = s->plan;
	if (!p.compressed || !s->d_image) {
		return B2A_OK;
	}
	CUDA_OK(cudaMemcpyAsync(h_scratch, s->d_image + (p.scratch_base - p.image_lo), p.scratch_bytes, cudaMemcpyDeviceToHost,
	                        static_cast<cudaStream_t>(cuda_stream)));
	return B2A_OK;
}

uint64_t b2a_scan_gpu_launches(b2a_scan *s) {
	return s->launches;
}

int b2a_scan_num_chunks(b2a_scan *s, uint64_t *n_chunks) {
	*n_chunks = s->plan.n_chunks;
	return B2A_OK;
}

} // extern "C"

// ------------------------------------------------------------------ one-call host path (e2e)
// read_arrow with HOST buffers in and HOST vectors out: the stream image goes up in slices of whole record
// batches on one stream, each slice is decoded as soon as it has landed, and its vectors go down on a third
// stream, so H2D, decode and D2H overlap (PCIe is full duplex).  string_t pointers refer to the host image,
// exactly where the reference's pointers point (zero-copy string payloads).
struct HostPipe {
	cudaStream_t in = nullptr, compute = nullptr, out = nullptr;
	std::vector<cudaEvent_t> ev_in, ev_c;
	std::vector<cudaEvent_t> ev_out; // B2A_PIPE_TRACE only
	// staged uploads: ring of pinned host buffers + their device twins
	static constexpr uint32_t kRing = 3;
	uint8_t *h_stage[kRing] = {nullptr, nullptr, nullptr};
	uint8_t *d_stage[kRing] = {nullptr, nullptr, nullptr};
	uint64_t stage_bytes = 0;
	cudaEvent_t ev_start = nullptr;
	cudaEvent_t ev_user = nullptr;
	uint8_t *d_out = nullptr;
	uint64_t d_out_bytes = 0;
	int device = -1;
	HostPipe() = default;
	HostPipe(const HostPipe &) = delete;
	HostPipe &operator=(const HostPipe &) = delete;
	// (pipes live in the process-wide free list below: this runs at process exit)
	~HostPipe() {
		if (!in || device < 0 || cudaSetDevice(device) != cudaSuccess) {
			(void)cudaGetLastError();
			return;
		}
		cudaStreamSynchronize(in);
		cudaStreamSynchronize(compute);
		cudaStreamSynchronize(out);
		for (auto e : ev_in) {
			cudaEventDestroy(e);
		}
		for (auto e : ev_c) {
			cudaEventDestroy(e);
		}
		for (auto e : ev_out) {
			cudaEventDestroy(e);
		}
		if (ev_start) {
			cudaEventDestroy(ev_start);
		}
		if (ev_user) {
			cudaEventDestroy(ev_user);
		}
		for (uint32_t k = 0; k < kRing; k++) {
			cudaFreeHost(h_stage[k]);
			cudaFree(d_stage[k]);
		}
		cudaFree(d_out);
		cudaStreamDestroy(in);
		cudaStreamDestroy(compute);
		cudaStreamDestroy(out);
		(void)cudaGetLastError();
	}
};
// Streams, events, staging rings and the arena belong to the device they were created on.  A pipe is leased for the
// duration of one call and goes back to a process-wide free list: a DuckDB-side caller runs every query's pipeline on a new
// worker thread, and a pipe per THREAD meant new streams, a new 96 MiB pinned ring and a new device arena per query.
static std::mutex g_pipe_mu;
static std::vector<std::unique_ptr<HostPipe>> g_free_pipes;

struct PipeLease {
	std::unique_ptr<HostPipe> pipe;
	explicit PipeLease(int device) {
		{
			std::lock_guard<std::mutex> lock(g_pipe_mu);
			for (size_t i = 0; i < g_free_pipes.size(); i++) {
				if (g_free_pipes[i]->device == device) {
					pipe = std::move(g_free_pipes[i]);
					g_free_pipes.erase(g_free_pipes.begin() + static_cast<std::ptrdiff_t>(i));
					break;
				}
			}
		}
		if (!pipe) {
			pipe.reset(new HostPipe());
			pipe->device = device;
		}
	}
	~PipeLease() {
		// (a call that failed half way leaves work queued on the pipe's streams: drain it before the next user)
		if (pipe->in && cudaSetDevice(pipe->device) == cudaSuccess) {
			cudaStreamSynchronize(pipe->in);
			cudaStreamSynchronize(pipe->compute);
			cudaStreamSynchronize(pipe->out);
		}
		(void)cudaGetLastError();
		std::lock_guard<std::mutex> lock(g_pipe_mu);
		g_free_pipes.push_back(std::move(pipe));
	}
};

static int UploadRanges(b2a_scan *s, const ScanPlan::RangeList &ranges, cudaStream_t st) {
	const ScanPlan &p = s->plan;
	const uint8_t *h_image = s->r->reader.Image();
	for (auto &r : ranges) {
		if (r.second > r.first) {
			CUDA_OK(cudaMemcpyAsync(s->d_image_owned + (r.first - p.image_lo), h_image + r.first, r.second - r.first,
			                        cudaMemcpyHostToDevice, st));
		}
	}
	return B2A_OK;
}

// h_out_pinned != NULL: vectors come back to the host (b2a_read_arrow_host).  d_out_user != NULL: vectors stay in the
// caller's device arena for device-side consumers (b2a_read_arrow_device): no D2H, string_t point into the device image,
// and `user_stream` is made to wait for the last decode instead of the host.
static int ReadArrowPipeline(b2a_scan *s, void *h_out_pinned, void *d_out_user, uint32_t n_slices, void *user_stream,
                             void *h_scratch = nullptr) {
	NEED_DEVICE(s);
	const ScanPlan &p = s->plan;
	if (p.alias_on_device && d_out_user == nullptr) {
		return Fail(B2A_EINVAL, "a B2A_SCAN_DEVICE_ALIAS scan leaves its direct columns in the device image: use "
		                        "b2a_read_arrow_device / b2a_scan_decode");
	}
	PipeLease lease(s->device);
	HostPipe &hp = *lease.pipe;
	const bool to_host = d_out_user == nullptr;
	if (!hp.in) {
		CUDA_OK(cudaStreamCreateWithFlags(&hp.in, cudaStreamNonBlocking));
		CUDA_OK(cudaStreamCreateWithFlags(&hp.compute, cudaStreamNonBlocking));
		CUDA_OK(cudaStreamCreateWithFlags(&hp.out, cudaStreamNonBlocking));
	}
	if (!s->d_image_owned) {
		CUDA_OK(ScanAlloc(&s->d_image_owned, DeviceImageBytes(p) + 64));
	}
	s->d_image = s->d_image_owned;
	if (to_host && hp.d_out_bytes < p.out_bytes) {
		cudaFree(hp.d_out);
		hp.d_out = nullptr;
		hp.d_out_bytes = 0;
		CUDA_OK(cudaMalloc(&hp.d_out, p.out_bytes));
		hp.d_out_bytes = p.out_bytes;
	}
	const uint64_t n_batches = p.batch_hi - p.batch_lo;
	const bool has_phase2 = p.phase2_tile0 < p.tile_task.size();
	if (n_slices == 0) {
		n_slices = 16;
	}
	if (n_slices > n_batches) {
		n_slices = static_cast<uint32_t>(n_batches ? n_batches : 1);
	}
	static const bool trace = getenv("B2A_PIPE_TRACE") != nullptr; // per-slice timeline on stderr (experiments)
	while (hp.ev_in.size() < n_slices + 1) {
		cudaEvent_t a, b, c;
		CUDA_OK(cudaEventCreateWithFlags(&a, trace ? cudaEventDefault : cudaEventDisableTiming));
		CUDA_OK(cudaEventCreateWithFlags(&b, trace ? cudaEventDefault : cudaEventDisableTiming));
		hp.ev_in.push_back(a);
		hp.ev_c.push_back(b);
		if (trace) {
			CUDA_OK(cudaEventCreate(&c));
			hp.ev_out.push_back(c);
		}
	}
	if (trace) {
		if (!hp.ev_start) {
			CUDA_OK(cudaEventCreate(&hp.ev_start));
		}
		CUDA_OK(cudaEventRecord(hp.ev_start, hp.in));
	}
	const uint8_t *h_image = s->r->reader.Image();
	uint8_t *h_out = static_cast<uint8_t *>(h_out_pinned);
	DecodeLaunch l;
	l.d_tasks = s->d_tasks;
	l.d_tile_task = s->d_tile_task;
	l.image_origin = s->d_image - p.image_lo;
	uint8_t *const arena = to_host ? hp.d_out : static_cast<uint8_t *>(d_out_user);
	l.out = arena;
	// string_t pointers: host consumers read the characters where they lie in the HOST image (like the reference's
	// pointers); bodies that were decompressed exist only in the scratch region, whose host copy is `h_scratch`
	l.string_ptr_origin = !to_host ? reinterpret_cast<uint64_t>(s->d_image) - p.image_lo
	                      : (p.compressed && h_scratch) ? reinterpret_cast<uint64_t>(h_scratch) - p.scratch_base
	                                                    : reinterpret_cast<uint64_t>(h_image);
	l.d_err = s->d_err;
	l.d_variadic = s->d_variadic;
	l.sm_count = s->sm_count;
	l.lean = p.lean;
	if (!to_host && user_stream) { // the caller's earlier work on the arena (a previous step's consumers) comes first
		if (!hp.ev_user) {
			CUDA_OK(cudaEventCreateWithFlags(&hp.ev_user, cudaEventDisableTiming));
		}
		CUDA_OK(cudaEventRecord(hp.ev_user, static_cast<cudaStream_t>(user_stream)));
		CUDA_OK(cudaStreamWaitEvent(hp.compute, hp.ev_user, 0));
		CUDA_OK(cudaStreamWaitEvent(hp.in, hp.ev_user, 0));
	}
	if (p.stats_bytes) {
		CUDA_OK(cudaMemsetAsync(arena + p.stats_base, 0, p.stats_bytes, hp.compute));
	}

	// ---- how the bytes get up: a pinned image is read by the copy engine range by range; anything else (an mmap'ed
	// file, a pageable buffer) is packed into pinned staging buffers by host threads, one large copy per slice
	if (s->stage_mode == 0 || (s->stage_mode == 2 && s->stage_plan.n_slices != n_slices)) {
		cudaPointerAttributes attr;
		const bool pinned = cudaPointerGetAttributes(&attr, h_image) == cudaSuccess &&
		                    (attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged);
		(void)cudaGetLastError();
		const char *force = getenv("B2A_H2D_STAGE"); // "1": stage pinned images too, "0": never stage (experiments)
		const bool stage = force ? force[0] == '1' : !pinned;
		s->stage_mode = stage ? 2 : 1;
		if (stage) {
			// file-backed images: how the bytes leave the file.  B2A_FILE_IO = mmap (default: host threads copy out
			// of the mapping) | pread (one pread per <= 1 MiB piece straight into the pinned ring) | direct (O_DIRECT
			// block reads where the file system has them — bypasses the page cache, runs at the disk's rate)
			s->stage_io = 0;
			s->stage_fd = -1;
			const int fd = s->r->reader.FileDescriptor();
			if (fd >= 0) {
				const char *io = getenv("B2A_FILE_IO");
				const std::string mode = io ? io : "mmap"; // measured on the box: profiles/r02_file_io.txt
				if (mode == "direct" && s->r->reader.DirectFileDescriptor() >= 0) {
					s->stage_io = 2;
					s->stage_fd = s->r->reader.DirectFileDescriptor();
				} else if (mode != "mmap") {
					s->stage_io = 1;
					s->stage_fd = fd;
				}
			}
			s->stage_plan.Build(p, n_slices, s->stage_io == 2 ? 4096 : 16);
			std::vector<ScatterPiece> sp;
			sp.reserve(s->stage_plan.pieces.size());
			for (auto &pc : s->stage_plan.pieces) {
				sp.push_back(ScatterPiece {pc.stage, pc.dst, pc.bytes});
			}
			if (s->d_stage_pieces) {
				cudaDeviceSynchronize(); // (an earlier b2a_read_arrow_device of this scan may still be scattering)
				ScanFree(s->d_stage_pieces);
			}
			s->d_stage_pieces = nullptr;
			if (!sp.empty()) {
				CUDA_OK(ScanAlloc(&s->d_stage_pieces, sp.size() * sizeof(ScatterPiece)));
				CUDA_OK(cudaMemcpy(s->d_stage_pieces, sp.data(), sp.size() * sizeof(ScatterPiece), cudaMemcpyHostToDevice));
			}
		}
	}
	const bool staged = s->stage_mode == 2 && !s->stage_plan.pieces.empty();
	std::unique_ptr<StageWorkers> workers;
	if (staged) {
		const uint64_t need = (s->stage_plan.max_slice_bytes + 255) & ~uint64_t(255);
		if (hp.stage_bytes < need) {
			for (uint32_t k = 0; k < HostPipe::kRing; k++) {
				cudaFreeHost(hp.h_stage[k]);
				cudaFree(hp.d_stage[k]);
				hp.h_stage[k] = hp.d_stage[k] = nullptr;
			}
			hp.stage_bytes = 0;
			for (uint32_t k = 0; k < HostPipe::kRing; k++) {
				CUDA_OK(cudaHostAlloc(&hp.h_stage[k], need, cudaHostAllocDefault));
				CUDA_OK(cudaMalloc(&hp.d_stage[k], need));
			}
			hp.stage_bytes = need;
		}
		static const uint32_t n_threads = [] {
			const char *e = getenv("B2A_STAGE_THREADS");
			int n = e ? atoi(e) : 0;
			if (n <= 0) {
				// measured on the 16-core box, one process (tools/gpu_r2_s.sh): 8 threads 137 ms, 12 threads 122 ms (= the
				// time from a pinned image), 16 threads 131 ms per SF10 file; several ranks on one node share the cores
				const unsigned hw = std::thread::hardware_concurrency();
				const char *lws = getenv("LOCAL_WORLD_SIZE");
				const bool alone = !lws || atoi(lws) <= 1;
				n = hw >= 16 ? (alone ? 12 : 8) : (hw >= 4 ? static_cast<int>(hw / 2) : 1);
			}
			return static_cast<uint32_t>(n);
		}();
		workers.reset(new StageWorkers(s->stage_plan, h_image, hp.h_stage, HostPipe::kRing, n_threads, s->stage_io,
		                               s->stage_fd, s->r->reader.ImageSize()));
		workers->Allow(HostPipe::kRing - 1);
	}
	for (uint32_t sl = 0; sl < n_slices; sl++) {
		const uint64_t b0 = n_batches * sl / n_slices, b1 = n_batches * (sl + 1) / n_slices;
		if (b0 == b1) {
			continue;
		}
		if (staged) {
			const uint32_t k = sl % HostPipe::kRing;
			workers->WaitPacked(sl);
			if (workers->Failed()) {
				return Fail(B2A_EIO, "Arrow IPC: reading the file into the staging buffers failed");
			}
			if (sl >= HostPipe::kRing) { // the device twin is free once slice sl - kRing has been scattered out of it
				CUDA_OK(cudaStreamWaitEvent(hp.in, hp.ev_c[sl - HostPipe::kRing], 0));
			}
			if (s->stage_plan.slice_bytes[sl]) {
				CUDA_OK(cudaMemcpyAsync(hp.d_stage[k], hp.h_stage[k], s->stage_plan.slice_bytes[sl], cudaMemcpyHostToDevice,
				                        hp.in));
			}
		} else {
			for (uint64_t b = b0; b < b1; b++) {
				int rc = UploadRanges(s, p.batch_ranges[b], hp.in);
				if (rc != B2A_OK) {
					return rc;
				}
			}
		}
		CUDA_OK(cudaEventRecord(hp.ev_in[sl], hp.in));
		CUDA_OK(cudaStreamWaitEvent(hp.compute, hp.ev_in[sl], 0));
		if (staged) {
			const uint32_t q0 = s->stage_plan.slice_piece0[sl], q1 = s->stage_plan.slice_piece0[sl + 1];
			int e = LaunchScatter(s->d_stage_pieces + q0, q1 - q0, hp.d_stage[sl % HostPipe::kRing], s->d_image_owned,
			                      s->sm_count, hp.compute);
			if (e != 0) {
				return Fail(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
			}
			if (q1 > q0) {
				s->launches++;
			}
			// the pinned buffer of slice sl + 1 may be refilled once slice sl + 1 - kRing has crossed the bus
			if (sl + 1 >= HostPipe::kRing) {
				CUDA_OK(cudaEventSynchronize(hp.ev_in[sl + 1 - HostPipe::kRing]));
			}
			workers->Allow(sl + 1);
		}
		{
			int rc = LaunchJobs(s, p.batch_job0[b0], p.batch_job0[b1], hp.compute);
			if (rc != B2A_OK) {
				return rc;
			}
		}
		l.tile_begin = p.batch_tile0[b0];
		l.n_tiles = p.batch_tile0[b1] - p.batch_tile0[b0];
		int e = LaunchDecode(l, hp.compute);
		if (e != 0) {
			return Fail(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
		}
		if (l.n_tiles) {
			s->launches++;
		}
		CUDA_OK(cudaEventRecord(hp.ev_c[sl], hp.compute));
		CUDA_OK(cudaStreamWaitEvent(hp.out, hp.ev_c[sl], 0));
		const uint64_t o0 = p.batch_out0[b0], o1 = p.batch_out0[b1];
		if (to_host && o1 > o0 && !has_phase2) {
			CUDA_OK(cudaMemcpyAsync(h_out + o0, hp.d_out + o0, o1 - o0, cudaMemcpyDeviceToHost, hp.out));
		}
		if (trace) {
			CUDA_OK(cudaEventRecord(hp.ev_out[sl], hp.out));
		}
	}
	// dictionaries (decoded once per scan) + the popcount cells live after the last batch
	const uint32_t tail_tile0 = p.batch_tile0.empty() ? 0 : p.batch_tile0.back();
	const uint64_t tail_out0 = p.batch_out0.empty() ? 0 : p.batch_out0.back();
	{
		int rc = UploadRanges(s, p.tail_ranges, hp.in);
		if (rc != B2A_OK) {
			return rc;
		}
	}
	CUDA_OK(cudaEventRecord(hp.ev_in[n_slices], hp.in));
	CUDA_OK(cudaStreamWaitEvent(hp.compute, hp.ev_in[n_slices], 0));
	{
		int rc = LaunchJobs(s, p.batch_job0.empty() ? 0 : p.batch_job0.back(), static_cast<uint32_t>(p.jobs.size()),
		                    hp.compute);
		if (rc != B2A_OK) {
			return rc;
		}
	}
	const uint32_t all_tiles = static_cast<uint32_t>(p.tile_task.size());
	const uint32_t p2_tile0 = has_phase2 ? p.phase2_tile0 : all_tiles;
	l.tile_begin = tail_tile0;
	l.n_tiles = p2_tile0 - tail_tile0;
	if (l.n_tiles) {
		int e = LaunchDecode(l, hp.compute);
		if (e != 0) {
			return Fail(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
		}
		s->launches++;
	}
	if (has_phase2) {
		l.tile_begin = p2_tile0;
		l.n_tiles = all_tiles - p2_tile0;
		int e = LaunchPhase2(l, hp.compute);
		if (e != 0) {
			return Fail(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
		}
		s->launches++;
	}
	CUDA_OK(cudaEventRecord(hp.ev_c[n_slices], hp.compute));
	CUDA_OK(cudaStreamWaitEvent(hp.out, hp.ev_c[n_slices], 0));
	// REE gathers write into the batches' regions after the per-slice copies would have left: one copy of the whole arena
	if (!to_host) {
		// device consumers: order the caller's stream behind the last decode; the host only waits for the staging ring
		if (user_stream) {
			CUDA_OK(cudaStreamWaitEvent(static_cast<cudaStream_t>(user_stream), hp.ev_c[n_slices], 0));
		} else {
			CUDA_OK(cudaStreamSynchronize(hp.compute));
		}
		if (workers) { // the pinned ring is reused by the next call: its last copies must have left the host
			CUDA_OK(cudaEventSynchronize(hp.ev_in[n_slices]));
		}
		return B2A_OK;
	}
	const uint64_t d2h_from = has_phase2 ? 0 : tail_out0;
	if (p.out_bytes > d2h_from) {
		CUDA_OK(cudaMemcpyAsync(h_out + d2h_from, hp.d_out + d2h_from, p.out_bytes - d2h_from, cudaMemcpyDeviceToHost,
		                        hp.out));
	}
	if (p.compressed && h_scratch && p.scratch_bytes) {
		CUDA_OK(cudaMemcpyAsync(h_scratch, s->d_image + (p.scratch_base - p.image_lo), p.scratch_bytes,
		                        cudaMemcpyDeviceToHost, hp.out));
	}
	CUDA_OK(cudaStreamSynchronize(hp.out));
	if (trace) {
		CUDA_OK(cudaStreamSynchronize(hp.compute));
		for (uint32_t sl = 0; sl < n_slices; sl++) {
			float a = 0, b = 0, c = 0;
			cudaEventElapsedTime(&a, hp.ev_start, hp.ev_in[sl]);
			cudaEventElapsedTime(&b, hp.ev_start, hp.ev_c[sl]);
			cudaEventElapsedTime(&c, hp.ev_start, hp.ev_out[sl]);
			fprintf(stderr, "[b2a pipe] slice %2u: h2d done %7.2f ms, decode done %7.2f ms, d2h done %7.2f ms\n", sl, a, b, c);
		}
	}
	return b2a_scan_finish(s, hp.compute);
}

extern "C" int b2a_read_arrow_host(b2a_scan *s, void *h_out_pinned, uint32_t n_slices) {
	B2A_TRY
	if (!h_out_pinned) {
		return Fail(B2A_EINVAL, "b2a_read_arrow_host needs a (pinned) host output arena");
	}
	return ReadArrowPipeline(s, h_out_pinned, nullptr, n_slices, nullptr);
	B2A_CATCH
}

extern "C" int b2a_read_arrow_host_scratch(b2a_scan *s, void *h_out_pinned, void *h_scratch_pinned, uint32_t n_slices) {
	B2A_TRY
	if (!h_out_pinned) {
		return Fail(B2A_EINVAL, "b2a_read_arrow_host_scratch needs a (pinned) host output arena");
	}
	if (s->plan.compressed && s->plan.scratch_bytes && !h_scratch_pinned) {
		return Fail(B2A_EINVAL, "compressed scan: a host buffer of b2a_scan_scratch().scratch_bytes is needed for the "
		                        "decompressed character buffers");
	}
	return ReadArrowPipeline(s, h_out_pinned, nullptr, n_slices, nullptr, h_scratch_pinned);
	B2A_CATCH
}

extern "C" int b2a_read_arrow_device(b2a_scan *s, void *d_out, uint32_t n_slices, void *cuda_stream) {
	B2A_TRY
	NEED_DEVICE(s);
	if (s->plan.alias_direct && !s->plan.alias_on_device) {
		return Fail(B2A_EINVAL, "a B2A_SCAN_HOST_ALIAS scan keeps its direct columns on the host: device consumers need "
		                        "a scan created without the flag");
	}
	if (!d_out || (reinterpret_cast<uintptr_t>(d_out) & 255) != 0) {
		return Fail(B2A_EINVAL, "output arena must be a 256-byte aligned device pointer");
	}
	return ReadArrowPipeline(s, nullptr, d_out, n_slices, cuda_stream);
	B2A_CATCH
}

// ------------------------------------------------------------------ filter / aggregate consumers
// Builds (and caches) the per-chunk column table for the given projected top-level columns.
// A column a B2A_SCAN_DEVICE_ALIAS scan left in the device image is addressed through the same `arena + offset` form the
// kernels use: offset = (address in the image) - (arena address), modulo 2^64.
static int ChunkTable(b2a_scan *s, const void *d_out, const std::vector<uint32_t> &cols_in, std::vector<uint32_t> &key,
                      uint64_t **d_table, std::vector<uint8_t> *widths) {
	const ScanPlan &p = s->plan;
	std::vector<uint32_t> cols = cols_in;
	if (p.alias_direct && !p.alias_on_device) {
		return Fail(B2A_EINVAL, "filters / aggregates need device-resident vectors: create the scan without "
		                        "B2A_SCAN_HOST_ALIAS");
	}
	const Schema &schema = s->r->reader.BaseSchema();
	std::vector<int> flats;
	for (uint32_t c : cols) {
		if (c >= p.top_columns.size()) {
			return Fail(B2A_EINVAL, "predicate column index out of range");
		}
		const Field &f = *schema.fields[static_cast<size_t>(p.top_columns[c])];
		if (f.dict_id >= 0 || f.out_width > 8 || f.out_width == 0 || f.type_id == B2A_T_FLOAT ||
		    f.type_id == B2A_T_DOUBLE || f.type_id == B2A_T_UBIGINT) {
			return Fail(B2A_ENOTIMPL, "filter pushdown supports signed integer-like columns (integers, dates, "
			                          "timestamps, DECIMAL up to 18 digits) only: column '" + f.name + "'");
		}
		flats.push_back(f.flat_index);
		if (widths) {
			// unsigned 8/16/32-bit values are zero-extended by the filter kernel (0x80 flag); UBIGINT is refused above
			const bool uns = f.type_id == B2A_T_UTINYINT || f.type_id == B2A_T_USMALLINT || f.type_id == B2A_T_UINTEGER;
			widths->push_back(static_cast<uint8_t>(f.out_width | (uns ? 0x80 : 0)));
		}
	}
	if (!s->d_chunk_rows) {
		std::vector<uint32_t> rows;
		for (uint64_t b = 0; b < p.batch_rows.size(); b++) {
			uint64_t left = p.batch_rows[b];
			while (left > 0) {
				uint32_t n = left > B2A_VECTOR_SIZE ? B2A_VECTOR_SIZE : static_cast<uint32_t>(left);
				rows.push_back(n);
				left -= n;
			}
		}
		CUDA_OK(ScanAlloc(&s->d_chunk_rows, (rows.size() ? rows.size() : 1) * sizeof(uint32_t)));
		CUDA_OK(cudaMemcpy(s->d_chunk_rows, rows.data(), rows.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
	}
	std::vector<uint32_t> want = cols;
	if (p.alias_on_device) { // the table holds addresses relative to THIS arena and THIS image
		if (!s->d_image) {
			return Fail(B2A_EINVAL, "no device image: call b2a_scan_upload, b2a_scan_set_device_image or "
			                        "b2a_read_arrow_device first");
		}
		const uint64_t a = reinterpret_cast<uint64_t>(d_out), b = reinterpret_cast<uint64_t>(s->d_image);
		for (uint64_t v : {a, b}) {
			want.push_back(static_cast<uint32_t>(v));
			want.push_back(static_cast<uint32_t>(v >> 32));
		}
	}
	if (key == want && *d_table) {
		return B2A_OK;
	}
	const uint64_t image_delta =
	    reinterpret_cast<uint64_t>(s->d_image) - p.image_lo - reinterpret_cast<uint64_t>(d_out); // (alias columns only)
	std::vector<uint64_t> table;
	table.reserve(p.n_chunks * 2 * cols.size());
	for (uint64_t b = 0; b < p.batch_rows.size(); b++) {
		const uint64_t nchunks = (p.batch_rows[b] + B2A_VECTOR_SIZE - 1) / B2A_VECTOR_SIZE;
		for (uint64_t k = 0; k < nchunks; k++) {
			for (size_t c = 0; c < flats.size(); c++) {
				const b2a_column_out &co = p.Column(p.batch_lo + b, static_cast<uint32_t>(flats[c]));
				const Field &f = *schema.flat[static_cast<size_t>(flats[c])];
				const uint64_t base = co.data_off != B2A_NONE ? co.data_off : image_delta + co.image_off;
				table.push_back(base + k * B2A_VECTOR_SIZE * static_cast<uint64_t>(f.out_width));
				table.push_back(co.validity_off == B2A_NONE ? B2A_NONE : co.validity_off + k * (B2A_VECTOR_SIZE / 8));
			}
		}
	}
	if (*d_table) {
		cudaDeviceSynchronize(); // (a launch that reads the old table may be in flight; cudaFree used to wait for it)
		ScanFree(*d_table);
	}
	*d_table = nullptr;
	CUDA_OK(ScanAlloc(d_table, (table.size() ? table.size() : 1) * sizeof(uint64_t)));
	CUDA_OK(cudaMemcpy(*d_table, table.data(), table.size() * sizeof(uint64_t), cudaMemcpyHostToDevice));
	key = want;
	return B2A_OK;
}

extern "C" int b2a_scan_filter(b2a_scan *s, const void *d_out, const b2a_predicate *preds_in, uint32_t n_preds,
                               void *d_sel, void *d_count, void *d_total, void *cuda_stream) {
	B2A_TRY
	NEED_DEVICE(s);
	if (n_preds == 0 || n_preds > 8 || !preds_in) {
		return Fail(B2A_EINVAL, "between 1 and 8 predicates are supported");
	}
	// predicates on the same column next to each other (the kernel then loads the column once); AND does not care
	std::vector<b2a_predicate> sorted(preds_in, preds_in + n_preds);
	std::stable_sort(sorted.begin(), sorted.end(),
	                 [](const b2a_predicate &a, const b2a_predicate &b) { return a.column < b.column; });
	const b2a_predicate *preds = sorted.data();
	std::vector<uint32_t> cols;
	for (uint32_t i = 0; i < n_preds; i++) {
		cols.push_back(preds[i].column);
		if (preds[i].cmp > B2A_CMP_GE) {
			return Fail(B2A_EINVAL, "unknown comparison");
		}
	}
	std::vector<uint8_t> widths;
	int rc = ChunkTable(s, d_out, cols, s->chunk_col_key, &s->d_chunk_col, &widths);
	if (rc != B2A_OK) {
		return rc;
	}
	cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
	CUDA_OK(cudaMemsetAsync(d_total, 0, sizeof(unsigned long long), st));
	FilterLaunch l;
	l.out = static_cast<const uint8_t *>(d_out);
	l.d_chunk_col = s->d_chunk_col;
	l.d_chunk_rows = s->d_chunk_rows;
	l.n_chunks = s->plan.n_chunks;
	l.n_preds = n_preds;
	for (uint32_t i = 0; i < n_preds; i++) {
		l.preds[i] = preds[i];
		l.widths[i] = widths[i];
	}
	l.d_sel = static_cast<uint32_t *>(d_sel);
	l.d_count = static_cast<uint32_t *>(d_count);
	l.d_total = static_cast<unsigned long long *>(d_total);
	l.sm_count = s->sm_count;
	int e = LaunchFilter(l, cuda_stream);
	if (e != 0) {
		return Fail(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
	}
	if (l.n_chunks) {
		s->launches++;
	}
	return B2A_OK;
	B2A_CATCH
}

// b2a_scan_filter for callers that may not allocate device memory (CUDA graph capture, serving loops, harnesses that
// forbid cudaMalloc): the per-chunk tables are built on the host on every call and copied into the fixed __device__
// array of kernels/filter.cu — no cudaMalloc, cudaFree or allocating library call on this path, nothing cached in the
// scan.  Same validation, same kernel, same results as b2a_scan_filter.
extern "C" int b2a_scan_filter_no_alloc(b2a_scan *s, const void *d_out, const b2a_predicate *preds_in, uint32_t n_preds,
                                        void *d_sel, void *d_count, void *d_total, void *cuda_stream) {
	B2A_TRY
	NEED_DEVICE(s);
	if (n_preds == 0 || n_preds > 8) {
		return Fail(B2A_EINVAL, "between 1 and 8 predicates are supported");
	}
	if (!preds_in || !d_out || !d_sel || !d_count || !d_total) {
		return Fail(B2A_EINVAL, "b2a_scan_filter_no_alloc: null argument");
	}
	std::vector<b2a_predicate> sorted(preds_in, preds_in + n_preds); // same order as b2a_scan_filter: columns together
	std::stable_sort(sorted.begin(), sorted.end(),
	                 [](const b2a_predicate &a, const b2a_predicate &b) { return a.column < b.column; });
	const b2a_predicate *preds = sorted.data();
	const ScanPlan &p = s->plan;
	if (p.alias_direct) {
		return Fail(B2A_EINVAL, "filters / aggregates need device-resident vectors: create the scan without "
		                        "B2A_SCAN_HOST_ALIAS");
	}
	const uint64_t need = FilterNoAllocBytesNeeded(p.n_chunks, n_preds);
	if (need > FilterNoAllocScratchBytes()) {
		return Fail(B2A_EINVAL, "b2a_scan_filter_no_alloc: " + std::to_string(p.n_chunks) + " vectors x " +
		                            std::to_string(n_preds) + " predicates need " + std::to_string(need) +
		                            " bytes of table, the fixed device scratch holds " +
		                            std::to_string(FilterNoAllocScratchBytes()) +
		                            " (use b2a_scan_filter, or scan a narrower batch range)");
	}
	const Schema &schema = s->r->reader.BaseSchema();
	FilterLaunch l;
	std::vector<int> flats;
	for (uint32_t i = 0; i < n_preds; i++) {
		if (preds[i].cmp > B2A_CMP_GE) {
			return Fail(B2A_EINVAL, "unknown comparison");
		}
		if (preds[i].column >= p.top_columns.size()) {
			return Fail(B2A_EINVAL, "predicate column index out of range");
		}
		const Field &f = *schema.fields[static_cast<size_t>(p.top_columns[preds[i].column])];
		if (f.dict_id >= 0 || f.out_width > 8 || f.out_width == 0 || f.type_id == B2A_T_FLOAT ||
		    f.type_id == B2A_T_DOUBLE || f.type_id == B2A_T_UBIGINT) {
			return Fail(B2A_ENOTIMPL, "filter pushdown supports signed integer-like columns (integers, dates, "
			                          "timestamps, DECIMAL up to 18 digits) only: column '" + f.name + "'");
		}
		flats.push_back(f.flat_index);
		const bool uns = f.type_id == B2A_T_UTINYINT || f.type_id == B2A_T_USMALLINT || f.type_id == B2A_T_UINTEGER;
		l.preds[i] = preds[i];
		l.widths[i] = static_cast<uint8_t>(f.out_width | (uns ? 0x80 : 0));
	}
	std::vector<uint32_t> rows;
	std::vector<uint64_t> table;
	rows.reserve(p.n_chunks);
	table.reserve(p.n_chunks * 2 * n_preds);
	for (uint64_t b = 0; b < p.batch_rows.size(); b++) {
		const uint64_t nchunks = (p.batch_rows[b] + B2A_VECTOR_SIZE - 1) / B2A_VECTOR_SIZE;
		for (uint64_t k = 0; k < nchunks; k++) {
			const uint64_t left = p.batch_rows[b] - k * B2A_VECTOR_SIZE;
			rows.push_back(left > B2A_VECTOR_SIZE ? B2A_VECTOR_SIZE : static_cast<uint32_t>(left));
			for (size_t c = 0; c < flats.size(); c++) {
				const b2a_column_out &co = p.Column(p.batch_lo + b, static_cast<uint32_t>(flats[c]));
				const Field &f = *schema.flat[static_cast<size_t>(flats[c])];
				table.push_back(co.data_off + k * B2A_VECTOR_SIZE * static_cast<uint64_t>(f.out_width));
				table.push_back(co.validity_off == B2A_NONE ? B2A_NONE : co.validity_off + k * (B2A_VECTOR_SIZE / 8));
			}
		}
	}
	if (rows.size() != p.n_chunks) {
		return Fail(B2A_EINTERNAL, "b2a_scan_filter_no_alloc: chunk count mismatch");
	}
	cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
	CUDA_OK(cudaMemsetAsync(d_total, 0, sizeof(unsigned long long), st));
	l.out = static_cast<const uint8_t *>(d_out);
	l.n_chunks = p.n_chunks;
	l.n_preds = n_preds;
	l.d_sel = static_cast<uint32_t *>(d_sel);
	l.d_count = static_cast<uint32_t *>(d_count);
	l.d_total = static_cast<unsigned long long *>(d_total);
	l.sm_count = s->sm_count;
	int e = LaunchFilterNoAlloc(l, table.data(), rows.data(), cuda_stream);
	if (e == -1) {
		return Fail(B2A_EINVAL, "b2a_scan_filter_no_alloc: tables exceed the fixed device scratch");
	}
	if (e != 0) {
		return Fail(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
	}
	if (l.n_chunks) {
		s->launches++;
	}
	return B2A_OK;
	B2A_CATCH
}

// The Q6 shape in one launch: count and sum(a * b) over the rows that pass the predicates, no selection vectors in between.
extern "C" int b2a_scan_filter_sum_product(b2a_scan *s, const void *d_out, const b2a_predicate *preds_in, uint32_t n_preds,
                                           uint32_t col_a, uint32_t col_b, void *d_total, void *d_result_i128,
                                           void *cuda_stream) {
	B2A_TRY
	NEED_DEVICE(s);
	if (n_preds == 0 || n_preds > 8 || !preds_in) {
		return Fail(B2A_EINVAL, "between 1 and 8 predicates are supported");
	}
	if (!d_total || !d_result_i128) {
		return Fail(B2A_EINVAL, "b2a_scan_filter_sum_product: null output");
	}
	std::vector<b2a_predicate> sorted(preds_in, preds_in + n_preds);
	std::stable_sort(sorted.begin(), sorted.end(),
	                 [](const b2a_predicate &a, const b2a_predicate &b) { return a.column < b.column; });
	std::vector<uint32_t> cols;
	for (auto &pr : sorted) {
		cols.push_back(pr.column);
		if (pr.cmp > B2A_CMP_GE) {
			return Fail(B2A_EINVAL, "unknown comparison");
		}
	}
	cols.push_back(col_a);
	cols.push_back(col_b);
	std::vector<uint8_t> widths;
	int rc = ChunkTable(s, d_out, cols, s->chunk_col3_key, &s->d_chunk_col3, &widths);
	if (rc != B2A_OK) {
		return rc;
	}
	if (widths[n_preds] != 8 || widths[n_preds + 1] != 8) {
		return Fail(B2A_ENOTIMPL, "sum_product needs two 8-byte integer columns (BIGINT / DECIMAL(<=18))");
	}
	cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
	CUDA_OK(cudaMemsetAsync(d_total, 0, sizeof(unsigned long long), st));
	CUDA_OK(cudaMemsetAsync(d_result_i128, 0, 16, st));
	FilterLaunch l;
	l.out = static_cast<const uint8_t *>(d_out);
	l.d_chunk_col = s->d_chunk_col3;
	l.d_chunk_rows = s->d_chunk_rows;
	l.n_chunks = s->plan.n_chunks;
	l.n_preds = n_preds;
	for (uint32_t i = 0; i < n_preds; i++) {
		l.preds[i] = sorted[i];
		l.widths[i] = widths[i];
	}
	l.d_total = static_cast<unsigned long long *>(d_total);
	l.sm_count = s->sm_count;
	int e = LaunchFilterSum(l, static_cast<unsigned long long *>(d_result_i128), cuda_stream);
	if (e != 0) {
		return Fail(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
	}
	if (l.n_chunks) {
		s->launches++;
	}
	return B2A_OK;
	B2A_CATCH
}

extern "C" int b2a_scan_sum_product(b2a_scan *s, const void *d_out, uint32_t col_a, uint32_t col_b, const void *d_sel,
                                    const void *d_count, void *d_result_i128, void *cuda_stream) {
	B2A_TRY
	NEED_DEVICE(s);
	std::vector<uint32_t> cols = {col_a, col_b};
	std::vector<uint8_t> widths;
	int rc = ChunkTable(s, d_out, cols, s->chunk_col2_key, &s->d_chunk_col2, &widths);
	if (rc != B2A_OK) {
		return rc;
	}
	if (widths[0] != 8 || widths[1] != 8) {
		return Fail(B2A_ENOTIMPL, "sum_product needs two 8-byte integer columns (BIGINT / DECIMAL(<=18))");
	}
	cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
	CUDA_OK(cudaMemsetAsync(d_result_i128, 0, 16, st));
	SumProductLaunch l;
	l.out = static_cast<const uint8_t *>(d_out);
	l.d_chunk_col = s->d_chunk_col2;
	l.d_chunk_rows = s->d_chunk_rows;
	l.d_sel = static_cast<const uint32_t *>(d_sel);
	l.d_count = static_cast<const uint32_t *>(d_count);
	l.n_chunks = s->plan.n_chunks;
	l.d_result = static_cast<unsigned long long *>(d_result_i128);
	l.sm_count = s->sm_count;
	int e = LaunchSumProduct(l, cuda_stream);
	if (e != 0) {
		return Fail(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
	}
	if (l.n_chunks) {
		s->launches++;
	}
	return B2A_OK;
	B2A_CATCH
}
