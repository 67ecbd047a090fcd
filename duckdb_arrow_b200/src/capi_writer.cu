// C ABI of the writer direction (include/b2arrow.h: b2a_writer_*): device memory plumbing around the write planner
// and the encode kernels.  No CPU fallback: without a CUDA device every compute entry point fails with B2A_ECUDA.
#include "b2arrow.h"
#include "kernels/encode.cuh"
#include "writer/write_plan.hpp"

#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>

using namespace b2a;

namespace b2a {
int SetLastError(int code, const std::string &msg); // capi.cu
}

struct b2a_writer {
	WritePlan plan;
	int device = 0;
	int sm_count = 148;
	std::vector<void *> allocs; // selection vectors, scratch chunks
	b2a_enc_task *d_tasks = nullptr;
	uint32_t *d_tile_task = nullptr;
	size_t cap_tasks = 0, cap_tiles = 0;
	uint32_t n_enc_tasks = 0, n_enc_tiles = 0;
	uint64_t *d_null_counts = nullptr;
	uint32_t *d_err = nullptr;
	uint64_t launches = 0;
	bool measured = false, finished = false;
	bool enc_full = false; // the encode tasks hold an op of b2a_encode_kernel<true> only
};

// B2A_WRITER_TRACE=1: wall-clock phases of b2a_writer_measure on stderr (where pass 1's milliseconds go)
struct PhaseTrace {
	bool on;
	std::chrono::steady_clock::time_point t0;
	PhaseTrace() : on(std::getenv("B2A_WRITER_TRACE") != nullptr), t0(std::chrono::steady_clock::now()) {
	}
	void Mark(const char *what) {
		if (!on) {
			return;
		}
		auto t1 = std::chrono::steady_clock::now();
		std::fprintf(stderr, "[b2a writer] %-28s %8.3f ms\n", what,
		             std::chrono::duration<double, std::milli>(t1 - t0).count());
		t0 = t1;
	}
};

#define W_TRY try {
#define W_CATCH                                                                                                        \
	}                                                                                                                  \
	catch (const Error &e) {                                                                                           \
		return SetLastError(e.code, e.msg);                                                                            \
	}                                                                                                                  \
	catch (const std::bad_alloc &) {                                                                                   \
		return SetLastError(B2A_EOOM, "out of host memory");                                                           \
	}                                                                                                                  \
	catch (const std::exception &e) {                                                                                  \
		return SetLastError(B2A_EINTERNAL, e.what());                                                                  \
	}

#define W_CUDA(expr)                                                                                                   \
	do {                                                                                                               \
		cudaError_t _e = (expr);                                                                                       \
		if (_e != cudaSuccess) {                                                                                       \
			return SetLastError(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " #expr);      \
		}                                                                                                              \
	} while (0)

#define W_NEED_DEVICE(w)                                                                                               \
	do {                                                                                                               \
		if ((w)->device < 0) {                                                                                         \
			return SetLastError(B2A_ECUDA, "writer was planned without a CUDA device; there is no CPU fallback");      \
		}                                                                                                              \
		W_CUDA(cudaSetDevice((w)->device));                                                                            \
	} while (0)

// Device memory of a writer comes from the device's default stream-ordered pool with the release threshold raised once
// per device: a COPY creates one writer per prepared batch collection, and cudaMalloc / cudaFree (0.1-0.5 ms each, six
// per writer) were a third of pass 1 (B2A_WRITER_TRACE).  Pool allocations cost microseconds once the pool is warm.
static void KeepPoolWarm(int device) {
	static std::mutex mu;
	static std::vector<int> done;
	std::lock_guard<std::mutex> lock(mu);
	for (int d : done) {
		if (d == device) {
			return;
		}
	}
	cudaMemPool_t pool;
	if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
		uint64_t threshold = UINT64_MAX;
		cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
	}
	done.push_back(device);
}

static int ReserveTasks(b2a_writer *w, size_t n_tasks, size_t n_tiles, cudaStream_t st) {
	if (n_tasks > w->cap_tasks) {
		if (w->d_tasks) {
			W_CUDA(cudaFreeAsync(w->d_tasks, st));
		}
		w->d_tasks = nullptr;
		w->cap_tasks = 0;
		W_CUDA(cudaMallocAsync(&w->d_tasks, n_tasks * sizeof(b2a_enc_task), st));
		w->cap_tasks = n_tasks;
	}
	if (n_tiles > w->cap_tiles) {
		if (w->d_tile_task) {
			W_CUDA(cudaFreeAsync(w->d_tile_task, st));
		}
		w->d_tile_task = nullptr;
		w->cap_tiles = 0;
		W_CUDA(cudaMallocAsync(&w->d_tile_task, n_tiles * sizeof(uint32_t), st));
		w->cap_tiles = n_tiles;
	}
	return B2A_OK;
}

static int UploadTasks(b2a_writer *w, const std::vector<b2a_enc_task> &tasks, const std::vector<uint32_t> &tiles,
                       cudaStream_t st) {
	int rc = ReserveTasks(w, tasks.size(), tiles.size(), st);
	if (rc != B2A_OK) {
		return rc;
	}
	// the host vectors are pageable and die with the caller's scope: plain synchronous copies, ordered after `st`
	W_CUDA(cudaStreamSynchronize(st));
	if (!tasks.empty()) {
		W_CUDA(cudaMemcpy(w->d_tasks, tasks.data(), tasks.size() * sizeof(b2a_enc_task), cudaMemcpyHostToDevice));
	}
	if (!tiles.empty()) {
		W_CUDA(cudaMemcpy(w->d_tile_task, tiles.data(), tiles.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
	}
	return B2A_OK;
}

static int DeviceAlloc(b2a_writer *w, size_t bytes, void **out, cudaStream_t st) {
	W_CUDA(cudaMallocAsync(out, bytes ? bytes : 8, st));
	w->allocs.push_back(*out);
	return B2A_OK;
}

extern "C" {

int b2a_writer_create(const b2a_vector *columns, uint32_t n_columns, uint64_t rows, const b2a_write_options *opt,
                      int device, b2a_writer **out) {
	W_TRY
	std::unique_ptr<b2a_writer> w(new b2a_writer());
	w->device = device;
	w->plan.Build(columns, n_columns, rows, opt);
	if (device >= 0) {
		W_CUDA(cudaSetDevice(device));
		int sms = 0; // (cudaGetDeviceProperties fills the whole struct: a millisecond per writer)
		W_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
		w->sm_count = sms;
		KeepPoolWarm(device);
		W_CUDA(cudaMallocAsync(&w->d_err, sizeof(uint32_t), 0));
		W_CUDA(cudaMemsetAsync(w->d_err, 0, sizeof(uint32_t), 0));
		W_CUDA(cudaStreamSynchronize(0));
	}
	*out = w.release();
	return B2A_OK;
	W_CATCH
}

void b2a_writer_destroy(b2a_writer *w) {
	if (!w) {
		return;
	}
	if (w->device >= 0) {
		cudaSetDevice(w->device);
		cudaDeviceSynchronize(); // whatever could still read these buffers has finished; the frees go back to the pool
		for (void *p : w->allocs) {
			cudaFreeAsync(p, 0);
		}
		for (void *p : {static_cast<void *>(w->d_tasks), static_cast<void *>(w->d_tile_task),
		                static_cast<void *>(w->d_null_counts), static_cast<void *>(w->d_err)}) {
			if (p) {
				cudaFreeAsync(p, 0);
			}
		}
	}
	delete w;
}

int b2a_writer_measure(b2a_writer *w, void *cuda_stream) {
	W_TRY
	W_NEED_DEVICE(w);
	if (w->measured) {
		return B2A_OK;
	}
	cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
	WritePlan &p = w->plan;
	PhaseTrace trace;
	EncodeLaunch l;
	l.sm_count = w->sm_count;
	l.d_err = w->d_err;
	{ // one allocation for both task uploads of the common case (the encode tasks are the larger set)
		size_t est_tiles = 16;
		for (auto &n : p.nodes) {
			est_tiles += static_cast<size_t>(n.rows / B2A_ENC_STR_TILE + 3);
		}
		int rc = ReserveTasks(w, 2 * p.nodes.size() + 16, est_tiles, st);
		if (rc != B2A_OK) {
			return rc;
		}
	}
	// ---- UNION members: mask = union row valid AND tag == member AND member value valid, over the input rows
	for (size_t i = 0; i < p.vecs.size(); i++) {
		if (p.vecs[i].union_member < 0 || p.vecs[i].union_mask_ready) {
			continue;
		}
		const WVec &u = p.vecs[static_cast<size_t>(p.vecs[i].union_vec)];
		void *mask = nullptr;
		int rc = DeviceAlloc(w, ((u.rows + 63) / 64 + 1) * 8, &mask, st);
		if (rc != B2A_OK) {
			return rc;
		}
		int e = LaunchEncUnionMask(reinterpret_cast<const void *>(u.data), reinterpret_cast<const void *>(u.validity),
		                           reinterpret_cast<const void *>(p.vecs[i].validity),
		                           static_cast<uint32_t>(p.vecs[i].union_member), u.rows, mask, w->sm_count, st);
		if (e != 0) {
			return SetLastError(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
		}
		if (u.rows) {
			w->launches++;
		}
		p.vecs[i].validity = reinterpret_cast<uint64_t>(mask);
		p.vecs[i].union_mask_ready = true;
	}
	std::vector<b2a_enc_task> enc_tasks;
	std::vector<uint32_t> enc_tiles;
	bool skeleton_ready = false;
	for (;;) {
		// ---- fixed-size lists under a selection: child selection = sel[i / k] * k + i % k
		for (;;) {
			std::vector<b2a_enc_task> tasks;
			std::vector<uint32_t> tiles;
			const size_t n_before = p.nodes.size();
			for (size_t i = 0; i < n_before; i++) {
				if (!p.nodes[i].needs_expand_array || !p.nodes[i].children.empty()) {
					continue;
				}
				const WVec v = p.vecs[static_cast<size_t>(p.nodes[i].vec)];
				const uint64_t k = static_cast<uint64_t>(v.param), child_rows = p.nodes[i].rows * k;
				void *sel = nullptr;
				int rc = DeviceAlloc(w, child_rows * 4, &sel, st);
				if (rc != B2A_OK) {
					return rc;
				}
				b2a_enc_task t;
				std::memset(&t, 0, sizeof(t));
				t.op = B2A_ENC_EXPAND_ARRAY;
				t.sel = p.nodes[i].sel;
				t.rows = child_rows;
				t.param = v.param;
				t.dst = reinterpret_cast<uint64_t>(sel);
				t.tile_rows = 16384;
				t.tile0 = static_cast<uint32_t>(tiles.size());
				tiles.insert(tiles.end(), (child_rows + t.tile_rows - 1) / t.tile_rows, static_cast<uint32_t>(tasks.size()));
				tasks.push_back(t);
				int child = p.AddNode(v.children.at(0), child_rows, 0, reinterpret_cast<uint64_t>(sel));
				p.nodes[i].children.push_back(child);
			}
			if (tasks.empty()) {
				break;
			}
			int rc = UploadTasks(w, tasks, tiles, st);
			if (rc != B2A_OK) {
				return rc;
			}
			l.d_tasks = w->d_tasks;
			l.d_tile_task = w->d_tile_task;
			l.n_tasks = static_cast<uint32_t>(tasks.size());
			l.n_tiles = static_cast<uint32_t>(tiles.size());
			if (l.n_tiles) {
				int e = LaunchEncExpand(l, st);
				if (e != 0) {
					return SetLastError(B2A_ECUDA, std::string("CUDA error: ") +
					                                   cudaGetErrorString(static_cast<cudaError_t>(e)));
				}
				w->launches++;
			}
		}
		// ---- strings and lists: per-tile sums, per-task scan, totals back to the host
		std::vector<int> pending = p.PendingScanNodes();
		if (pending.empty()) {
			break;
		}
		uint64_t n_tiles_total = 0;
		for (int ni : pending) {
			n_tiles_total += WritePlan::ScanTiles(p.nodes[static_cast<size_t>(ni)].rows);
		}
		void *chunk = nullptr;
		int rc = DeviceAlloc(w, (n_tiles_total + pending.size() + 1) * 8, &chunk, st);
		if (rc != B2A_OK) {
			return rc;
		}
		uint64_t *base = static_cast<uint64_t *>(chunk);
		uint64_t *d_heap_bytes = base + n_tiles_total + pending.size();
		W_CUDA(cudaMemsetAsync(d_heap_bytes, 0, 8, st));
		uint64_t cursor = 0;
		for (size_t j = 0; j < pending.size(); j++) {
			WNode &n = p.nodes[static_cast<size_t>(pending[j])];
			n.scratch = reinterpret_cast<uint64_t>(base + cursor);
			n.counter = reinterpret_cast<uint64_t>(base + n_tiles_total + j);
			cursor += WritePlan::ScanTiles(n.rows);
		}
		std::vector<b2a_enc_task> tasks;
		std::vector<uint32_t> tiles;
		trace.Mark("scratch alloc");
		p.ScanTasks(pending, tasks, tiles);
		trace.Mark("scan tasks (host)");
		rc = UploadTasks(w, tasks, tiles, st);
		if (rc != B2A_OK) {
			return rc;
		}
		trace.Mark("scan tasks upload");
		l.d_tasks = w->d_tasks;
		l.d_tile_task = w->d_tile_task;
		l.n_tasks = static_cast<uint32_t>(tasks.size());
		l.n_tiles = static_cast<uint32_t>(tiles.size());
		l.d_heap_bytes = d_heap_bytes;
		int e = LaunchEncMeasure(l, st);
		if (e != 0) {
			return SetLastError(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
		}
		w->launches += 2;
		// While the measure kernels read the string_t records: if no list is pending the node set is final, and
		// everything about the encode tasks except their stream offsets can be built now (the D2H copy of the totals
		// below returns when the kernels are done).
		if (!skeleton_ready) {
			bool final_nodes = true;
			for (int ni : pending) {
				final_nodes = final_nodes && !WritePlan::IsListOp(p.vecs[static_cast<size_t>(p.nodes[static_cast<size_t>(ni)].vec)].op);
			}
			if (final_nodes) {
				p.BuildPreorders();
				p.EncodeTaskSkeleton(enc_tasks, enc_tiles);
				skeleton_ready = true;
				trace.Mark("encode task skeleton (host, overlapped)");
			}
		}
		std::vector<uint64_t> totals(pending.size() + 1);
		W_CUDA(cudaMemcpyAsync(totals.data(), base + n_tiles_total, (pending.size() + 1) * 8, cudaMemcpyDeviceToHost, st));
		W_CUDA(cudaStreamSynchronize(st));
		p.heap_bytes += totals[pending.size()];
		trace.Mark("measure kernels + totals");
		// ---- lists: child selection vectors (ArrowListData::AppendOffsets child_sel), children become new nodes
		std::vector<b2a_enc_task> xtasks;
		std::vector<uint32_t> xtiles;
		for (size_t j = 0; j < pending.size(); j++) {
			const int ni = pending[j];
			p.nodes[static_cast<size_t>(ni)].total = totals[j];
			p.nodes[static_cast<size_t>(ni)].measured = true;
			const WVec v = p.vecs[static_cast<size_t>(p.nodes[static_cast<size_t>(ni)].vec)];
			if (!WritePlan::IsListOp(v.op)) {
				continue;
			}
			const WVec &cv = p.vecs[static_cast<size_t>(v.children.at(0))];
			if (cv.rows > 0xFFFFFFFFull) {
				return SetLastError(B2A_EINVAL, "writer: list child vector exceeds the sel_t (uint32) range");
			}
			void *sel = nullptr;
			rc = DeviceAlloc(w, totals[j] * 4, &sel, st);
			if (rc != B2A_OK) {
				return rc;
			}
			b2a_enc_task t = tasks[j];
			t.dst = reinterpret_cast<uint64_t>(sel);
			t.tile0 = static_cast<uint32_t>(xtiles.size());
			xtiles.insert(xtiles.end(), WritePlan::ScanTiles(t.rows), static_cast<uint32_t>(xtasks.size()));
			xtasks.push_back(t);
			int child = p.AddNode(v.children.at(0), totals[j], 0, reinterpret_cast<uint64_t>(sel));
			p.nodes[static_cast<size_t>(ni)].children.push_back(child);
		}
		if (!xtasks.empty()) {
			rc = UploadTasks(w, xtasks, xtiles, st);
			if (rc != B2A_OK) {
				return rc;
			}
			l.d_tasks = w->d_tasks;
			l.d_tile_task = w->d_tile_task;
			l.n_tasks = static_cast<uint32_t>(xtasks.size());
			l.n_tiles = static_cast<uint32_t>(xtiles.size());
			e = LaunchEncExpand(l, st);
			if (e != 0) {
				return SetLastError(B2A_ECUDA,
				                    std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)));
			}
			w->launches++;
		}
	}
	// ---- everything is sized: lay the stream out and stage the encode tasks
	trace.Mark("list expansion");
	p.Layout();
	trace.Mark("layout (host)");
	if (w->d_null_counts) {
		W_CUDA(cudaFreeAsync(w->d_null_counts, st));
	}
	w->d_null_counts = nullptr;
	W_CUDA(cudaMallocAsync(&w->d_null_counts, (p.nodes.size() ? p.nodes.size() : 1) * 8, st));
	for (size_t i = 0; i < p.nodes.size(); i++) {
		p.nodes[i].null_counter = reinterpret_cast<uint64_t>(w->d_null_counts + i);
	}
	std::vector<b2a_enc_task> &tasks = enc_tasks;
	std::vector<uint32_t> &tiles = enc_tiles;
	trace.Mark("null count alloc");
	if (skeleton_ready) {
		p.PatchEncodeTasks(tasks);
	} else {
		p.EncodeTasks(tasks, tiles);
	}
	trace.Mark("encode tasks (host)");
	int rc = UploadTasks(w, tasks, tiles, st);
	if (rc != B2A_OK) {
		return rc;
	}
	trace.Mark("encode tasks upload");
	w->n_enc_tasks = static_cast<uint32_t>(tasks.size());
	w->n_enc_tiles = static_cast<uint32_t>(tiles.size());
	w->enc_full = false;
	for (auto &t : tasks) {
		w->enc_full = w->enc_full || B2A_ENC_OP_IS_RARE(t.op);
	}
	w->measured = true;
	return B2A_OK;
	W_CATCH
}

int b2a_writer_sizes(b2a_writer *w, uint64_t *stream_bytes, uint64_t *n_batches, uint64_t *body_bytes,
                     uint64_t *algorithmic_bytes) {
	if (!w->plan.laid_out) {
		return SetLastError(B2A_EINVAL, "writer: call b2a_writer_measure first");
	}
	*stream_bytes = w->plan.stream_bytes;
	*n_batches = w->plan.batches.size();
	*body_bytes = w->plan.body_bytes;
	*algorithmic_bytes = w->plan.algorithmic_in + w->plan.algorithmic_out;
	return B2A_OK;
}

int b2a_writer_encode(b2a_writer *w, void *d_stream, void *cuda_stream) {
	W_TRY
	W_NEED_DEVICE(w);
	if (!w->measured) {
		return SetLastError(B2A_EINVAL, "writer: call b2a_writer_measure first");
	}
	if (!d_stream || (reinterpret_cast<uintptr_t>(d_stream) & 255) != 0) {
		return SetLastError(B2A_EINVAL, "writer: the stream buffer must be a 256-byte aligned device pointer");
	}
	cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
	W_CUDA(cudaMemsetAsync(w->d_null_counts, 0, (w->plan.nodes.size() ? w->plan.nodes.size() : 1) * 8, st));
	EncodeLaunch l;
	l.sm_count = w->sm_count;
	l.d_err = w->d_err;
	l.d_tasks = w->d_tasks;
	l.d_tile_task = w->d_tile_task;
	l.n_tasks = w->n_enc_tasks;
	l.n_tiles = w->n_enc_tiles;
	l.stream_out = static_cast<uint8_t *>(d_stream);
	l.full = w->enc_full;
	int e = LaunchEncode(l, st);
	if (e != 0) {
		return SetLastError(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)) +
		                                   " launching b2a_encode_kernel");
	}
	if (l.n_tiles) {
		w->launches++;
	}
	w->finished = false;
	return B2A_OK;
	W_CATCH
}

int b2a_writer_finish(b2a_writer *w, void *cuda_stream) {
	W_TRY
	W_NEED_DEVICE(w);
	cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
	WritePlan &p = w->plan;
	std::vector<uint64_t> nulls(p.nodes.size() ? p.nodes.size() : 1, 0);
	uint32_t err = 0;
	W_CUDA(cudaMemcpyAsync(nulls.data(), w->d_null_counts, nulls.size() * 8, cudaMemcpyDeviceToHost, st));
	W_CUDA(cudaMemcpyAsync(&err, w->d_err, sizeof(err), cudaMemcpyDeviceToHost, st));
	W_CUDA(cudaStreamSynchronize(st));
	if (err != 0) {
		W_CUDA(cudaMemsetAsync(w->d_err, 0, sizeof(uint32_t), st));
		switch (err) {
		case B2A_DEV_ENC_OFFSET_I32:
			return SetLastError(B2A_EINVAL, "Arrow Appender: The maximum total string size for regular string buffers "
			                                "is 2147483647 but the offset exceeds this.\n* SET "
			                                "arrow_large_buffer_size=true to use large string buffers");
		case B2A_DEV_ENC_LIST_I32:
			return SetLastError(B2A_EINVAL, "Arrow Appender: The maximum combined list offset for regular list buffers "
			                                "is 2147483647 but the offset exceeds this.\n* SET "
			                                "arrow_large_buffer_size=true to use large list buffers");
		default:
			return SetLastError(B2A_EINTERNAL, "unknown device error code " + std::to_string(err));
		}
	}
	for (size_t i = 0; i < p.nodes.size(); i++) {
		const WVec &v = p.vecs[static_cast<size_t>(p.nodes[i].vec)];
		p.nodes[i].null_count = v.type_id == B2A_T_SQLNULL ? static_cast<int64_t>(p.nodes[i].rows)
		                                                   : static_cast<int64_t>(nulls[i]);
	}
	// "Arrow doesn't accept NULL keys on Maps" (map_data.hpp:97-100)
	for (size_t i = 0; i < p.nodes.size(); i++) {
		const WVec &v = p.vecs[static_cast<size_t>(p.nodes[i].vec)];
		if (v.type_id == B2A_T_MAP && !p.nodes[i].children.empty()) {
			const WNode &entries = p.nodes[static_cast<size_t>(p.nodes[i].children[0])];
			if (!entries.children.empty() && p.nodes[static_cast<size_t>(entries.children[0])].null_count > 0) {
				return SetLastError(B2A_EINVAL, "Arrow doesn't accept NULL keys on Maps");
			}
		}
	}
	w->finished = true;
	return B2A_OK;
	W_CATCH
}

int b2a_writer_headers(b2a_writer *w, void *h_stream) {
	W_TRY
	if (!w->finished) {
		return SetLastError(B2A_EINVAL, "writer: call b2a_writer_finish first (null counts are part of the headers)");
	}
	const WritePlan &p = w->plan;
	uint8_t *h = static_cast<uint8_t *>(h_stream);
	std::memcpy(h, p.schema_message.data(), p.schema_message.size());
	for (size_t bi = 0; bi < p.dict_batches.size() + p.batches.size(); bi++) {
		const bool is_dict = bi < p.dict_batches.size();
		const WBatch &b = is_dict ? p.dict_batches[bi] : p.batches[bi - p.dict_batches.size()];
		std::vector<NodeSlice> nodes;
		std::vector<BufferSlice> buffers;
		for (int ni : b.preorder) {
			const WNode &n = p.nodes[static_cast<size_t>(ni)];
			const WVec &v = p.vecs[static_cast<size_t>(n.vec)];
			nodes.push_back(NodeSlice {static_cast<int64_t>(n.rows), n.null_count});
			for (int k = 0; k < v.n_buffers; k++) {
				buffers.push_back(n.buf[k]);
			}
		}
		std::vector<uint8_t> msg = is_dict
		                               ? EncodeDictionaryBatchMessage(b.dictionary_id, b.rows, nodes, buffers, b.body_length)
		                               : EncodeRecordBatchMessage(b.rows, nodes, buffers, b.body_length, p.n_variadic);
		if (msg.size() != (is_dict ? p.dict_header_bytes : p.batch_header_bytes)) {
			return SetLastError(B2A_EINTERNAL, "writer: record batch header size changed after layout");
		}
		std::memcpy(h + b.meta_offset, msg.data(), msg.size());
	}
	std::vector<uint8_t> eos = EncodeEndOfStream();
	std::memcpy(h + p.stream_bytes - eos.size(), eos.data(), eos.size());
	return B2A_OK;
	W_CATCH
}

int b2a_writer_batch(b2a_writer *w, uint64_t i, b2a_batch_info *out) {
	const WritePlan &p = w->plan;
	if (!p.laid_out || i >= p.batches.size()) {
		return SetLastError(B2A_EINVAL, "writer: batch index out of range (or b2a_writer_measure not called)");
	}
	std::memset(out, 0, sizeof(*out));
	out->rows = p.batches[i].rows;
	out->meta_offset = p.batches[i].meta_offset;
	out->body_offset = p.batches[i].body_offset;
	out->body_length = p.batches[i].body_length;
	out->buffer_bytes = p.batches[i].body_length;
	return B2A_OK;
}

int b2a_writer_preamble(b2a_writer *w, uint64_t *bytes) {
	if (!w || !bytes) {
		return SetLastError(B2A_EINVAL, "writer: null argument");
	}
	if (!w->plan.laid_out) {
		return SetLastError(B2A_EINVAL, "writer: call b2a_writer_measure first");
	}
	*bytes = w->plan.preamble_bytes;
	return B2A_OK;
}

uint64_t b2a_writer_gpu_launches(b2a_writer *w) {
	return w->launches;
}

int b2a_writer_schema_message(b2a_writer *w, const void **data, uint64_t *size) {
	if (!w || !data || !size) {
		return SetLastError(B2A_EINVAL, "writer: null argument");
	}
	*data = w->plan.schema_message.data();
	*size = w->plan.schema_message.size();
	return B2A_OK;
}

} // extern "C"
