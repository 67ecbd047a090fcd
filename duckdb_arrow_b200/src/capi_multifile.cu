// C ABI of the multi-file scan (include/b2arrow.h: b2a_multifile_*): binding on the host (file_scanner/multi_file.cpp),
// column casts on the device (kernels/cast.cu).  No CPU fallback: b2a_multifile_finalize_run needs a CUDA device.
#include "capi_internal.hpp"
#include "file_scanner/multi_file.hpp"
#include "kernels/cast.cuh"

#include <cuda_runtime.h>

#include <cstring>
#include <memory>

using namespace b2a;

struct b2a_multifile {
	MultiFile mf;
};

struct b2a_mf_finalize {
	b2a_multifile *m = nullptr;
	b2a_scan *scan = nullptr;
	uint32_t file = 0;
	uint64_t final_bytes = 0;
	std::vector<b2a_cast_task> tasks;
	std::vector<uint32_t> tile_task;
	// [(batch - batch_lo) * n_columns + global column] -> cast result location (data_off / chars_off in the FINAL arena)
	std::vector<b2a_column_out> cols;
	b2a_cast_task *d_tasks = nullptr;
	uint32_t *d_tile_task = nullptr;
	uint32_t *d_err = nullptr;
	uint64_t launches = 0;
};

#define M_TRY try {
#define M_CATCH                                                                                                        \
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

#define M_CUDA(expr)                                                                                                   \
	do {                                                                                                               \
		cudaError_t _e = (expr);                                                                                       \
		if (_e != cudaSuccess) {                                                                                       \
			return SetLastError(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " #expr);      \
		}                                                                                                              \
	} while (0)

static void CopyName(char *dst, size_t cap, const std::string &s) {
	size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
	std::memcpy(dst, s.data(), n);
	dst[n] = 0;
}

extern "C" {

int b2a_multifile_open(const char *const *paths, uint32_t n_paths, const b2a_multifile_options *opt,
                       b2a_multifile **out) {
	M_TRY
	std::vector<std::string> patterns;
	for (uint32_t i = 0; i < n_paths; i++) {
		patterns.emplace_back(paths[i]);
	}
	MfOptions o;
	if (opt) {
		o.union_by_name = opt->union_by_name != 0;
		o.filename = opt->filename != 0;
		o.hive_partitioning = opt->hive_partitioning != 0;
		o.hive_types_autocast = opt->hive_types_autocast != 0;
		if (opt->filename_column && opt->filename_column[0]) {
			o.filename_column = opt->filename_column;
		}
	}
	std::unique_ptr<b2a_multifile> m(new b2a_multifile());
	m->mf.Open(patterns, o);
	*out = m.release();
	return B2A_OK;
	M_CATCH
}

void b2a_multifile_close(b2a_multifile *m) {
	delete m;
}

int b2a_multifile_counts(b2a_multifile *m, uint32_t *n_files, uint32_t *n_columns, uint32_t *n_projected) {
	*n_files = static_cast<uint32_t>(m->mf.files.size());
	*n_columns = static_cast<uint32_t>(m->mf.columns.size());
	*n_projected = static_cast<uint32_t>(m->mf.projection.size());
	return B2A_OK;
}

int b2a_multifile_column(b2a_multifile *m, uint32_t column, b2a_mf_column_info *out) {
	if (column >= m->mf.columns.size()) {
		return SetLastError(B2A_EINVAL, "column index out of range");
	}
	const MfColumn &c = m->mf.columns[column];
	std::memset(out, 0, sizeof(*out));
	CopyName(out->name, sizeof(out->name), c.name);
	CopyName(out->type_name, sizeof(out->type_name), c.type.name);
	out->type_id = c.type.type_id;
	out->decimal_width = c.type.decimal_width;
	out->decimal_scale = c.type.decimal_scale;
	out->kind = c.kind;
	return B2A_OK;
}

int b2a_multifile_projected(b2a_multifile *m, uint32_t i, uint32_t *column) {
	if (i >= m->mf.projection.size()) {
		return SetLastError(B2A_EINVAL, "projection index out of range");
	}
	*column = static_cast<uint32_t>(m->mf.projection[i]);
	return B2A_OK;
}

int b2a_multifile_set_projection(b2a_multifile *m, const char *const *names, uint32_t n) {
	M_TRY
	std::vector<std::string> v;
	for (uint32_t i = 0; i < n; i++) {
		v.emplace_back(names[i]);
	}
	m->mf.SetProjection(v);
	return B2A_OK;
	M_CATCH
}

int b2a_multifile_file(b2a_multifile *m, uint32_t file, const char **path, b2a_reader **reader) {
	if (file >= m->mf.files.size()) {
		return SetLastError(B2A_EINVAL, "file index out of range");
	}
	*path = m->mf.files[file].path.c_str();
	*reader = m->mf.files[file].reader;
	return B2A_OK;
}

int b2a_multifile_map(b2a_multifile *m, uint32_t file, uint32_t column, b2a_mf_map_info *out) {
	if (file >= m->mf.files.size() || column >= m->mf.columns.size()) {
		return SetLastError(B2A_EINVAL, "file / column index out of range");
	}
	const MfMap &mp = m->mf.files[file].maps[column];
	std::memset(out, 0, sizeof(*out));
	out->kind = mp.kind;
	out->local_flat = mp.local_flat;
	out->from_type_id = mp.from.type_id;
	out->to_type_id = m->mf.columns[column].type.type_id;
	out->constant_is_null = mp.constant_is_null;
	out->constant = mp.constant.c_str();
	return B2A_OK;
}

// ------------------------------------------------------------------ finalize: the cast expressions of one file
int b2a_multifile_finalize_create(b2a_multifile *m, uint32_t file, b2a_scan *scan, b2a_mf_finalize **out) {
	M_TRY
	if (file >= m->mf.files.size()) {
		return SetLastError(B2A_EINVAL, "file index out of range");
	}
	const MfFile &f = m->mf.files[file];
	if (scan->r != f.reader) {
		return SetLastError(B2A_EINVAL, "the scan does not belong to this file of the multi-file scan");
	}
	if (scan->plan.alias_direct) {
		return SetLastError(B2A_EINVAL, "multi-file casts need device-resident vectors: create the scan without "
		                                "B2A_SCAN_HOST_ALIAS");
	}
	std::unique_ptr<b2a_mf_finalize> fin(new b2a_mf_finalize());
	fin->m = m;
	fin->scan = scan;
	fin->file = file;
	const ScanPlan &p = scan->plan;
	const size_t ncol = m->mf.columns.size();
	const uint64_t nb = p.batch_hi - p.batch_lo;
	b2a_column_out none;
	std::memset(&none, 0xFF, sizeof(none));
	fin->cols.assign(nb * ncol, none);
	uint64_t off = 0;
	auto alloc = [&](uint64_t bytes) {
		uint64_t o = off;
		off += (bytes + 255) & ~255ull;
		return o;
	};
	uint32_t tile = 0;
	for (int c : m->mf.projection) {
		const MfMap &mp = f.maps[static_cast<size_t>(c)];
		if (mp.kind != MF_MAP_CAST) {
			continue;
		}
		const MfType &to = m->mf.columns[static_cast<size_t>(c)].type;
		const bool from_decimal = mp.from.type_id == B2A_T_DECIMAL;
		const int sw = from_decimal ? DecimalPhysicalWidth(mp.from.decimal_width) : TypeWidth(mp.from.type_id);
		const bool to_decimal = to.type_id == B2A_T_DECIMAL;
		const int dw = to_decimal ? DecimalPhysicalWidth(to.decimal_width) : TypeWidth(to.type_id);
		// VARCHAR results: a fixed slot per row, sized for the longest text of the source type (kernels/cast.cuh)
		const uint64_t stride = (mp.from.type_id == B2A_T_TIMESTAMP || mp.from.type_id == B2A_T_HUGEINT ||
		                         (from_decimal && mp.from.decimal_width > 18))
		                            ? B2A_CAST_HEAP_STRIDE_WIDE
		                            : B2A_CAST_HEAP_STRIDE;
		for (uint64_t b = 0; b < nb; b++) {
			const b2a_column_out &src = p.Column(p.batch_lo + b, static_cast<uint32_t>(mp.local_flat));
			b2a_column_out &dst = fin->cols[b * ncol + static_cast<size_t>(c)];
			dst = none;
			dst.rows = src.rows;
			dst.null_count = src.null_count;
			dst.validity_off = src.validity_off; // the mask is shared with the source vector (scan arena)
			dst.stats_off = src.stats_off;
			dst.data_off = alloc(src.rows * static_cast<uint64_t>(dw));
			b2a_cast_task t;
			std::memset(&t, 0, sizeof(t));
			t.src = src.data_off;
			t.validity = src.validity_off;
			t.dst = dst.data_off;
			t.heap = B2A_NONE;
			if (to.type_id == B2A_T_VARCHAR) {
				dst.chars_off = alloc(src.rows * stride);
				dst.chars_len = src.rows * stride;
				t.heap = dst.chars_off;
				t.heap_stride = static_cast<uint8_t>(stride);
			}
			if (from_decimal) {
				t.decimal_width = static_cast<uint8_t>(mp.from.decimal_width);
				t.decimal_scale = static_cast<uint8_t>(mp.from.decimal_scale);
			}
			if (to_decimal) {
				t.dst_decimal_width = static_cast<uint8_t>(to.decimal_width);
				t.dst_decimal_scale = static_cast<uint8_t>(to.decimal_scale);
			}
			t.rows = src.rows;
			t.tile0 = tile;
			t.from = static_cast<uint16_t>(mp.from.type_id);
			t.to = static_cast<uint16_t>(to.type_id);
			t.src_width = static_cast<uint8_t>(sw);
			t.dst_width = static_cast<uint8_t>(dw);
			const uint32_t ntiles = static_cast<uint32_t>((src.rows + B2A_VECTOR_SIZE - 1) / B2A_VECTOR_SIZE);
			for (uint32_t k = 0; k < ntiles; k++) {
				fin->tile_task.push_back(static_cast<uint32_t>(fin->tasks.size()));
			}
			tile += ntiles;
			fin->tasks.push_back(t);
		}
	}
	fin->final_bytes = off;
	*out = fin.release();
	return B2A_OK;
	M_CATCH
}

void b2a_multifile_finalize_destroy(b2a_mf_finalize *fin) {
	if (!fin) {
		return;
	}
	cudaFree(fin->d_tasks);
	cudaFree(fin->d_tile_task);
	cudaFree(fin->d_err);
	delete fin;
}

int b2a_multifile_finalize_sizes(b2a_mf_finalize *fin, uint64_t *final_bytes, uint64_t *n_casts) {
	*final_bytes = fin->final_bytes;
	*n_casts = fin->tasks.size();
	return B2A_OK;
}

int b2a_multifile_finalize_run(b2a_mf_finalize *fin, const void *d_out, void *d_final, uint64_t src_string_ptr_base,
                               uint64_t dst_string_ptr_base, void *cuda_stream) {
	M_TRY
	b2a_scan *s = fin->scan;
	if (s->device < 0) {
		return SetLastError(B2A_ECUDA, "scan was planned without a CUDA device; there is no CPU fallback");
	}
	if (fin->tasks.empty()) {
		return B2A_OK;
	}
	M_CUDA(cudaSetDevice(s->device));
	cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
	if (!fin->d_tasks) {
		M_CUDA(cudaMalloc(&fin->d_tasks, fin->tasks.size() * sizeof(b2a_cast_task)));
		M_CUDA(cudaMalloc(&fin->d_tile_task, (fin->tile_task.size() + 1) * sizeof(uint32_t)));
		M_CUDA(cudaMalloc(&fin->d_err, sizeof(uint32_t)));
		M_CUDA(cudaMemcpy(fin->d_tasks, fin->tasks.data(), fin->tasks.size() * sizeof(b2a_cast_task),
		                  cudaMemcpyHostToDevice));
		M_CUDA(cudaMemcpy(fin->d_tile_task, fin->tile_task.data(), fin->tile_task.size() * sizeof(uint32_t),
		                  cudaMemcpyHostToDevice));
		M_CUDA(cudaMemset(fin->d_err, 0, sizeof(uint32_t)));
	}
	CastLaunch l;
	l.d_tasks = fin->d_tasks;
	l.d_tile_task = fin->d_tile_task;
	l.n_tiles = static_cast<uint32_t>(fin->tile_task.size());
	l.scan_arena = static_cast<const uint8_t *>(d_out);
	l.final_arena = static_cast<uint8_t *>(d_final);
	// string_t pointers of the scan were emitted relative to src_string_ptr_base (0: device image addresses)
	const uint64_t dev_base = reinterpret_cast<uint64_t>(s->d_image);
	l.src_ptr_delta = src_string_ptr_base ? static_cast<int64_t>(dev_base) - static_cast<int64_t>(src_string_ptr_base) : 0;
	l.dst_ptr_origin = dst_string_ptr_base ? dst_string_ptr_base : reinterpret_cast<uint64_t>(d_final);
	l.d_err = fin->d_err;
	l.sm_count = s->sm_count;
	int e = LaunchCast(l, st);
	if (e != 0) {
		return SetLastError(B2A_ECUDA, std::string("CUDA error: ") + cudaGetErrorString(static_cast<cudaError_t>(e)) +
		                                   " launching b2a_cast_kernel");
	}
	fin->launches++;
	return B2A_OK;
	M_CATCH
}

int b2a_multifile_finalize_finish(b2a_mf_finalize *fin, void *cuda_stream) {
	M_TRY
	if (!fin->d_err) {
		return B2A_OK;
	}
	cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
	uint32_t err = 0;
	M_CUDA(cudaMemcpyAsync(&err, fin->d_err, sizeof(err), cudaMemcpyDeviceToHost, st));
	M_CUDA(cudaStreamSynchronize(st));
	if (err != B2A_CAST_DEV_OK) {
		M_CUDA(cudaMemsetAsync(fin->d_err, 0, sizeof(uint32_t), st));
		const std::string &path = fin->m->mf.files[fin->file].path;
		switch (err) {
		case B2A_CAST_DEV_RANGE:
			return SetLastError(B2A_ECONVERSION, "file \"" + path + "\": a value is out of range for the destination "
			                                                        "type of its column cast");
		case B2A_CAST_DEV_PARSE:
			return SetLastError(B2A_ECONVERSION, "file \"" + path + "\": could not convert a string to the "
			                                                        "destination type of its column cast");
		default:
			return SetLastError(B2A_ENOTIMPL, "file \"" + path + "\": VARCHAR -> DOUBLE cast of a string on which the "
			                                                     "reference cast reads past the end of its input ('_' "
			                                                     "separators in a long integer) is not implemented on the device");
		}
	}
	return B2A_OK;
	M_CATCH
}

int b2a_multifile_finalize_column(b2a_mf_finalize *fin, uint64_t batch, uint32_t column, b2a_column_out *out) {
	const ScanPlan &p = fin->scan->plan;
	const size_t ncol = fin->m->mf.columns.size();
	if (batch < p.batch_lo || batch >= p.batch_hi || column >= ncol) {
		return SetLastError(B2A_EINVAL, "batch / column out of range");
	}
	*out = fin->cols[(batch - p.batch_lo) * ncol + column];
	return B2A_OK;
}

uint64_t b2a_multifile_finalize_gpu_launches(b2a_mf_finalize *fin) {
	return fin->launches;
}

} // extern "C"
