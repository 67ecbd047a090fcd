// C ABI, writer side with HOST vectors in and HOST stream bytes out — the twin of b2a_read_arrow_host.
//
// What a DuckDB COPY ... (FORMAT ARROWS) sink holds is host memory: the flat vectors of one row group (the reference
// concatenates the ColumnDataCollection into one DataChunk first, src/writer/column_data_collection_serializer.cpp:97-115,
// then runs ArrowAppender over it).  b2a_write_arrow_host takes exactly that: the vector tree is uploaded, the
// conversion is the b2a_writer kernels (src/kernels/encode.cu), and the finished stream comes back in pinned memory.
// No byte of the conversion is done on the host; the ONE host pass is over string_t records whose payload lives in
// DuckDB's string heap (pointers the device cannot follow): the payloads of strings longer than 12 bytes are packed
// into a pinned heap by host threads and the records re-pointed at its device twin.
#include "capi_internal.hpp"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

using namespace b2a;

namespace b2a {
int SetLastError(int code, const std::string &msg); // capi.cu
}

namespace {

struct HostWrite {
	int device = 0;
	std::vector<void *> d_allocs;  // device copies of the vector tree
	std::vector<void *> h_allocs;  // pinned staging (string records + heaps)
	std::vector<std::unique_ptr<b2a_vector[]>> child_arrays;
	b2a_writer *writer = nullptr;
	void *d_stream = nullptr;
	void *h_stream = nullptr;
	cudaStream_t st = nullptr;

	~HostWrite() {
		if (device >= 0) {
			cudaSetDevice(device);
		}
		if (writer) {
			b2a_writer_destroy(writer);
		}
		for (void *p : d_allocs) {
			cudaFree(p);
		}
		if (d_stream) {
			cudaFree(d_stream);
		}
		for (void *p : h_allocs) {
			cudaFreeHost(p);
		}
		if (h_stream) {
			cudaFreeHost(h_stream);
		}
		if (st) {
			cudaStreamDestroy(st);
		}
	}
};

struct CudaFail {
	cudaError_t e;
};
#define HW_CUDA(x)                                                                                                     \
	do {                                                                                                               \
		cudaError_t e__ = (x);                                                                                         \
		if (e__ != cudaSuccess) {                                                                                      \
			throw CudaFail {e__};                                                                                      \
		}                                                                                                              \
	} while (0)

int FixedWidth(const b2a_vector &v) {
	switch (v.type_id) {
	case B2A_T_BOOLEAN:
	case B2A_T_TINYINT:
	case B2A_T_UTINYINT:
	case B2A_T_UNION: // the UTINYINT tags
		return 1;
	case B2A_T_SMALLINT:
	case B2A_T_USMALLINT:
		return 2;
	case B2A_T_INTEGER:
	case B2A_T_UINTEGER:
	case B2A_T_FLOAT:
	case B2A_T_DATE:
		return 4;
	case B2A_T_BIGINT:
	case B2A_T_UBIGINT:
	case B2A_T_DOUBLE:
	case B2A_T_TIME:
	case B2A_T_TIME_TZ:
	case B2A_T_TIMESTAMP:
	case B2A_T_TIMESTAMP_SEC:
	case B2A_T_TIMESTAMP_MS:
	case B2A_T_TIMESTAMP_NS:
	case B2A_T_TIMESTAMP_TZ:
		return 8;
	case B2A_T_HUGEINT:
	case B2A_T_UHUGEINT:
	case B2A_T_UUID:
	case B2A_T_INTERVAL:
	case B2A_T_LIST: // list_entry_t
	case B2A_T_MAP:
	case B2A_T_VARCHAR: // string_t
	case B2A_T_BLOB:
	case B2A_T_BIT:
	case B2A_T_VARINT:
	case B2A_T_JSON:
		return 16;
	case B2A_T_DECIMAL:
		return v.decimal_width <= 4 ? 2 : v.decimal_width <= 9 ? 4 : v.decimal_width <= 18 ? 8 : 16;
	case B2A_T_ENUM: { // codes sized by the dictionary (EnumTypeInfo::DictType)
		const uint64_t n = v.n_children == 1 && v.children ? v.children[0].rows : 0;
		return n <= 0xFF ? 1 : n <= 0xFFFF ? 2 : 4;
	}
	default:
		return 0; // STRUCT / ARRAY / SQLNULL: no data of their own
	}
}

bool IsStringLike(int type_id) {
	return type_id == B2A_T_VARCHAR || type_id == B2A_T_BLOB || type_id == B2A_T_BIT || type_id == B2A_T_VARINT ||
	       type_id == B2A_T_JSON;
}

void *DeviceCopy(HostWrite &hw, const void *src, size_t bytes) {
	void *d = nullptr;
	HW_CUDA(cudaMalloc(&d, bytes ? bytes : 16));
	hw.d_allocs.push_back(d);
	if (bytes) {
		HW_CUDA(cudaMemcpyAsync(d, src, bytes, cudaMemcpyHostToDevice, hw.st));
	}
	return d;
}

inline bool RowValid(const uint64_t *validity, uint64_t i) {
	return !validity || ((validity[i >> 6] >> (i & 63)) & 1);
}

// string_t records of a host vector -> device records whose long strings point into a device heap
void *UploadStrings(HostWrite &hw, const b2a_vector &v) {
	const uint64_t n = v.rows;
	const uint8_t *rec = static_cast<const uint8_t *>(v.data);
	uint8_t *h_rec = nullptr;
	HW_CUDA(cudaHostAlloc(&h_rec, n ? n * 16 : 16, cudaHostAllocDefault));
	hw.h_allocs.push_back(h_rec);
	const unsigned hw_threads = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
	const unsigned T = n < (1u << 16) ? 1 : hw_threads;
	std::vector<uint64_t> part(T + 1, 0);
	auto range = [&](unsigned t) { return std::make_pair(n * t / T, n * (t + 1) / T); };
	{ // pass 1: bytes of long strings per part
		std::vector<std::thread> th;
		for (unsigned t = 0; t < T; t++) {
			th.emplace_back([&, t] {
				auto r = range(t);
				uint64_t sum = 0;
				for (uint64_t i = r.first; i < r.second; i++) {
					uint32_t len;
					std::memcpy(&len, rec + i * 16, 4);
					if (len > 12 && RowValid(v.validity, i)) {
						sum += len;
					}
				}
				part[t + 1] = sum;
			});
		}
		for (auto &x : th) {
			x.join();
		}
	}
	for (unsigned t = 0; t < T; t++) {
		part[t + 1] += part[t];
	}
	const uint64_t heap_bytes = part[T];
	uint8_t *h_heap = nullptr, *d_heap = nullptr;
	HW_CUDA(cudaHostAlloc(&h_heap, heap_bytes + 16, cudaHostAllocDefault));
	hw.h_allocs.push_back(h_heap);
	HW_CUDA(cudaMalloc(&d_heap, heap_bytes + 16));
	hw.d_allocs.push_back(d_heap);
	{ // pass 2: pack payloads, re-point the records
		std::vector<std::thread> th;
		for (unsigned t = 0; t < T; t++) {
			th.emplace_back([&, t] {
				auto r = range(t);
				uint64_t pos = part[t];
				for (uint64_t i = r.first; i < r.second; i++) {
					uint32_t len;
					std::memcpy(&len, rec + i * 16, 4);
					if (len > 12 && RowValid(v.validity, i)) {
						const char *src;
						std::memcpy(&src, rec + i * 16 + 8, 8);
						std::memcpy(h_heap + pos, src, len);
						const uint64_t dptr = reinterpret_cast<uint64_t>(d_heap) + pos;
						std::memcpy(h_rec + i * 16, rec + i * 16, 8);
						std::memcpy(h_rec + i * 16 + 8, &dptr, 8);
						pos += len;
					} else {
						std::memcpy(h_rec + i * 16, rec + i * 16, 16);
					}
				}
			});
		}
		for (auto &x : th) {
			x.join();
		}
	}
	if (heap_bytes) {
		HW_CUDA(cudaMemcpyAsync(d_heap, h_heap, heap_bytes, cudaMemcpyHostToDevice, hw.st));
	}
	void *d_rec = nullptr;
	HW_CUDA(cudaMalloc(&d_rec, n ? n * 16 : 16));
	hw.d_allocs.push_back(d_rec);
	if (n) {
		HW_CUDA(cudaMemcpyAsync(d_rec, h_rec, n * 16, cudaMemcpyHostToDevice, hw.st));
	}
	return d_rec;
}

void Mirror(HostWrite &hw, const b2a_vector &src, b2a_vector &dst, int depth) {
	if (depth > 32) {
		throw Error(B2A_EINVAL, "writer: type nesting too deep");
	}
	dst = src;
	dst.data = nullptr;
	dst.validity = nullptr;
	dst.children = nullptr;
	if (src.validity && src.rows) {
		dst.validity = static_cast<const uint64_t *>(DeviceCopy(hw, src.validity, ((src.rows + 63) / 64) * 8));
	}
	const int w = FixedWidth(src);
	if (w && src.data) {
		dst.data = IsStringLike(src.type_id) ? UploadStrings(hw, src) : DeviceCopy(hw, src.data, src.rows * static_cast<size_t>(w));
	} else if (w && src.rows) {
		throw Error(B2A_EINVAL, "writer: vector '" + std::string(src.name ? src.name : "") + "' has rows but no data");
	}
	if (src.n_children > 0) {
		std::unique_ptr<b2a_vector[]> kids(new b2a_vector[static_cast<size_t>(src.n_children)]);
		for (int32_t c = 0; c < src.n_children; c++) {
			Mirror(hw, src.children[c], kids[static_cast<size_t>(c)], depth + 1);
		}
		dst.children = kids.get();
		hw.child_arrays.push_back(std::move(kids));
	}
}

} // namespace

extern "C" {

int b2a_pinned_alloc(void **out, uint64_t bytes) {
	if (!out) {
		return SetLastError(B2A_EINVAL, "b2a_pinned_alloc: null output");
	}
	cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 16, cudaHostAllocDefault);
	if (e != cudaSuccess) {
		(void)cudaGetLastError();
		return SetLastError(e == cudaErrorMemoryAllocation ? B2A_EOOM : B2A_ECUDA,
		                    std::string("cudaHostAlloc: ") + cudaGetErrorString(e));
	}
	return B2A_OK;
}

void b2a_pinned_free(void *p) {
	if (p) {
		cudaFreeHost(p);
	}
}

int b2a_write_arrow_host(const b2a_vector *host_columns, uint32_t n_columns, uint64_t rows, const b2a_write_options *opt,
                         int device, b2a_host_stream *out) {
	if (!out) {
		return SetLastError(B2A_EINVAL, "b2a_write_arrow_host: null output");
	}
	std::memset(out, 0, sizeof(*out));
	if (device < 0) {
		return SetLastError(B2A_ECUDA, "b2a_write_arrow_host needs a CUDA device; there is no CPU fallback");
	}
	std::unique_ptr<HostWrite> hw(new HostWrite());
	hw->device = device;
	try {
		HW_CUDA(cudaSetDevice(device));
		HW_CUDA(cudaStreamCreateWithFlags(&hw->st, cudaStreamNonBlocking));
		std::vector<b2a_vector> cols(n_columns);
		for (uint32_t c = 0; c < n_columns; c++) {
			Mirror(*hw, host_columns[c], cols[c], 0);
		}
		HW_CUDA(cudaStreamSynchronize(hw->st)); // the pinned staging may be reused / freed by the caller's next call
		int rc = b2a_writer_create(cols.data(), n_columns, rows, opt, device, &hw->writer);
		if (rc != B2A_OK) {
			return rc;
		}
		rc = b2a_writer_measure(hw->writer, hw->st);
		if (rc != B2A_OK) {
			return rc;
		}
		uint64_t stream_bytes = 0, n_batches = 0, body_bytes = 0, alg = 0;
		rc = b2a_writer_sizes(hw->writer, &stream_bytes, &n_batches, &body_bytes, &alg);
		if (rc != B2A_OK) {
			return rc;
		}
		HW_CUDA(cudaMalloc(&hw->d_stream, stream_bytes + 256));
		rc = b2a_writer_encode(hw->writer, hw->d_stream, hw->st);
		if (rc != B2A_OK) {
			return rc;
		}
		rc = b2a_writer_finish(hw->writer, hw->st);
		if (rc != B2A_OK) {
			return rc;
		}
		HW_CUDA(cudaHostAlloc(&hw->h_stream, stream_bytes + 16, cudaHostAllocDefault));
		HW_CUDA(cudaMemcpyAsync(hw->h_stream, hw->d_stream, stream_bytes, cudaMemcpyDeviceToHost, hw->st));
		HW_CUDA(cudaStreamSynchronize(hw->st));
		rc = b2a_writer_headers(hw->writer, hw->h_stream);
		if (rc != B2A_OK) {
			return rc;
		}
		out->data = hw->h_stream;
		out->size = stream_bytes;
		out->n_batches = n_batches;
		out->writer = hw->writer;
		out->gpu_launches = b2a_writer_gpu_launches(hw->writer);
		out->priv = hw.release();
		return B2A_OK;
	} catch (const CudaFail &f) {
		(void)cudaGetLastError();
		return SetLastError(f.e == cudaErrorMemoryAllocation ? B2A_EOOM : B2A_ECUDA,
		                    std::string("CUDA error: ") + cudaGetErrorString(f.e));
	} catch (const Error &e) {
		return SetLastError(e.code, e.msg);
	} catch (const std::bad_alloc &) {
		return SetLastError(B2A_EOOM, "out of host memory");
	} catch (const std::exception &e) {
		return SetLastError(B2A_EINTERNAL, e.what());
	}
}

void b2a_host_stream_release(b2a_host_stream *s) {
	if (s && s->priv) {
		delete static_cast<HostWrite *>(s->priv);
		std::memset(s, 0, sizeof(*s));
	}
}

} // extern "C"
