// sm_100a encode kernels: DuckDB vectors (resident in HBM) -> Arrow IPC record-batch bodies, every buffer written
// at its final stream offset.  Mirror image of decode.cu; each op restates one ArrowAppender routine
// (duckdb/src/include/duckdb/common/arrow/appender/*.hpp, duckdb/src/common/arrow/appender/*.cpp).
//
//   pass 1  b2a_enc_measure_kernel : per 1 024-row tile, sum of string lengths / list lengths of valid rows
//           b2a_enc_scan_kernel    : per task, exclusive scan of its tile sums (tile bases) + total
//           b2a_enc_expand_kernel  : child selection vectors of lists (ArrowListData::AppendOffsets child_sel)
//   pass 2  b2a_encode_kernel      : ONE launch for all buffers of all batches (persistent CTAs over a tile table)
//
// All integer/byte work, HBM-bound; no tensor cores.
#include "encode.cuh"

#include <cuda_runtime.h>

#include <cstdlib>

namespace b2a {

static constexpr int kThreads = 256;
static constexpr int kRowsPerThread = B2A_ENC_STR_TILE / kThreads; // 4
static constexpr uint32_t kStageBytes = 37 * 1024;                 // shared-memory staging of one tile's characters
static constexpr uint32_t kRecBytes = B2A_ENC_STR_TILE * 16;       // the string_t records of one tile (prefetched)
static constexpr uint32_t kEncStageArea = (kStageBytes + 64 + 127) & ~127u;
static constexpr uint32_t kEncDynSmem = kEncStageArea + kRecBytes; // dynamic shared memory of b2a_encode_kernel

__device__ __forceinline__ void enc_raise(uint32_t *err, uint32_t code) {
	atomicMax(err, code);
}
__device__ __forceinline__ uint64_t map_row(const b2a_enc_task &t, uint64_t i) {
	return t.sel ? static_cast<uint64_t>(__ldg(reinterpret_cast<const uint32_t *>(t.sel) + i)) : t.row0 + i;
}
__device__ __forceinline__ bool src_valid(const b2a_enc_task &t, uint64_t src_row) {
	if (!t.validity) {
		return true;
	}
	const uint64_t w = __ldg(reinterpret_cast<const unsigned long long *>(t.validity) + (src_row >> 6));
	return (w >> (src_row & 63)) & 1;
}
__device__ __forceinline__ uint4 ld16(const void *p) {
	uint4 r;
	asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
	             : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
	             : "l"(p));
	return r;
}
__device__ __forceinline__ uint2 ld8(const void *p) {
	uint2 r;
	asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
	return r;
}
__device__ __forceinline__ void st16(void *p, uint4 v) {
	asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
	             : "memory");
}
__device__ __forceinline__ void st8(void *p, uint2 v) {
	asm volatile("st.global.cs.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// zero bytes [from, to) of the stream (buffer padding up to the next 8-byte boundary; nanoarrow pads with zeros)
__device__ __forceinline__ void zero_range(uint8_t *p, uint64_t from, uint64_t to) {
	for (uint64_t k = from + threadIdx.x; k < to; k += kThreads) {
		p[k] = 0;
	}
}
__device__ __forceinline__ uint64_t align8(uint64_t v) {
	return (v + 7) & ~uint64_t(7);
}

// block-wide exclusive scan of one uint64 per thread; returns the exclusive prefix, *total = block sum
__device__ __forceinline__ uint64_t block_exclusive_scan(uint64_t v, uint64_t *total) {
	__shared__ uint64_t warp_sums[kThreads / 32];
	__shared__ uint64_t block_total;
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint64_t inc = v;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		uint64_t n = __shfl_up_sync(0xffffffffu, inc, o);
		if (lane >= static_cast<uint32_t>(o)) {
			inc += n;
		}
	}
	__syncthreads(); // previous use of warp_sums is over
	if (lane == 31) {
		warp_sums[warp] = inc;
	}
	__syncthreads();
	uint64_t base = 0;
#pragma unroll
	for (int w = 0; w < kThreads / 32; w++) {
		if (static_cast<uint32_t>(w) < warp) {
			base += warp_sums[w];
		}
	}
	if (threadIdx.x == kThreads - 1) {
		block_total = base + inc;
	}
	__syncthreads();
	*total = block_total;
	return base + inc - v;
}

// length contributed by output row i of a STRING / LIST task (0 for NULL rows), and the raw 16-byte entry
__device__ __forceinline__ uint64_t entry_len(const b2a_enc_task &t, uint64_t i, uint4 &e, bool &valid) {
	const uint64_t sr = map_row(t, i);
	valid = src_valid(t, sr);
	e = __ldg(reinterpret_cast<const uint4 *>(t.src) + sr); // L1-cached: a thread reads 4 adjacent entries
	if (!valid) {
		return 0;
	}
	if (t.op == B2A_ENC_STRING32 || t.op == B2A_ENC_STRING64) {
		return e.x; // string_t.length
	}
	if (t.op == B2A_ENC_STRINGVIEW) {
		return e.x > 12u ? e.x : 0; // only the non-inlined strings reach the data buffer
	}
	if (t.op == B2A_ENC_UUID32 || t.op == B2A_ENC_UUID64) {
		return 36; // UUID::STRING_SIZE (ArrowUUIDConverter::GetLength, varchar_data.hpp:33-37); e = the hugeint_t
	}
	return (static_cast<uint64_t>(e.w) << 32) | e.z; // list_entry_t.length
}

// ------------------------------------------------------------------------------------------------
// pass 1a: per-tile sums.  Simple grid-stride over the tile table (tiles are small and uniform).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
    b2a_enc_measure_kernel(const b2a_enc_task *__restrict__ tasks, const uint32_t *__restrict__ tile_task,
                           uint32_t n_tiles, unsigned long long *__restrict__ heap_bytes) {
	__shared__ uint64_t warp_sums[kThreads / 32];
	// bytes pass 2 will have to fetch from the string heap (strings longer than the 12 inlined bytes): part of the
	// algorithmic traffic of the encode launch; one atomic per CTA
	uint64_t heap_acc = 0;
	for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
		const b2a_enc_task &t = tasks[__ldg(tile_task + tile)];
		const uint32_t k = tile - t.tile0;
		const uint64_t row0 = static_cast<uint64_t>(k) * B2A_ENC_STR_TILE;
		uint64_t sum = 0;
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			const uint64_t i = row0 + u * kThreads + threadIdx.x; // any assignment works for a sum: coalesced
			if (i < t.rows) {
				uint4 e;
				bool valid;
				const uint64_t len = entry_len(t, i, e, valid);
				sum += len;
				if (len > 12 && (t.op == B2A_ENC_STRING32 || t.op == B2A_ENC_STRING64 || t.op == B2A_ENC_STRINGVIEW)) {
					heap_acc += len;
				}
			}
		}
		for (int o = 16; o > 0; o >>= 1) {
			sum += __shfl_xor_sync(0xffffffffu, sum, o);
		}
		if ((threadIdx.x & 31) == 0) {
			warp_sums[threadIdx.x >> 5] = sum;
		}
		__syncthreads();
		if (threadIdx.x == 0) {
			uint64_t s = 0;
			for (int w = 0; w < kThreads / 32; w++) {
				s += warp_sums[w];
			}
			reinterpret_cast<uint64_t *>(t.scratch)[k] = s;
		}
		__syncthreads();
	}
	if (heap_bytes) {
		for (int o = 16; o > 0; o >>= 1) {
			heap_acc += __shfl_xor_sync(0xffffffffu, heap_acc, o);
		}
		if ((threadIdx.x & 31) == 0 && heap_acc) {
			atomicAdd(heap_bytes, static_cast<unsigned long long>(heap_acc));
		}
	}
}

// pass 1b: one warp per task: exclusive scan of its tile sums in place, total -> counters[t.counter]
__global__ void __launch_bounds__(32)
    b2a_enc_scan_kernel(const b2a_enc_task *__restrict__ tasks, uint32_t n_tasks) {
	for (uint32_t ti = blockIdx.x; ti < n_tasks; ti += gridDim.x) {
		const b2a_enc_task &t = tasks[ti];
		const uint64_t n = (t.rows + B2A_ENC_STR_TILE - 1) / B2A_ENC_STR_TILE;
		uint64_t *s = reinterpret_cast<uint64_t *>(t.scratch);
		uint64_t carry = 0;
		for (uint64_t base = 0; base < n; base += 32) {
			const uint64_t i = base + threadIdx.x;
			const uint64_t v = i < n ? s[i] : 0;
			uint64_t inc = v;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				uint64_t x = __shfl_up_sync(0xffffffffu, inc, o);
				if (threadIdx.x >= static_cast<uint32_t>(o)) {
					inc += x;
				}
			}
			if (i < n) {
				s[i] = carry + inc - v;
			}
			carry += __shfl_sync(0xffffffffu, inc, 31);
		}
		if (threadIdx.x == 0) {
			*reinterpret_cast<uint64_t *>(t.counter) = carry;
		}
	}
}

// pass 1c: child selection vectors.  LIST: sel[new_offset(i) + k] = entry.offset + k for every valid row i
// (list_data.hpp:93-95); EXPAND_ARRAY: sel[i] = parent_sel[i / size] * size + i % size.
__global__ void __launch_bounds__(kThreads)
    b2a_enc_expand_kernel(const b2a_enc_task *__restrict__ tasks, const uint32_t *__restrict__ tile_task,
                          uint32_t n_tiles) {
	for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
		const b2a_enc_task &t = tasks[__ldg(tile_task + tile)];
		const uint32_t k = tile - t.tile0;
		uint32_t *out = reinterpret_cast<uint32_t *>(t.dst);
		if (t.op == B2A_ENC_EXPAND_ARRAY) {
			const uint64_t row0 = static_cast<uint64_t>(k) * t.tile_rows;
			const uint64_t end = min(t.rows, row0 + t.tile_rows);
			const uint64_t size = static_cast<uint64_t>(t.param);
			for (uint64_t i = row0 + threadIdx.x; i < end; i += kThreads) {
				out[i] = static_cast<uint32_t>(map_row(t, i / size) * size + i % size);
			}
			continue;
		}
		const uint64_t row0 = static_cast<uint64_t>(k) * B2A_ENC_STR_TILE;
		uint64_t len[kRowsPerThread], off0[kRowsPerThread], sum = 0;
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			const uint64_t i = row0 + threadIdx.x * kRowsPerThread + u;
			len[u] = 0;
			off0[u] = 0;
			if (i < t.rows) {
				uint4 e;
				bool valid;
				len[u] = entry_len(t, i, e, valid);
				off0[u] = (static_cast<uint64_t>(e.y) << 32) | e.x; // list_entry_t.offset
			}
			sum += len[u];
		}
		uint64_t total;
		uint64_t pos = reinterpret_cast<const uint64_t *>(t.scratch)[k] + block_exclusive_scan(sum, &total);
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			for (uint64_t j = 0; j < len[u]; j++) {
				out[pos + j] = static_cast<uint32_t>(off0[u] + j);
			}
			pos += len[u];
		}
	}
}

// ------------------------------------------------------------------------------------------------
// pass 2 ops
// ------------------------------------------------------------------------------------------------

// K10: validity.  Arrow bitmap bit i = validity of input row map(i); the buffer is created 0xFF-filled
// (ResizeValidity) so bits past `rows` in the last byte stay 1; bytes past ceil(rows/8) up to the 8-byte
// boundary are zero padding.  One thread per 64-bit output word.  null_count -> counters[t.counter].
__device__ __forceinline__ void enc_validity(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint64_t row0,
                                             uint64_t nrows) {
	const unsigned long long *mask = reinterpret_cast<const unsigned long long *>(t.validity);
	const uint32_t *sel = reinterpret_cast<const uint32_t *>(t.sel);
	unsigned long long *dst = reinterpret_cast<unsigned long long *>(stream + t.dst);
	const uint64_t w0 = row0 >> 6, w1 = (row0 + nrows + 63) >> 6;
	const uint64_t nbytes = (t.rows + 7) >> 3;
	unsigned long long nulls = 0;
	for (uint64_t w = w0 + threadIdx.x; w < w1; w += kThreads) {
		const uint64_t first = w << 6;
		uint64_t v = ~uint64_t(0);
		if (mask) {
			if (sel) {
				v = 0;
				for (int b = 0; b < 64 && first + b < t.rows; b++) {
					const uint64_t sr = __ldg(sel + first + b);
					v |= ((__ldg(mask + (sr >> 6)) >> (sr & 63)) & 1ull) << b;
				}
			} else {
				const uint64_t s0 = t.row0 + first;
				const uint32_t sh = static_cast<uint32_t>(s0 & 63);
				v = __ldg(mask + (s0 >> 6)) >> sh;
				if (sh && first + (64 - sh) < t.rows) { // the next word exists only if rows reach into it
					v |= __ldg(mask + (s0 >> 6) + 1) << (64 - sh);
				}
			}
		}
		const uint64_t live = t.rows - first >= 64 ? 64 : t.rows - first; // rows in this word
		const uint64_t live_mask = live == 64 ? ~uint64_t(0) : ((uint64_t(1) << live) - 1);
		nulls += __popcll(~v & live_mask);
		v |= ~live_mask; // bits past the last row: 1 ...
		const uint64_t bytes_here = nbytes - (w << 3) >= 8 ? 8 : nbytes - (w << 3);
		if (bytes_here < 8) {
			v &= (uint64_t(1) << (8 * bytes_here)) - 1; // ... but padding bytes are zero
		}
		dst[w] = v;
	}
	for (int o = 16; o > 0; o >>= 1) {
		nulls += __shfl_xor_sync(0xffffffffu, nulls, o);
	}
	if ((threadIdx.x & 31) == 0 && nulls) {
		atomicAdd(reinterpret_cast<unsigned long long *>(t.counter), nulls);
	}
}

// K11: scalar values (ArrowScalarBaseData::Append).  Identity mapping = a straight copy (16-byte vectors when both
// sides allow, else 8-byte); through a selection = element gather.  NULL slots are copied like the reference does.
__device__ __forceinline__ void enc_copy(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint64_t row0,
                                         uint64_t nrows) {
	const uint64_t w = t.src_width;
	uint8_t *dst = stream + t.dst + row0 * w;
	if (t.sel) {
		const uint32_t *sel = reinterpret_cast<const uint32_t *>(t.sel) + row0;
		const uint8_t *src = reinterpret_cast<const uint8_t *>(t.src);
		for (uint64_t i = threadIdx.x; i < nrows; i += kThreads) {
			const uint64_t sr = __ldg(sel + i);
			switch (w) {
			case 1:
				dst[i] = __ldg(src + sr);
				break;
			case 2:
				reinterpret_cast<uint16_t *>(dst)[i] = __ldg(reinterpret_cast<const uint16_t *>(src) + sr);
				break;
			case 4:
				reinterpret_cast<uint32_t *>(dst)[i] = __ldg(reinterpret_cast<const uint32_t *>(src) + sr);
				break;
			case 8:
				st8(reinterpret_cast<uint2 *>(dst) + i, ld8(reinterpret_cast<const uint2 *>(src) + sr));
				break;
			default: { // 16
				const uint2 *p = reinterpret_cast<const uint2 *>(src) + 2 * sr;
				st8(reinterpret_cast<uint2 *>(dst) + 2 * i, ld8(p));
				st8(reinterpret_cast<uint2 *>(dst) + 2 * i + 1, ld8(p + 1));
				break;
			}
			}
		}
		return;
	}
	const uint8_t *src = reinterpret_cast<const uint8_t *>(t.src) + (t.row0 + row0) * w;
	const uint64_t bytes = nrows * w;
	const uintptr_t a = reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst);
	uint64_t done = 0;
	if ((a & 15) == 0) {
		const uint64_t nvec = bytes >> 4;
		constexpr int U = 4;
		uint64_t i = threadIdx.x;
		for (; i + (U - 1) * kThreads < nvec; i += U * kThreads) {
			uint4 v[U];
#pragma unroll
			for (int u = 0; u < U; u++) {
				v[u] = ld16(src + ((i + u * kThreads) << 4));
			}
#pragma unroll
			for (int u = 0; u < U; u++) {
				st16(dst + ((i + u * kThreads) << 4), v[u]);
			}
		}
		for (; i < nvec; i += kThreads) {
			st16(dst + (i << 4), ld16(src + (i << 4)));
		}
		done = nvec << 4;
	} else if ((a & 7) == 0) {
		const uint64_t nvec = bytes >> 3;
		constexpr int U = 4;
		uint64_t i = threadIdx.x;
		for (; i + (U - 1) * kThreads < nvec; i += U * kThreads) {
			uint2 v[U];
#pragma unroll
			for (int u = 0; u < U; u++) {
				v[u] = ld8(src + ((i + u * kThreads) << 3));
			}
#pragma unroll
			for (int u = 0; u < U; u++) {
				st8(dst + ((i + u * kThreads) << 3), v[u]);
			}
		}
		for (; i < nvec; i += kThreads) {
			st8(dst + (i << 3), ld8(src + (i << 3)));
		}
		done = nvec << 3;
	}
	for (uint64_t k = done + threadIdx.x; k < bytes; k += kThreads) {
		dst[k] = src[k];
	}
}

// ArrowScalarData<int32_t,int16_t> (DECIMAL width <= 4) and ArrowIntervalConverter (NULL rows skipped -> 0 here)
__device__ __forceinline__ void enc_convert(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint64_t row0,
                                            uint64_t nrows) {
	for (uint64_t i = row0 + threadIdx.x; i < row0 + nrows; i += kThreads) {
		const uint64_t sr = map_row(t, i);
		if (t.op == B2A_ENC_WIDEN16) {
			reinterpret_cast<int32_t *>(stream + t.dst)[i] =
			    static_cast<int32_t>(__ldg(reinterpret_cast<const int16_t *>(t.src) + sr));
		} else { // interval_t{i32 months, i32 days, i64 micros} -> ArrowInterval{i32, i32, i64 nanos}
			uint2 a = make_uint2(0, 0), b = make_uint2(0, 0);
			if (src_valid(t, sr)) {
				const uint2 *p = reinterpret_cast<const uint2 *>(t.src) + 2 * sr;
				a = ld8(p);
				const uint2 m = ld8(p + 1);
				const int64_t micros = static_cast<int64_t>((static_cast<uint64_t>(m.y) << 32) | m.x);
				const uint64_t nanos = static_cast<uint64_t>(micros) * 1000ull; // Interval::NANOS_PER_MICRO
				b = make_uint2(static_cast<uint32_t>(nanos), static_cast<uint32_t>(nanos >> 32));
			}
			uint2 *d = reinterpret_cast<uint2 *>(stream + t.dst) + 2 * i;
			st8(d, a);
			st8(d + 1, b);
		}
	}
}

// ArrowTimeTzConverter (scalar_data.hpp:54-68): dtime_tz_t keeps the time in the upper 40 bits -> int64 micros;
// SkipNulls: NULL slots stay unwritten in the reference (0 here)
__device__ __forceinline__ void enc_timetz(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint64_t row0,
                                           uint64_t nrows) {
	for (uint64_t i = row0 + threadIdx.x; i < row0 + nrows; i += kThreads) {
		const uint64_t sr = map_row(t, i);
		uint64_t v = 0;
		if (src_valid(t, sr)) {
			v = __ldg(reinterpret_cast<const unsigned long long *>(t.src) + sr) >> 24; // dtime_tz_t::OFFSET_BITS
		}
		reinterpret_cast<uint64_t *>(stream + t.dst)[i] = v;
	}
}

// ArrowUUIDBlobConverter (scalar_data.hpp:70-87), arrow_lossless_conversion: hugeint_t{u64 lower; i64 upper} -> the 16
// bytes of the UUID, most significant first (the stored sign bit of `upper` flipped back); SkipNulls: 0 here
__device__ __forceinline__ void enc_uuid_blob(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint64_t row0,
                                              uint64_t nrows) {
	for (uint64_t i = row0 + threadIdx.x; i < row0 + nrows; i += kThreads) {
		const uint64_t sr = map_row(t, i);
		uint2 hi = make_uint2(0, 0), lo = make_uint2(0, 0);
		if (src_valid(t, sr)) {
			const uint4 e = __ldg(reinterpret_cast<const uint4 *>(t.src) + sr);
			hi.x = __byte_perm(e.w ^ 0x80000000u, 0, 0x0123);
			hi.y = __byte_perm(e.z, 0, 0x0123);
			lo.x = __byte_perm(e.y, 0, 0x0123);
			lo.y = __byte_perm(e.x, 0, 0x0123);
		}
		uint2 *out = reinterpret_cast<uint2 *>(stream + t.dst) + 2 * i; // body buffers are 8-byte aligned, not 16
		out[0] = hi;
		out[1] = lo;
	}
}

// ArrowUnionData::Append (union_data.cpp:34-47): type id = tag of a valid union row, 0 for a NULL one
__device__ __forceinline__ void enc_union_tag(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint64_t row0,
                                              uint64_t nrows) {
	for (uint64_t i = row0 + threadIdx.x; i < row0 + nrows; i += kThreads) {
		const uint64_t sr = map_row(t, i);
		stream[t.dst + i] = src_valid(t, sr) ? __ldg(reinterpret_cast<const uint8_t *>(t.src) + sr) : uint8_t(0);
	}
}

// UUID::ToString (duckdb/src/common/types/uuid.cpp:70-101): the sign bit of `upper` is flipped back, 8-4-4-4-12
// lower-case hex digits.  e = hugeint_t{u64 lower; i64 upper}.
__device__ __forceinline__ void put_uuid(uint8_t *dst, const uint4 &e) {
	const uint64_t upper = ((static_cast<uint64_t>(e.w) << 32) | e.z) ^ (1ull << 63);
	const uint64_t lower = (static_cast<uint64_t>(e.y) << 32) | e.x;
	uint32_t pos = 0;
#pragma unroll
	for (int b = 0; b < 16; b++) {
		const uint32_t byte = static_cast<uint32_t>((b < 8 ? upper >> (56 - 8 * b) : lower >> (120 - 8 * b)) & 0xFF);
		const uint32_t hi = byte >> 4, lo = byte & 15;
		dst[pos++] = static_cast<uint8_t>(hi < 10 ? '0' + hi : 'a' + hi - 10);
		dst[pos++] = static_cast<uint8_t>(lo < 10 ? '0' + lo : 'a' + lo - 10);
		if (b == 3 || b == 5 || b == 7 || b == 9) {
			dst[pos++] = '-';
		}
	}
}

// K13: bool bytes -> bits (ArrowBoolData::Append): the bit buffer starts all ones, a bit is cleared only for a VALID
// false; bits past `rows` in the last byte stay 1.  One warp ballot per 32 rows.
__device__ __forceinline__ void enc_bool(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint64_t row0,
                                         uint64_t nrows) {
	uint32_t *dst = reinterpret_cast<uint32_t *>(stream + t.dst);
	const uint8_t *src = reinterpret_cast<const uint8_t *>(t.src);
	const uint64_t end = row0 + nrows;
	const uint64_t nbytes = (t.rows + 7) >> 3;
	for (uint64_t base = row0 + (threadIdx.x & ~31u); base < end; base += kThreads) {
		const uint64_t i = base + (threadIdx.x & 31);
		bool bit = true;
		if (i < t.rows) {
			const uint64_t sr = map_row(t, i);
			bit = !src_valid(t, sr) || __ldg(src + sr) != 0;
		}
		uint32_t word = __ballot_sync(0xffffffffu, bit);
		if ((threadIdx.x & 31) == 0) {
			const uint64_t byte0 = base >> 3;
			const uint64_t bytes_here = nbytes - byte0 >= 4 ? 4 : nbytes - byte0;
			if (bytes_here < 4) {
				word &= (1u << (8 * bytes_here)) - 1;
			}
			dst[base >> 5] = word;
		}
	}
}

// Writes the `len` bytes of one string to dst.  Inlined strings (len <= 12) come out of the string_t registers;
// longer ones are fetched from the pointer as aligned 32-bit words (one load per word, funnel-shifted into place).
// Every thread of a warp copies its own string, so a warp moves 32 strings per loop trip.
__device__ __forceinline__ void put_string(uint8_t *dst, const uint4 &e, uint32_t len) {
	if (len == 0) {
		return;
	}
	if (len <= 12) {
#pragma unroll
		for (int w = 0; w < 3; w++) {
			const uint32_t word = w == 0 ? e.y : (w == 1 ? e.z : e.w);
#pragma unroll
			for (int b = 0; b < 4; b++) {
				if (static_cast<uint32_t>(4 * w + b) < len) {
					dst[4 * w + b] = static_cast<uint8_t>(word >> (8 * b));
				}
			}
		}
		return;
	}
	const uintptr_t p = (static_cast<uintptr_t>(e.w) << 32) | e.z;
	const uint32_t *wp = reinterpret_cast<const uint32_t *>(p & ~uintptr_t(3));
	const uint32_t shift = (static_cast<uint32_t>(p) & 3u) * 8u;
	const uint32_t total_words = (len + (shift >> 3) + 3u) >> 2; // aligned words that cover the string
	uint32_t lo = __ldg(wp);
	uint32_t widx = 1;
	for (uint32_t k = 0; k < len; k += 4) {
		uint32_t hi = 0;
		if (widx < total_words) {
			hi = __ldg(wp + widx);
		}
		widx++;
		const uint32_t word = __funnelshift_r(lo, hi, shift);
		const uint32_t nb = len - k;
		dst[k] = static_cast<uint8_t>(word);
		if (nb > 1) {
			dst[k + 1] = static_cast<uint8_t>(word >> 8);
		}
		if (nb > 2) {
			dst[k + 2] = static_cast<uint8_t>(word >> 16);
		}
		if (nb > 3) {
			dst[k + 3] = static_cast<uint8_t>(word >> 24);
		}
		lo = hi;
	}
}

// Same contract as put_string, for the shared-memory stage: the bytes leave in aligned 32-bit stores (plus at most three
// byte stores at either end, because neighbouring strings share those words), and the source words of a long string are
// fetched eight at a time before anything is stored — a warp issues in order, so a loop that stores each word right after
// loading it pays one memory latency per word.
__device__ __forceinline__ void emit_words(uint8_t *dst, const uint32_t (&w)[9], uint32_t r, uint32_t n) {
#pragma unroll
	for (uint32_t i = 0; i < 8; i++) {
		if (i < n) {
			*reinterpret_cast<uint32_t *>(dst + 4 * i) = __funnelshift_r(w[i], w[i + 1], r);
		}
	}
}

__device__ __forceinline__ void put_string_words(uint8_t *dst, const uint4 &e, uint32_t len) {
	if (len == 0) {
		return;
	}
	const uint32_t dalign = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(dst)) & 3u;
	uint32_t head = (4u - dalign) & 3u;
	if (head > len) {
		head = len;
	}
	const uint32_t body = (len - head) >> 2, tail = (len - head) & 3u;
	if (len <= 12) {
		// source = the 12 inlined bytes in registers
		const uint32_t w[5] = {e.y, e.z, e.w, 0u, 0u};
#pragma unroll
		for (uint32_t j = 0; j < 3; j++) {
			if (j < head) {
				dst[j] = static_cast<uint8_t>(w[0] >> (8 * j));
			}
		}
		const uint32_t r = head * 8; // source byte `head` is byte (head & 3) of word 0 (head <= 3)
		uint8_t *d = dst + head;
#pragma unroll
		for (uint32_t k = 0; k < 3; k++) {
			if (k < body) {
				*reinterpret_cast<uint32_t *>(d + 4 * k) = __funnelshift_r(w[k], w[k + 1], r);
			}
		}
		const uint32_t v = body == 0 ? __funnelshift_r(w[0], w[1], r)
		                             : (body == 1 ? __funnelshift_r(w[1], w[2], r) : __funnelshift_r(w[2], w[3], r));
#pragma unroll
		for (uint32_t j = 0; j < 3; j++) {
			if (j < tail) {
				d[4 * body + j] = static_cast<uint8_t>(v >> (8 * j));
			}
		}
		return;
	}
	const uintptr_t p = (static_cast<uintptr_t>(e.w) << 32) | e.z;
	const uint32_t *wp = reinterpret_cast<const uint32_t *>(p & ~uintptr_t(3));
	const uint32_t sshift = static_cast<uint32_t>(p) & 3u;
	const uint32_t last = (sshift + len - 1) >> 2; // index of the last aligned source word that holds string bytes
	// head bytes (<= 3) come out of source words 0 and 1
	const uint32_t w0 = __ldg(wp), w1 = last >= 1 ? __ldg(wp + 1) : 0u;
#pragma unroll
	for (uint32_t j = 0; j < 3; j++) {
		if (j < head) {
			const uint32_t b = sshift + j;
			dst[j] = static_cast<uint8_t>((b < 4 ? w0 : w1) >> (8 * (b & 3)));
		}
	}
	const uint32_t s = sshift + head;
	const uint32_t r = (s & 3u) * 8u;
	uint32_t idx = (s >> 2) + 1; // next source word to fetch
	uint32_t lo = (s >> 2) == 0 ? w0 : w1;
	uint8_t *d = dst + head;
	uint32_t left = body;
	while (left > 0) {
		uint32_t w[9];
		w[0] = lo;
#pragma unroll
		for (uint32_t i = 0; i < 8; i++) {
			w[i + 1] = idx + i <= last ? __ldg(wp + idx + i) : 0u;
		}
		const uint32_t n = left < 8 ? left : 8;
		emit_words(d, w, r, n);
		// the word after the last emitted one becomes the new low word
		lo = n == 8 ? w[8] : (n == 7 ? w[7] : (n == 6 ? w[6] : (n == 5 ? w[5] : (n == 4 ? w[4] : (n == 3 ? w[3] : (n == 2 ? w[2] : w[1]))))));
		idx += n;
		d += 4 * n;
		left -= n;
	}
	if (tail) {
		const uint32_t hi = idx <= last ? __ldg(wp + idx) : 0u;
		const uint32_t v = __funnelshift_r(lo, hi, r);
#pragma unroll
		for (uint32_t j = 0; j < 3; j++) {
			if (j < tail) {
				d[j] = static_cast<uint8_t>(v >> (8 * j));
			}
		}
	}
}

// K12: string_t -> offsets (running sum of lengths, NULL rows repeat the previous offset) + character bytes
// (ArrowVarcharData::AppendTemplated).  Tile = 1 024 rows, thread = 4 consecutive rows.  The tile's characters form
// one contiguous window of the data buffer: they are assembled in shared memory (inlined strings from registers,
// long strings by warp-cooperative coalesced loads) and leave with aligned 16-byte stores.  Windows larger than the
// staging area are written straight to global memory.
template <typename OFF, bool UUID = false>
__device__ __forceinline__ void enc_string(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint32_t k,
                                           uint8_t *stage, uint32_t *err) {
	const uint64_t row0 = static_cast<uint64_t>(k) * B2A_ENC_STR_TILE;
	uint4 e[kRowsPerThread];
	uint32_t len[kRowsPerThread];
	uint64_t sum = 0;
#pragma unroll
	for (int u = 0; u < kRowsPerThread; u++) {
		const uint64_t i = row0 + threadIdx.x * kRowsPerThread + u;
		len[u] = 0;
		e[u] = make_uint4(0, 0, 0, 0);
		if (i < t.rows) {
			bool valid;
			len[u] = static_cast<uint32_t>(entry_len(t, i, e[u], valid));
		}
		sum += len[u];
	}
	uint64_t window;
	const uint64_t tile_base = reinterpret_cast<const uint64_t *>(t.scratch)[k];
	const uint64_t rel0 = block_exclusive_scan(sum, &window);
	// ---- offsets: out[i + 1] = running sum through row i; out[0] = 0.  Thread t owns the ALIGNED group of positions
	// row0 + 4t .. row0 + 4t + 3 = {its exclusive prefix, the sums through its first three rows}: one 16-byte (int32) or
	// two 16-byte (int64) coalesced stores instead of four scattered 4-byte ones; the sum through its fourth row is the
	// next thread's (or the next tile's first thread's) prefix, and the owner of the last row writes out[rows] itself.
	OFF *offs = reinterpret_cast<OFF *>(stream + t.dst);
	{
		uint64_t v[kRowsPerThread + 1];
		v[0] = tile_base + rel0;
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			v[u + 1] = v[u] + len[u];
		}
		if (sizeof(OFF) == 4 && v[kRowsPerThread] > 0x7FFFFFFFull) {
			enc_raise(err, B2A_DEV_ENC_OFFSET_I32);
		}
		const uint64_t p0 = row0 + static_cast<uint64_t>(threadIdx.x) * kRowsPerThread;
		OFF *o = offs + p0;
		static_assert(kRowsPerThread == 4, "offset groups are four positions wide");
		if (p0 + 3 <= t.rows && (reinterpret_cast<uintptr_t>(o) & 15u) == 0) {
			if (sizeof(OFF) == 4) {
				st16(o, make_uint4(static_cast<uint32_t>(v[0]), static_cast<uint32_t>(v[1]), static_cast<uint32_t>(v[2]),
				                   static_cast<uint32_t>(v[3])));
			} else {
				st16(o, make_uint4(static_cast<uint32_t>(v[0]), static_cast<uint32_t>(v[0] >> 32),
				                   static_cast<uint32_t>(v[1]), static_cast<uint32_t>(v[1] >> 32)));
				st16(o + 2, make_uint4(static_cast<uint32_t>(v[2]), static_cast<uint32_t>(v[2] >> 32),
				                       static_cast<uint32_t>(v[3]), static_cast<uint32_t>(v[3] >> 32)));
			}
		} else {
#pragma unroll
			for (int u = 0; u < kRowsPerThread; u++) {
				if (p0 + u <= t.rows) {
					o[u] = static_cast<OFF>(v[u]);
				}
			}
		}
		if (p0 + kRowsPerThread == t.rows) {
			o[kRowsPerThread] = static_cast<OFF>(v[kRowsPerThread]);
		}
	}
	if (row0 + B2A_ENC_STR_TILE >= t.rows) { // last tile: zero padding of the character buffer up to 8 bytes
		const uint64_t total = tile_base + window;
		zero_range(stream, t.dst2 + total, t.dst2 + align8(total));
	}
	if (window == 0) {
		return;
	}
	// ---- characters
	uint8_t *gdst = stream + t.dst2 + tile_base; // window start in the stream
	const uint32_t phase = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(gdst) & 15);
	const bool staged = window + phase <= kStageBytes;
	{
		uint64_t rel = rel0;
		if (staged) { // separate paths so that the stores compile to st.shared / st.global, not generic stores
			uint8_t *wdst = stage + phase;
			// ---- contiguous-heap fast path.  Vectors that come out of a scan keep their strings end to end in row order
			// (Arrow character buffers, DuckDB string heaps filled by one append): then pointer - (offset of the string in
			// this tile's window) is the same address for every non-inlined string, and the window is one memcpy from
			// there — coalesced word loads instead of one word loop per string.  Inlined strings need not exist in that
			// memory: their bytes are patched in from the string_t registers afterwards.  The copy spans first .. last
			// non-inlined string only, and a tile qualifies only if every warp holds a non-inlined string: any run of
			// inlined strings in between is then shorter than 256 rows x 12 B < 4 KiB, the smallest page, so every byte
			// read lies in the page of a byte some string_t points at.
			__shared__ unsigned long long s_base, s_lo, s_hi;
			// (tiles that average under 8 B per row — flags, codes — rarely qualify and skip all of this)
			const bool try_fast = !UUID && window >= 8ull * B2A_ENC_STR_TILE;
			bool fast = false;
			if (try_fast) {
				unsigned long long my_base = 0;
				bool has_long = false, bad = false;
				{
					uint64_t r = rel0;
	#pragma unroll
					for (int u = 0; u < kRowsPerThread; u++) {
						if (len[u] > 12) {
							const unsigned long long b = ((static_cast<unsigned long long>(e[u].w) << 32) | e[u].z) - r;
							bad = bad || (has_long && b != my_base);
							my_base = b;
							has_long = true;
						}
						r += len[u];
					}
				}
				const uint32_t longs = __ballot_sync(0xffffffffu, has_long);
				const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
				if (longs == 0) {
					bad = true;
				} else if (lane == static_cast<uint32_t>(__ffs(longs) - 1)) {
					if (warp == 0) { // first non-inlined string of the tile
						uint64_t r = rel0;
						bool done = false;
	#pragma unroll
						for (int u = 0; u < kRowsPerThread; u++) {
							if (!done && len[u] > 12) {
								s_lo = r;
								done = true;
							}
							r += len[u];
						}
						s_base = my_base;
					}
				}
				if (longs != 0 && warp == kThreads / 32 - 1 && lane == 31u - static_cast<uint32_t>(__clz(longs))) {
					uint64_t r = rel0, hi = 0; // end of the last non-inlined string of the tile
	#pragma unroll
					for (int u = 0; u < kRowsPerThread; u++) {
						r += len[u];
						if (len[u] > 12) {
							hi = r;
						}
					}
					s_hi = hi;
				}
				__syncthreads();
				bad = bad || (has_long && my_base != s_base);
				fast = !__syncthreads_or(bad);
			}
			if (fast) {
				const uint64_t lo = s_lo, n = s_hi - lo;
				const uint8_t *S = reinterpret_cast<const uint8_t *>(s_base) + lo;
				uint8_t *D = wdst + lo;
				const uint32_t head = min(n, static_cast<uint64_t>((4u - (static_cast<uint32_t>(
				                                                            reinterpret_cast<uintptr_t>(D)) & 3u)) & 3u));
				if (threadIdx.x < head) {
					D[threadIdx.x] = __ldg(S + threadIdx.x);
				}
				const uint64_t words = (n - head) >> 2;
				const uintptr_t sa = reinterpret_cast<uintptr_t>(S) + head;
				const uint32_t *wp = reinterpret_cast<const uint32_t *>(sa & ~uintptr_t(3));
				const uint32_t sh = (static_cast<uint32_t>(sa) & 3u) * 8u;
				uint32_t *dw = reinterpret_cast<uint32_t *>(D + head);
				// destination word j = funnel(source word j, j + 1).  Words up to the first 16-byte boundary of the
				// SOURCE one by one, then source-aligned chunks of four words: one 16-byte load + the next chunk's first
				// word per chunk, four chunks (80 B) in flight per thread
				const uint64_t j0 = min(words, static_cast<uint64_t>((16u - (static_cast<uint32_t>(
				                                                           reinterpret_cast<uintptr_t>(wp)) & 15u)) & 15u) >> 2);
				if (threadIdx.x < j0) {
					const uint32_t a = __ldg(wp + threadIdx.x), b = sh ? __ldg(wp + threadIdx.x + 1) : 0u;
					dw[threadIdx.x] = __funnelshift_r(a, b, sh);
				}
				const uint64_t chunks = (words - j0) >> 2;
				const uint4 *cp = reinterpret_cast<const uint4 *>(wp + j0);
				uint32_t *cd = dw + j0;
				constexpr uint32_t kU = 4;
				uint64_t c = threadIdx.x;
				for (; c + (kU - 1) * kThreads < chunks; c += kU * kThreads) {
					uint4 v[kU];
					uint32_t nx[kU];
#pragma unroll
					for (uint32_t k = 0; k < kU; k++) {
						v[k] = __ldg(cp + c + k * kThreads);
						nx[k] = sh ? __ldg(reinterpret_cast<const uint32_t *>(cp + c + k * kThreads + 1)) : 0u;
					}
#pragma unroll
					for (uint32_t k = 0; k < kU; k++) {
						uint32_t *o = cd + 4 * (c + k * kThreads);
						o[0] = __funnelshift_r(v[k].x, v[k].y, sh);
						o[1] = __funnelshift_r(v[k].y, v[k].z, sh);
						o[2] = __funnelshift_r(v[k].z, v[k].w, sh);
						o[3] = __funnelshift_r(v[k].w, nx[k], sh);
					}
				}
				for (; c < chunks; c += kThreads) {
					const uint4 v = __ldg(cp + c);
					const uint32_t nx = sh ? __ldg(reinterpret_cast<const uint32_t *>(cp + c + 1)) : 0u;
					uint32_t *o = cd + 4 * c;
					o[0] = __funnelshift_r(v.x, v.y, sh);
					o[1] = __funnelshift_r(v.y, v.z, sh);
					o[2] = __funnelshift_r(v.z, v.w, sh);
					o[3] = __funnelshift_r(v.w, nx, sh);
				}
				{
					const uint64_t j = j0 + (chunks << 2) + threadIdx.x;
					if (j < words) {
						const uint32_t a = __ldg(wp + j), b = sh ? __ldg(wp + j + 1) : 0u;
						dw[j] = __funnelshift_r(a, b, sh);
					}
				}
				const uint64_t tail0 = head + (words << 2);
				if (tail0 + threadIdx.x < n) {
					D[tail0 + threadIdx.x] = __ldg(S + tail0 + threadIdx.x);
				}
				__syncthreads(); // the inlined strings below overwrite bytes of the span just copied
#pragma unroll
				for (int u = 0; u < kRowsPerThread; u++) {
					if (len[u] <= 12) {
						put_string_words(wdst + rel, e[u], len[u]);
					}
					rel += len[u];
				}
			} else {
#pragma unroll
				for (int u = 0; u < kRowsPerThread; u++) {
					if (UUID) {
						if (len[u]) {
							put_uuid(wdst + rel, e[u]);
						}
					} else {
						put_string_words(wdst + rel, e[u], len[u]);
					}
					rel += len[u];
				}
			}
		} else {
#pragma unroll
			for (int u = 0; u < kRowsPerThread; u++) {
				if (UUID) { // (never taken: 1 024 x 36 bytes always fit the stage)
					if (len[u]) {
						put_uuid(gdst + rel, e[u]);
					}
				} else {
					put_string(gdst + rel, e[u], len[u]);
				}
				rel += len[u];
			}
		}
	}
	if (!staged) {
		return;
	}
	__syncthreads();
	// stage -> global: head bytes up to the first 16-byte boundary, aligned vectors, tail bytes
	const uint64_t total = window;
	uint64_t head = phase ? 16 - phase : 0;
	if (head > total) {
		head = total;
	}
	for (uint64_t b = threadIdx.x; b < head; b += kThreads) {
		gdst[b] = stage[phase + b];
	}
	const uint64_t nvec = (total - head) >> 4;
	const uint4 *svec = reinterpret_cast<const uint4 *>(stage + phase + head); // 16-byte aligned in shared memory
	uint4 *gvec = reinterpret_cast<uint4 *>(gdst + head);
	for (uint64_t v = threadIdx.x; v < nvec; v += kThreads) {
		st16(gvec + v, svec[v]);
	}
	for (uint64_t b = head + (nvec << 4) + threadIdx.x; b < total; b += kThreads) {
		gdst[b] = stage[phase + b];
	}
}

// ------------------------------------------------------------------------------------------------
// mbarrier + 1-D bulk async copy (TMA unit, SASS UBLKCP): global -> shared without passing through registers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t enc_smem_u32(const void *p) {
	return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void enc_mbar_init(uint64_t *bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(enc_smem_u32(bar)), "r"(count) : "memory");
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void enc_bulk_load(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
	// earlier generic-proxy accesses of the staging area (the previous tile) are ordered before the async-proxy writes
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(enc_smem_u32(bar)), "r"(bytes) : "memory");
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
	                 enc_smem_u32(dst_smem)),
	             "l"(src_gmem), "r"(bytes), "r"(enc_smem_u32(bar))
	             : "memory");
}
__device__ __forceinline__ void enc_mbar_wait(uint64_t *bar, uint32_t parity) {
	asm volatile("{\n"
	             ".reg .pred P1;\n"
	             "ENC_WAIT:\n"
	             "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
	             "@P1 bra ENC_DONE;\n"
	             "bra ENC_WAIT;\n"
	             "ENC_DONE:\n"
	             "}" ::"r"(enc_smem_u32(bar)),
	             "r"(parity)
	             : "memory");
}

// 16 bytes starting at byte offset `sb` of the staging area (any alignment): two aligned 16-byte shared loads and a
// funnel shift; `sb & 15` is the same for every thread of a store loop, so the switch is a uniform branch
__device__ __forceinline__ uint4 stage_load16(const uint8_t *stage, uint32_t sb) {
	const uint4 a = *reinterpret_cast<const uint4 *>(stage + (sb & ~15u));
	if ((sb & 15u) == 0) {
		return a;
	}
	const uint4 b = *reinterpret_cast<const uint4 *>(stage + (sb & ~15u) + 16);
	const uint32_t r = (sb & 3u) * 8u;
	switch ((sb & 15u) >> 2) {
	case 0:
		return make_uint4(__funnelshift_r(a.x, a.y, r), __funnelshift_r(a.y, a.z, r), __funnelshift_r(a.z, a.w, r),
		                  __funnelshift_r(a.w, b.x, r));
	case 1:
		return make_uint4(__funnelshift_r(a.y, a.z, r), __funnelshift_r(a.z, a.w, r), __funnelshift_r(a.w, b.x, r),
		                  __funnelshift_r(b.x, b.y, r));
	case 2:
		return make_uint4(__funnelshift_r(a.z, a.w, r), __funnelshift_r(a.w, b.x, r), __funnelshift_r(b.x, b.y, r),
		                  __funnelshift_r(b.y, b.z, r));
	default:
		return make_uint4(__funnelshift_r(a.w, b.x, r), __funnelshift_r(b.x, b.y, r), __funnelshift_r(b.y, b.z, r),
		                  __funnelshift_r(b.z, b.w, r));
	}
}

// Exclusive scan of the tile's 1 024 lengths in ROW order for the coalesced row mapping (row = u * 256 + thread): four
// warp scans, then the 32 warp totals are scanned by every warp for itself — one barrier.  32-bit sums: a tile whose
// strings could overflow them (one string of 4 MiB or more) reports `big` and is handled by the 64-bit generic path.
__device__ __forceinline__ void enc_tile_scan(const uint32_t (&len)[kRowsPerThread], uint32_t (&rel)[kRowsPerThread],
                                              uint32_t &window, bool &big, uint32_t *s_wsum, uint32_t *s_big) {
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	bool mine_big = false;
#pragma unroll
	for (int u = 0; u < kRowsPerThread; u++) {
		mine_big = mine_big || len[u] >= (1u << 22);
		uint32_t inc = len[u];
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
			if (lane >= static_cast<uint32_t>(o)) {
				inc += n;
			}
		}
		rel[u] = inc - len[u];
		if (lane == 31) {
			s_wsum[u * (kThreads / 32) + warp] = inc;
		}
	}
	if (mine_big) {
		*s_big = 1;
	}
	__syncthreads();
	big = *s_big != 0;
	uint32_t inc = s_wsum[lane]; // kRowsPerThread * kThreads / 32 == 32 entries
	const uint32_t own = inc;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
		if (lane >= static_cast<uint32_t>(o)) {
			inc += n;
		}
	}
	window = __shfl_sync(0xffffffffu, inc, 31);
	const uint32_t exc = inc - own;
#pragma unroll
	for (int u = 0; u < kRowsPerThread; u++) {
		rel[u] += __shfl_sync(0xffffffffu, exc, u * (kThreads / 32) + static_cast<int>(warp));
	}
}

// an inlined string (<= 12 bytes, in the string_t registers) into the staging area; one-character strings (flags, codes)
// are a single byte store
__device__ __forceinline__ void put_inlined(uint8_t *dst, const uint4 &e, uint32_t len) {
	if (len == 1) {
		*dst = static_cast<uint8_t>(e.y);
	} else {
		put_string_words(dst, e, len);
	}
}

// Staging area (window byte 0 at byte `ph`) -> stream: head bytes up to the first 16-byte boundary of the DESTINATION,
// aligned 16-byte vectors (shifted from the staging phase to the destination's), tail bytes
__device__ __forceinline__ void enc_stage_flush(const uint8_t *stage, uint32_t ph, uint8_t *gdst, uint32_t total) {
	const uint32_t pd = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(gdst) & 15);
	uint32_t head = (16u - pd) & 15u;
	if (head > total) {
		head = total;
	}
	if (threadIdx.x < head) {
		gdst[threadIdx.x] = stage[ph + threadIdx.x];
	}
	const uint32_t nvec = (total - head) >> 4;
	uint4 *gvec = reinterpret_cast<uint4 *>(gdst + head);
	const uint32_t sb0 = ph + head;
	for (uint32_t v = threadIdx.x; v < nvec; v += kThreads) {
		st16(gvec + v, stage_load16(stage, sb0 + 16u * v));
	}
	for (uint32_t b = head + (nvec << 4) + threadIdx.x; b < total; b += kThreads) {
		gdst[b] = stage[ph + b];
	}
}

// Cross-tile pipeline of the string ops: while a CTA works on tile n, the 16 KiB of string_t records of ITS next tile
// (n + gridDim.x) arrive in shared memory by one bulk async copy — a tile otherwise starts with every warp waiting for
// that first load (27 % of the stall samples of the first cut, profiles/r02_encode_kernel_v2_first_cut_details.txt).
struct EncRecPipe {
	uint8_t *rec;             // kRecBytes, 128-byte aligned
	uint64_t *bar;            // completion of the records copy
	uint32_t parity;          // phase the next wait expects
	bool ready;               // `rec` receives the CURRENT tile's records
	bool next_ready;          // ... the next tile's (set while the current tile is processed)
	const b2a_enc_task *next; // descriptor of the CTA's next tile (resident in shared memory), nullptr if none
	uint32_t next_tile;
};

// issues the copy for the next tile if it is a string tile whose rows are adjacent records (no selection vector);
// every thread evaluates the condition (it is uniform), one thread issues
__device__ __forceinline__ bool enc_prefetch_records(const EncRecPipe &rp) {
#ifdef B2A_ENC_STRING_V1
	return false;
#else
	if (!rp.next) {
		return false;
	}
	const b2a_enc_task &tn = *rp.next;
	if (!(tn.op == B2A_ENC_STRING32 || tn.op == B2A_ENC_STRING64 || tn.op == B2A_ENC_STRINGVIEW) || tn.sel != 0 ||
	    (tn.src & 15u) != 0) {
		return false;
	}
	const uint64_t r0 = static_cast<uint64_t>(rp.next_tile - tn.tile0) * B2A_ENC_STR_TILE;
	if (r0 >= tn.rows) {
		return false;
	}
	const uint64_t n = tn.rows - r0 < B2A_ENC_STR_TILE ? tn.rows - r0 : B2A_ENC_STR_TILE;
	if (threadIdx.x == 0) {
		enc_bulk_load(rp.rec, reinterpret_cast<const void *>(tn.src + (tn.row0 + r0) * 16), static_cast<uint32_t>(n * 16),
		              rp.bar);
	}
	return true;
#endif
}

// K12, second generation (VARCHAR / BLOB with int32 or int64 offsets; UUID keeps enc_string): string_t -> offsets +
// characters (ArrowVarcharData::AppendTemplated, varchar_data.hpp:53-111).  Tile = 1 024 rows.
//  * the 16 KiB of string_t are read fully coalesced (row = u * 256 + thread), the tile-wide exclusive scan of the lengths
//    is four warp scans + one 32-entry scan every warp repeats for itself (ONE barrier), offsets leave as coalesced 4 / 8
//    byte stores;
//  * contiguous-heap tiles (vectors that come out of a scan: pointer - window offset is the same address for every
//    non-inlined string): the span first .. last non-inlined string is ONE bulk async copy (cp.async.bulk -> UBLKCP) of
//    the enclosing 16-byte blocks into the staging area, issued by one thread, no registers in between; inlined strings
//    are patched in from the string_t registers afterwards.  The staging area then has the SOURCE's phase inside a
//    16-byte vector; the store loop shifts to the destination's phase (stage_load16) and writes aligned 16-byte vectors;
//  * any other tile assembles its characters string by string (put_string_words) at the destination's phase.
// Page safety of the bulk copy as in enc_string: a tile qualifies only if every group of 128 consecutive rows holds a
// non-inlined string, so any run of inlined strings inside the span is shorter than 256 x 12 B < 4 KiB.
template <typename OFF>
__device__ __forceinline__ void enc_string2(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint32_t k,
                                            uint8_t *stage, uint32_t *err, uint64_t *bar, uint32_t &parity,
                                            EncRecPipe &rp) {
	__shared__ uint32_t s_wsum[32];
	__shared__ unsigned long long s_base;
	__shared__ uint32_t s_lo, s_hi, s_big;
	__shared__ uint32_t s_grp[8];
	const uint64_t row0 = static_cast<uint64_t>(k) * B2A_ENC_STR_TILE;
	const uint32_t lane = threadIdx.x & 31;
	const uint64_t tile_base = __ldg(reinterpret_cast<const unsigned long long *>(t.scratch) + k); // (needed late: issued early)
	uint4 e[kRowsPerThread];
	uint32_t len[kRowsPerThread];
	uint32_t rel[kRowsPerThread];
	const bool from_pipe = rp.ready;
	if (from_pipe) { // the records were fetched while the previous tile was processed
		enc_mbar_wait(rp.bar, rp.parity);
		rp.parity ^= 1u;
		const uint4 *rec = reinterpret_cast<const uint4 *>(rp.rec);
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			const uint32_t r = static_cast<uint32_t>(u) * kThreads + threadIdx.x;
			len[u] = 0;
			e[u] = make_uint4(0, 0, 0, 0);
			if (row0 + r < t.rows) {
				e[u] = rec[r];
				len[u] = src_valid(t, t.row0 + row0 + r) ? e[u].x : 0u;
			}
		}
	} else {
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			const uint64_t i = row0 + static_cast<uint64_t>(u) * kThreads + threadIdx.x;
			len[u] = 0;
			e[u] = make_uint4(0, 0, 0, 0);
			if (i < t.rows) {
				bool valid;
				len[u] = static_cast<uint32_t>(entry_len(t, i, e[u], valid));
			}
		}
	}
	if (threadIdx.x < 8) {
		s_grp[threadIdx.x] = 0;
	}
	if (threadIdx.x == 0) {
		s_lo = ~0u;
		s_hi = 0;
		s_base = 0;
		s_big = 0;
	}
	uint32_t window;
	bool big;
	enc_tile_scan(len, rel, window, big, s_wsum, &s_big);
	if (from_pipe) { // every thread has its records in registers (the scan's barrier): the area takes the next tile's
		rp.next_ready = enc_prefetch_records(rp);
	}
	if (big) { // (a string of 4 MiB or more in this tile) the generic 64-bit path
		__syncthreads();
		enc_string<OFF>(t, stream, k, stage, err);
		return;
	}
	// ---- offsets: out[i] = running sum before row i, out[rows] = total
	{
		OFF *offs = reinterpret_cast<OFF *>(stream + t.dst);
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			const uint64_t i = row0 + static_cast<uint64_t>(u) * kThreads + threadIdx.x;
			if (i < t.rows) {
				const uint64_t v = tile_base + rel[u];
				if (sizeof(OFF) == 4 && v + len[u] > 0x7FFFFFFFull) {
					enc_raise(err, B2A_DEV_ENC_OFFSET_I32);
				}
				offs[i] = static_cast<OFF>(v);
				if (i + 1 == t.rows) {
					offs[i + 1] = static_cast<OFF>(v + len[u]);
				}
			}
		}
		if (t.rows == 0 && k == 0 && threadIdx.x == 0) {
			offs[0] = 0;
		}
	}
	if (row0 + B2A_ENC_STR_TILE >= t.rows) { // last tile: zero padding of the character buffer up to 8 bytes
		const uint64_t total = tile_base + window;
		zero_range(stream, t.dst2 + total, t.dst2 + align8(total));
	}
	if (window == 0) {
		return;
	}
	// ---- characters
	uint8_t *gdst = stream + t.dst2 + tile_base; // window start in the stream
	const uint32_t pd = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(gdst) & 15);
	if (window + 48 > kStageBytes) { // (a tile of very long strings) straight to global memory, string by string
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			put_string(gdst + rel[u], e[u], len[u]);
		}
		return;
	}
	bool fast = false;
	if (window >= 8u * B2A_ENC_STR_TILE) { // (tiles that average under 8 B per row — flags, codes — rarely qualify)
		unsigned long long my_base = 0;
		uint32_t my_lo = ~0u, my_hi = 0;
		bool has_long = false, bad = false;
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			if (len[u] > 12) {
				const unsigned long long b = ((static_cast<unsigned long long>(e[u].w) << 32) | e[u].z) - rel[u];
				bad = bad || (has_long && b != my_base);
				my_base = b;
				has_long = true;
				my_lo = my_lo < rel[u] ? my_lo : rel[u];
				my_hi = my_hi > rel[u] + len[u] ? my_hi : rel[u] + len[u];
				s_grp[u * 2 + (threadIdx.x >> 7)] = 1; // (same value from every writer)
			}
		}
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) {
			const uint32_t l2 = __shfl_xor_sync(0xffffffffu, my_lo, o), h2 = __shfl_xor_sync(0xffffffffu, my_hi, o);
			my_lo = my_lo < l2 ? my_lo : l2;
			my_hi = my_hi > h2 ? my_hi : h2;
		}
		const uint32_t longs = __ballot_sync(0xffffffffu, has_long);
		if (longs != 0 && lane == static_cast<uint32_t>(__ffs(longs) - 1)) {
			atomicMin(&s_lo, my_lo);
			atomicMax(&s_hi, my_hi);
			s_base = my_base; // any candidate: every thread checks its own against it below
		}
		__syncthreads();
		bad = bad || (has_long && my_base != s_base);
		if (threadIdx.x < 8) { // every 128-row group that has rows at all must hold a non-inlined string
			const uint64_t g0 = row0 + (threadIdx.x >> 1) * static_cast<uint64_t>(kThreads) + (threadIdx.x & 1) * 128u;
			bad = bad || (g0 < t.rows && s_grp[threadIdx.x] == 0);
		}
		fast = !__syncthreads_or(bad);
	}
	uint32_t ph; // phase of window byte 0 inside the staging area
	if (fast) {
		const unsigned long long S = s_base;
		const uint32_t lo = s_lo, hi = s_hi;
		ph = static_cast<uint32_t>(S & 15ull);
		if (threadIdx.x == 0) {
			const unsigned long long a0 = (S + lo) & ~15ull, a1 = (S + hi + 15ull) & ~15ull;
			enc_bulk_load(stage + (a0 - (S & ~15ull)), reinterpret_cast<const void *>(a0), static_cast<uint32_t>(a1 - a0), bar);
		}
		enc_mbar_wait(bar, parity);
		parity ^= 1u;
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			if (len[u] <= 12) {
				put_inlined(stage + ph + rel[u], e[u], len[u]);
			}
		}
	} else {
		ph = pd;
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			put_inlined(stage + ph + rel[u], e[u], len[u]);
		}
	}
	__syncthreads();
	enc_stage_flush(stage, ph, gdst, window);
}

__device__ __forceinline__ void enc_string_view(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint32_t k,
                                                uint32_t *err); // generic 64-bit path, below

// ArrowVarcharToStringViewData::Append (varchar_data.hpp:136-187), second generation: same tile, coalesced row mapping and
// scan as enc_string2; the 16-byte views leave as coalesced stores, and the tile's slice of the ONE data buffer (the
// strings longer than 12 bytes, end to end) is assembled in the staging area at the destination's phase — every thread
// fetches its own long strings eight words at a time (put_string_words) — and flushed with aligned 16-byte stores.
__device__ __forceinline__ void enc_string_view2(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint32_t k,
                                                 uint8_t *stage, uint32_t *err, EncRecPipe &rp) {
	__shared__ uint32_t s_wsum[32];
	__shared__ uint32_t s_big;
	const uint64_t row0 = static_cast<uint64_t>(k) * B2A_ENC_STR_TILE;
	const uint64_t tile_base = __ldg(reinterpret_cast<const unsigned long long *>(t.scratch) + k);
	uint4 e[kRowsPerThread];
	uint32_t len[kRowsPerThread]; // bytes that go to the data buffer (0 for NULL / inlined rows)
	bool valid[kRowsPerThread];
	uint32_t rel[kRowsPerThread];
	if (threadIdx.x == 0) {
		s_big = 0;
	}
	const bool from_pipe = rp.ready;
	if (from_pipe) {
		enc_mbar_wait(rp.bar, rp.parity);
		rp.parity ^= 1u;
		const uint4 *rec = reinterpret_cast<const uint4 *>(rp.rec);
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			const uint32_t r = static_cast<uint32_t>(u) * kThreads + threadIdx.x;
			len[u] = 0;
			valid[u] = false;
			e[u] = make_uint4(0, 0, 0, 0);
			if (row0 + r < t.rows) {
				e[u] = rec[r];
				valid[u] = src_valid(t, t.row0 + row0 + r);
				len[u] = (valid[u] && e[u].x > 12u) ? e[u].x : 0u;
			}
		}
	} else {
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			const uint64_t i = row0 + static_cast<uint64_t>(u) * kThreads + threadIdx.x;
			len[u] = 0;
			valid[u] = false;
			e[u] = make_uint4(0, 0, 0, 0);
			if (i < t.rows) {
				len[u] = static_cast<uint32_t>(entry_len(t, i, e[u], valid[u]));
			}
		}
	}
	uint32_t window;
	bool big;
	enc_tile_scan(len, rel, window, big, s_wsum, &s_big);
	if (from_pipe) {
		rp.next_ready = enc_prefetch_records(rp);
	}
	if (big) {
		__syncthreads();
		enc_string_view(t, stream, k, err);
		return;
	}
#pragma unroll
	for (int u = 0; u < kRowsPerThread; u++) {
		const uint64_t i = row0 + static_cast<uint64_t>(u) * kThreads + threadIdx.x;
		if (i >= t.rows) {
			continue;
		}
		uint4 v = make_uint4(0, 0, 0, 0);
		if (valid[u] && len[u] == 0) { // inlined: string_t and the view share the layout; bytes past the length are zeroed
			const uint32_t n = e[u].x;
			auto keep = [n](uint32_t w, uint32_t first) {
				return n >= first + 4 ? w : n <= first ? 0u : (w & ((1u << (8 * (n - first))) - 1u));
			};
			v = make_uint4(n, keep(e[u].y, 0), keep(e[u].z, 4), keep(e[u].w, 8));
		} else if (valid[u]) {
			const uint64_t run = tile_base + rel[u];
			if (run > 0x7FFFFFFFull) { // the view's offset is an int32 (UnsafeNumericCast<int32_t>(append_data.offset))
				enc_raise(err, B2A_DEV_ENC_OFFSET_I32);
			}
			v = make_uint4(e[u].x, e[u].y, 0u, static_cast<uint32_t>(run));
		}
		// buffers start on 8-byte boundaries of the stream (not 16): two 8-byte stores
		uint2 *o = reinterpret_cast<uint2 *>(stream + t.dst) + 2 * i;
		st8(o, make_uint2(v.x, v.y));
		st8(o + 1, make_uint2(v.z, v.w));
	}
	if (row0 + B2A_ENC_STR_TILE >= t.rows) { // last tile: zero padding of the data buffer up to 8 bytes
		const uint64_t total = tile_base + window;
		zero_range(stream, t.dst2 + total, t.dst2 + align8(total));
	}
	if (window == 0) {
		return;
	}
	uint8_t *gdst = stream + t.dst2 + tile_base;
	if (window + 48 > kStageBytes) {
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			if (len[u]) {
				put_string(gdst + rel[u], e[u], len[u]);
			}
		}
		return;
	}
	const uint32_t pd = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(gdst) & 15);
#pragma unroll
	for (int u = 0; u < kRowsPerThread; u++) {
		if (len[u]) {
			put_string_words(stage + pd + rel[u], e[u], len[u]);
		}
	}
	__syncthreads();
	enc_stage_flush(stage, pd, gdst, window);
}

// ArrowListData<BUFTYPE>::AppendOffsets: offsets only (children are appended through the selection built by
// b2a_enc_expand_kernel)
// VIEW: ArrowListViewData<BUFTYPE>::AppendListMetadata — offsets[i] = the running offset BEFORE row i, sizes[i] = its
// length (0 for a NULL row), no trailing offset; the child selection is the same
template <typename OFF, bool VIEW = false>
__device__ __forceinline__ void enc_list(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint32_t k,
                                         uint32_t *err) {
	const uint64_t row0 = static_cast<uint64_t>(k) * B2A_ENC_STR_TILE;
	uint64_t len[kRowsPerThread], sum = 0;
#pragma unroll
	for (int u = 0; u < kRowsPerThread; u++) {
		const uint64_t i = row0 + threadIdx.x * kRowsPerThread + u;
		len[u] = 0;
		if (i < t.rows) {
			uint4 e;
			bool valid;
			len[u] = entry_len(t, i, e, valid);
		}
		sum += len[u];
	}
	uint64_t window;
	uint64_t run = reinterpret_cast<const uint64_t *>(t.scratch)[k] + block_exclusive_scan(sum, &window);
	OFF *offs = reinterpret_cast<OFF *>(stream + t.dst);
	if (VIEW) {
		OFF *sizes = reinterpret_cast<OFF *>(stream + t.dst2);
#pragma unroll
		for (int u = 0; u < kRowsPerThread; u++) {
			const uint64_t i = row0 + threadIdx.x * kRowsPerThread + u;
			if (i < t.rows) {
				if (sizeof(OFF) == 4 && run + len[u] > 0x7FFFFFFFull) {
					enc_raise(err, B2A_DEV_ENC_LIST_I32);
				}
				offs[i] = static_cast<OFF>(run);
				sizes[i] = static_cast<OFF>(len[u]);
				if (sizeof(OFF) == 4 && i + 1 == t.rows && (t.rows & 1)) {
					sizes[i + 1] = 0; // zero padding of the sizes buffer up to 8 bytes (the main buffer: last-tile code)
				}
			}
			run += len[u];
		}
		return;
	}
	if (k == 0 && threadIdx.x == 0) {
		offs[0] = 0;
	}
#pragma unroll
	for (int u = 0; u < kRowsPerThread; u++) {
		const uint64_t i = row0 + threadIdx.x * kRowsPerThread + u;
		run += len[u];
		if (i < t.rows) {
			if (sizeof(OFF) == 4 && run > 0x7FFFFFFFull) {
				enc_raise(err, B2A_DEV_ENC_LIST_I32);
			}
			offs[i + 1] = static_cast<OFF>(run);
		}
	}
}

// ArrowVarcharToStringViewData::Append (varchar_data.hpp:136-187): one 16-byte view per row — NULL: zeros; length <= 12:
// {length, bytes, zero padding}; longer: {length, first 4 bytes, buffer index 0, int32 offset} with the characters
// appended to the single data buffer (dst2).  Rare op: the long strings are copied one at a time per warp, bytes strided over the lanes.
#ifndef B2A_ENC_VIEW_COPIES
#define B2A_ENC_VIEW_COPIES 4
#endif
static constexpr int kViewCopies = B2A_ENC_VIEW_COPIES; // long strings a warp copies at a time (loads in flight)
__device__ __forceinline__ void enc_string_view(const b2a_enc_task &t, uint8_t *__restrict__ stream, uint32_t k,
                                                uint32_t *err) {
	const uint64_t row0 = static_cast<uint64_t>(k) * B2A_ENC_STR_TILE;
	uint4 e[kRowsPerThread];
	uint32_t len[kRowsPerThread]; // bytes that go to the data buffer (0 for NULL / inlined rows)
	bool valid[kRowsPerThread];
	uint64_t sum = 0;
#pragma unroll
	for (int u = 0; u < kRowsPerThread; u++) {
		const uint64_t i = row0 + threadIdx.x * kRowsPerThread + u;
		len[u] = 0;
		valid[u] = false;
		e[u] = make_uint4(0, 0, 0, 0);
		if (i < t.rows) {
			len[u] = static_cast<uint32_t>(entry_len(t, i, e[u], valid[u]));
		}
		sum += len[u];
	}
	uint64_t window;
	const uint64_t tile_base = reinterpret_cast<const uint64_t *>(t.scratch)[k];
	uint64_t run = tile_base + block_exclusive_scan(sum, &window);
	unsigned long long cp_src[kRowsPerThread], cp_dst[kRowsPerThread];
	uint32_t cp_len[kRowsPerThread];
#pragma unroll
	for (int u = 0; u < kRowsPerThread; u++) {
		const uint64_t i = row0 + threadIdx.x * kRowsPerThread + u;
		cp_len[u] = 0;
		cp_src[u] = cp_dst[u] = 0;
		if (i >= t.rows) {
			continue;
		}
		uint4 v = make_uint4(0, 0, 0, 0);
		if (valid[u] && len[u] == 0) { // inlined: string_t and the view share the layout; bytes past the length are zeroed
			const uint32_t n = e[u].x;
			auto keep = [n](uint32_t w, uint32_t first) {
				return n >= first + 4 ? w : n <= first ? 0u : (w & ((1u << (8 * (n - first))) - 1u));
			};
			v = make_uint4(n, keep(e[u].y, 0), keep(e[u].z, 4), keep(e[u].w, 8));
		} else if (valid[u]) {
			if (run > 0x7FFFFFFFull) { // the view's offset is an int32 (UnsafeNumericCast<int32_t>(append_data.offset))
				enc_raise(err, B2A_DEV_ENC_OFFSET_I32);
			}
			v = make_uint4(e[u].x, e[u].y, 0u, static_cast<uint32_t>(run));
			cp_src[u] = (static_cast<unsigned long long>(e[u].w) << 32) | e[u].z;
			cp_dst[u] = t.dst2 + run;
			cp_len[u] = len[u];
			run += len[u];
		}
		// buffers start on 8-byte boundaries of the stream (not 16): two 8-byte stores
		uint2 *o = reinterpret_cast<uint2 *>(stream + t.dst) + 2 * i;
		o[0] = make_uint2(v.x, v.y);
		o[1] = make_uint2(v.z, v.w);
	}
	// the long strings: four at a time per warp (independent loads in flight), the bytes of each strided over the lanes
	// — coalesced loads and stores; bytes past the first 32 of a string follow in a per-string loop
	const uint32_t lane = threadIdx.x & 31u;
#pragma unroll
	for (int u = 0; u < kRowsPerThread; u++) {
		unsigned m = __ballot_sync(0xFFFFFFFFu, cp_len[u] != 0);
		while (m) {
			const uint8_t *src[kViewCopies];
			uint8_t *dst[kViewCopies];
			uint32_t n[kViewCopies];
			uint8_t byte[kViewCopies];
#pragma unroll
			for (int q = 0; q < kViewCopies; q++) {
				const int owner = m ? __ffs(m) - 1 : 0;
				const bool have = m != 0;
				m &= m - 1;
				src[q] = reinterpret_cast<const uint8_t *>(__shfl_sync(0xFFFFFFFFu, cp_src[u], owner));
				dst[q] = stream + __shfl_sync(0xFFFFFFFFu, cp_dst[u], owner);
				n[q] = have ? __shfl_sync(0xFFFFFFFFu, cp_len[u], owner) : (__shfl_sync(0xFFFFFFFFu, cp_len[u], owner), 0u);
			}
#pragma unroll
			for (int q = 0; q < kViewCopies; q++) {
				byte[q] = lane < n[q] ? __ldg(src[q] + lane) : 0;
			}
#pragma unroll
			for (int q = 0; q < kViewCopies; q++) {
				if (lane < n[q]) {
					dst[q][lane] = byte[q];
				}
				for (uint32_t b = lane + 32; b < n[q]; b += 32) {
					dst[q][b] = __ldg(src[q] + b);
				}
			}
		}
	}
	if (row0 + B2A_ENC_STR_TILE >= t.rows) { // last tile: zero padding of the data buffer up to 8 bytes
		const uint64_t total = tile_base + window;
		zero_range(stream, t.dst2 + total, t.dst2 + align8(total));
	}
}

// ------------------------------------------------------------------------------------------------
// the encode kernel: persistent CTAs stride over the tile table, next descriptor prefetched with cp.async
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void enc_cp_async16(void *smem, const void *gmem) {
	const uint32_t sa = static_cast<uint32_t>(__cvta_generic_to_shared(smem));
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void enc_cp_async_wait_all() {
	asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

static constexpr uint32_t kEncTaskVecs = sizeof(b2a_enc_task) / 16;
static_assert(sizeof(b2a_enc_task) % 16 == 0, "b2a_enc_task is copied in 16-byte pieces");

// FULL = false: the ops of the common types only (what a lineitem COPY runs); FULL = true adds the rare ones
// (TIME_TZ, UUID -> string, UNION type ids) — kept out of the common instantiation's switch like on the decode side
#ifndef B2A_ENC_MIN_BLOCKS
#define B2A_ENC_MIN_BLOCKS 4 // resident CTAs per SM the register budget is sized for (A/B: -DB2A_ENC_MIN_BLOCKS=5)
#endif
template <bool FULL>
__global__ void __launch_bounds__(kThreads, B2A_ENC_MIN_BLOCKS)
    b2a_encode_kernel(const b2a_enc_task *__restrict__ tasks, const uint32_t *__restrict__ tile_task, uint32_t n_tiles,
                      uint8_t *__restrict__ stream, uint32_t *err) {
	// descriptors: a ring of three — the current tile's, the next tile's (resident: the record prefetch reads it) and the
	// one after that (arriving by cp.async)
	__shared__ b2a_enc_task s_task[3];
	extern __shared__ __align__(128) uint8_t enc_dyn_smem[]; // kEncDynSmem bytes: character stage, then the record area
	uint8_t *const stage = enc_dyn_smem;
	uint8_t *const s_rec = enc_dyn_smem + kEncStageArea;
	__shared__ __align__(8) uint64_t s_bar;     // completion of the character bulk copies of enc_string2
	__shared__ __align__(8) uint64_t s_bar_rec; // completion of the record prefetch
	uint32_t bulk_parity = 0;
	uint32_t tile = blockIdx.x;
	if (tile >= n_tiles) {
		return;
	}
	if (threadIdx.x == 0) {
		enc_mbar_init(&s_bar, 1);
		enc_mbar_init(&s_bar_rec, 1);
	}
	const uint64_t stride = gridDim.x;
	const bool loader = threadIdx.x < kEncTaskVecs;
	uint32_t ti_next2 = 0; // task of tile + 2 * stride (loader threads)
	if (loader) {
		const uint32_t ti = __ldg(tile_task + tile);
		reinterpret_cast<uint4 *>(&s_task[0])[threadIdx.x] =
		    __ldg(reinterpret_cast<const uint4 *>(tasks + ti) + threadIdx.x);
		if (tile + stride < n_tiles) {
			const uint32_t ti1 = __ldg(tile_task + tile + stride);
			reinterpret_cast<uint4 *>(&s_task[1])[threadIdx.x] =
			    __ldg(reinterpret_cast<const uint4 *>(tasks + ti1) + threadIdx.x);
		}
		if (tile + 2 * stride < n_tiles) {
			ti_next2 = __ldg(tile_task + tile + 2 * stride);
		}
	}
	EncRecPipe rp;
	rp.rec = s_rec;
	rp.bar = &s_bar_rec;
	rp.parity = 0;
	rp.ready = false; // the CTA's first tile loads its records itself
	__syncthreads();
	for (uint32_t slot = 0;; slot = slot == 2 ? 0 : slot + 1) {
		const uint64_t next = tile + stride, next2 = tile + 2 * stride;
		if (loader && next2 < n_tiles) {
			enc_cp_async16(reinterpret_cast<uint4 *>(&s_task[slot == 0 ? 2 : slot - 1]) + threadIdx.x,
			               reinterpret_cast<const uint4 *>(tasks + ti_next2) + threadIdx.x);
			if (next2 + stride < n_tiles) {
				ti_next2 = __ldg(tile_task + next2 + stride);
			}
		}
		const b2a_enc_task &t = s_task[slot];
		rp.next = next < n_tiles ? &s_task[slot == 2 ? 0 : slot + 1] : nullptr;
		rp.next_tile = static_cast<uint32_t>(next);
		rp.next_ready = false;
		if (!rp.ready) { // the record area is not in use by this tile: the next tile's records may start right away
			rp.next_ready = enc_prefetch_records(rp);
		}
		const uint32_t k = tile - t.tile0;
		const uint64_t row0 = static_cast<uint64_t>(k) * t.tile_rows;
		uint64_t nrows = t.rows > row0 ? t.rows - row0 : 0;
		if (nrows > t.tile_rows) {
			nrows = t.tile_rows;
		}
		switch (t.op) {
		case B2A_ENC_VALIDITY:
		case B2A_ENC_FILL_VALID:
			enc_validity(t, stream, row0, nrows);
			break;
		case B2A_ENC_COPY:
			enc_copy(t, stream, row0, nrows);
			break;
		case B2A_ENC_WIDEN16:
		case B2A_ENC_INTERVAL:
			enc_convert(t, stream, row0, nrows);
			break;
		case B2A_ENC_BOOL:
			enc_bool(t, stream, row0, nrows);
			break;
#ifdef B2A_ENC_STRING_V1 // A/B builds only (tools/build_variants.py): the first-generation string path
		case B2A_ENC_STRING32:
			enc_string<int32_t>(t, stream, k, stage, err);
			break;
		case B2A_ENC_STRING64:
			enc_string<int64_t>(t, stream, k, stage, err);
			break;
#else
		case B2A_ENC_STRING32:
			enc_string2<int32_t>(t, stream, k, stage, err, &s_bar, bulk_parity, rp);
			break;
		case B2A_ENC_STRING64:
			enc_string2<int64_t>(t, stream, k, stage, err, &s_bar, bulk_parity, rp);
			break;
#endif
		case B2A_ENC_LIST32:
			enc_list<int32_t>(t, stream, k, err);
			break;
		case B2A_ENC_LIST64:
			enc_list<int64_t>(t, stream, k, err);
			break;
		default:
			if (FULL) {
				if (t.op == B2A_ENC_TIMETZ) {
					enc_timetz(t, stream, row0, nrows);
				} else if (t.op == B2A_ENC_UNION_TAG) {
					enc_union_tag(t, stream, row0, nrows);
				} else if (t.op == B2A_ENC_UUID_BLOB) {
					enc_uuid_blob(t, stream, row0, nrows);
				} else if (t.op == B2A_ENC_UUID32) {
					enc_string<int32_t, true>(t, stream, k, stage, err);
				} else if (t.op == B2A_ENC_UUID64) {
					enc_string<int64_t, true>(t, stream, k, stage, err);
				} else if (t.op == B2A_ENC_LISTVIEW32) {
					enc_list<int32_t, true>(t, stream, k, err);
				} else if (t.op == B2A_ENC_LISTVIEW64) {
					enc_list<int64_t, true>(t, stream, k, err);
				} else if (t.op == B2A_ENC_STRINGVIEW) {
					enc_string_view2(t, stream, k, stage, err, rp);
				}
			}
			break;
		}
		if (row0 + t.tile_rows >= t.rows) { // last tile of the task: zero padding of the main buffer up to 8 bytes
			zero_range(stream, t.dst + t.dst_len, t.dst + align8(t.dst_len));
		}
		if (next >= n_tiles) {
			break;
		}
		enc_cp_async_wait_all();
		__syncthreads();
		tile = static_cast<uint32_t>(next);
		rp.ready = rp.next_ready;
	}
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
static uint32_t PersistentGrid(const void *kernel, int sm_count, uint32_t n_tiles, size_t dyn_smem = 0) {
	int per_sm = 0;
	if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kThreads, dyn_smem) != cudaSuccess || per_sm <= 0) {
		per_sm = 2;
	}
	uint32_t grid = static_cast<uint32_t>(sm_count) * static_cast<uint32_t>(per_sm);
	return grid > n_tiles ? n_tiles : grid;
}

// more than 48 KiB of shared memory per CTA is an opt-in, per kernel and per device
static int EncodeSmemOptIn() {
	static bool done[64][2] = {};
	int dev = 0;
	if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
		dev = 0;
	}
	if (!done[dev][0]) {
		cudaError_t e = cudaFuncSetAttribute(b2a_encode_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
		                                     static_cast<int>(kEncDynSmem));
		if (e != cudaSuccess) {
			return static_cast<int>(e);
		}
		e = cudaFuncSetAttribute(b2a_encode_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
		                         static_cast<int>(kEncDynSmem));
		if (e != cudaSuccess) {
			return static_cast<int>(e);
		}
		done[dev][0] = true;
	}
	return 0;
}

int LaunchEncMeasure(const EncodeLaunch &l, void *stream) {
	cudaStream_t st = static_cast<cudaStream_t>(stream);
	if (l.n_tiles) {
		uint32_t grid = static_cast<uint32_t>(l.sm_count) * 8;
		grid = grid > l.n_tiles ? l.n_tiles : grid;
		b2a_enc_measure_kernel<<<grid, kThreads, 0, st>>>(l.d_tasks, l.d_tile_task, l.n_tiles,
		                                                  reinterpret_cast<unsigned long long *>(l.d_heap_bytes));
	}
	if (l.n_tasks) {
		uint32_t grid = l.n_tasks < 65535u ? l.n_tasks : 65535u;
		b2a_enc_scan_kernel<<<grid, 32, 0, st>>>(l.d_tasks, l.n_tasks);
	}
	return static_cast<int>(cudaGetLastError());
}

int LaunchEncExpand(const EncodeLaunch &l, void *stream) {
	if (l.n_tiles == 0) {
		return 0;
	}
	uint32_t grid = static_cast<uint32_t>(l.sm_count) * 8;
	grid = grid > l.n_tiles ? l.n_tiles : grid;
	b2a_enc_expand_kernel<<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(l.d_tasks, l.d_tile_task,
	                                                                              l.n_tiles);
	return static_cast<int>(cudaGetLastError());
}

int LaunchEncode(const EncodeLaunch &l, void *stream) {
	if (l.n_tiles == 0) {
		return 0;
	}
	if (int e = EncodeSmemOptIn()) {
		return e;
	}
	if (l.full) {
		const uint32_t grid =
		    PersistentGrid(reinterpret_cast<const void *>(b2a_encode_kernel<true>), l.sm_count, l.n_tiles, kEncDynSmem);
		b2a_encode_kernel<true><<<grid, kThreads, kEncDynSmem, static_cast<cudaStream_t>(stream)>>>(
		    l.d_tasks, l.d_tile_task, l.n_tiles, l.stream_out, l.d_err);
	} else {
		const uint32_t grid =
		    PersistentGrid(reinterpret_cast<const void *>(b2a_encode_kernel<false>), l.sm_count, l.n_tiles, kEncDynSmem);
		b2a_encode_kernel<false><<<grid, kThreads, kEncDynSmem, static_cast<cudaStream_t>(stream)>>>(
		    l.d_tasks, l.d_tile_task, l.n_tiles, l.stream_out, l.d_err);
	}
	return static_cast<int>(cudaGetLastError());
}

// one thread per input row of the UNION vector, one ballot per 32 rows -> the 32-bit halves of the member's
// ValidityMask words: union row valid AND tag == member AND member value valid
__global__ void __launch_bounds__(kThreads)
    b2a_enc_union_mask_kernel(const uint8_t *__restrict__ tags, const uint64_t *__restrict__ union_valid,
                              const uint64_t *__restrict__ member_valid, uint32_t member, uint64_t rows,
                              uint32_t *__restrict__ out) {
	const uint64_t padded = (rows + 63) & ~uint64_t(63);
	for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * kThreads + threadIdx.x; i < padded;
	     i += static_cast<uint64_t>(gridDim.x) * kThreads) {
		bool v = false;
		if (i < rows) {
			v = __ldg(tags + i) == member;
			if (union_valid) {
				v = v && ((__ldg(union_valid + (i >> 6)) >> (i & 63)) & 1);
			}
			if (member_valid) {
				v = v && ((__ldg(member_valid + (i >> 6)) >> (i & 63)) & 1);
			}
		}
		const uint32_t word = __ballot_sync(0xffffffffu, v);
		if ((threadIdx.x & 31) == 0) {
			out[i >> 5] = word;
		}
	}
}

int LaunchEncUnionMask(const void *tags, const void *union_valid, const void *member_valid, uint32_t member,
                       uint64_t rows, void *out_words, int sm_count, void *stream) {
	if (rows == 0) {
		return 0;
	}
	uint64_t blocks = (rows + kThreads - 1) / kThreads;
	const uint64_t cap = static_cast<uint64_t>(sm_count) * 8;
	b2a_enc_union_mask_kernel<<<static_cast<uint32_t>(blocks > cap ? cap : blocks), kThreads, 0,
	                            static_cast<cudaStream_t>(stream)>>>(
	    static_cast<const uint8_t *>(tags), static_cast<const uint64_t *>(union_valid),
	    static_cast<const uint64_t *>(member_valid), member, rows, static_cast<uint32_t *>(out_words));
	return static_cast<int>(cudaGetLastError());
}

} // namespace b2a
