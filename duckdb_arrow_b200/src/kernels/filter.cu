// K9: pushed-down predicates -> selection vectors (one per 2 048-row DuckDB vector) with warp-level stream
// compaction, and a Q6-style consumer (sum of products over the selected rows).
// Not in the reference (filter_pushdown = false, src/scanner/read_arrow.cpp:46-48); semantics = DuckDB's
// ConstantFilter on the scanned vectors (duckdb/src/include/duckdb/planner/filter/constant_filter.hpp:17-38):
// a row passes iff it is valid and `value <cmp> constant` holds for every predicate; sel_t = uint32 row index
// inside the vector, ascending (SelectionVector as produced by a DuckDB filter).
#include "decode.cuh"

#include <cuda_runtime.h>

namespace b2a {

static constexpr int kFilterThreads = 256;

struct FilterParams {
	b2a_predicate preds[8];
	uint8_t widths[8];
	uint32_t n_preds;
};

__device__ __forceinline__ bool cmp_i64(int64_t v, uint32_t cmp, int64_t c) {
	switch (cmp) {
	case B2A_CMP_EQ:
		return v == c;
	case B2A_CMP_NE:
		return v != c;
	case B2A_CMP_LT:
		return v < c;
	case B2A_CMP_LE:
		return v <= c;
	case B2A_CMP_GT:
		return v > c;
	default:
		return v >= c;
	}
}

// chunk_col: per chunk, per predicate: {arena offset of the chunk's first value, arena offset of the chunk's first
// validity word or B2A_NONE_OFF}
__global__ void __launch_bounds__(kFilterThreads)
    b2a_filter_kernel(const uint8_t *__restrict__ out, const uint64_t *__restrict__ chunk_col,
                      const uint32_t *__restrict__ chunk_rows, uint64_t n_chunks, FilterParams fp,
                      uint32_t *__restrict__ sel, uint32_t *__restrict__ count, unsigned long long *total) {
	__shared__ uint32_t warp_counts[kFilterThreads / 32];
	__shared__ uint32_t running;
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	for (uint64_t chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
		const uint32_t rows = chunk_rows[chunk];
		const uint64_t *cc = chunk_col + chunk * 2 * fp.n_preds;
		uint32_t *csel = sel + chunk * B2A_VECTOR_SIZE;
		if (threadIdx.x == 0) {
			running = 0;
		}
		__syncthreads();
		for (uint32_t base = 0; base < rows; base += kFilterThreads) {
			const uint32_t r = base + threadIdx.x;
			bool pass = r < rows;
			for (uint32_t p = 0; p < fp.n_preds && pass; p++) {
				const uint64_t doff = cc[2 * p], voff = cc[2 * p + 1];
				if (voff != B2A_NONE_OFF) {
					uint64_t w = reinterpret_cast<const uint64_t *>(out + voff)[r >> 6];
					pass = (w >> (r & 63)) & 1;
				}
				if (pass) {
					int64_t v;
					switch (fp.widths[p]) {
					case 1:
						v = reinterpret_cast<const int8_t *>(out + doff)[r];
						break;
					case 2:
						v = reinterpret_cast<const int16_t *>(out + doff)[r];
						break;
					case 4:
						v = reinterpret_cast<const int32_t *>(out + doff)[r];
						break;
					default:
						v = reinterpret_cast<const int64_t *>(out + doff)[r];
						break;
					}
					pass = cmp_i64(v, fp.preds[p].cmp, fp.preds[p].constant);
				}
			}
			// warp-level compaction: ballot + popc prefix
			const uint32_t ballot = __ballot_sync(0xffffffffu, pass);
			const uint32_t before = __popc(ballot & ((1u << lane) - 1u));
			if (lane == 0) {
				warp_counts[warp] = __popc(ballot);
			}
			__syncthreads();
			uint32_t woff = running;
			for (uint32_t w = 0; w < warp; w++) {
				woff += warp_counts[w];
			}
			if (pass) {
				csel[woff + before] = r;
			}
			__syncthreads();
			if (threadIdx.x == 0) {
				uint32_t s = 0;
				for (uint32_t w = 0; w < kFilterThreads / 32; w++) {
					s += warp_counts[w];
				}
				running += s;
			}
			__syncthreads();
		}
		if (threadIdx.x == 0) {
			count[chunk] = running;
			if (running) {
				atomicAdd(total, static_cast<unsigned long long>(running));
			}
		}
		__syncthreads();
	}
}

int LaunchFilter(const FilterLaunch &l, void *stream) {
	if (l.n_chunks == 0) {
		return 0;
	}
	FilterParams fp;
	fp.n_preds = l.n_preds;
	for (uint32_t i = 0; i < l.n_preds; i++) {
		fp.preds[i] = l.preds[i];
		fp.widths[i] = l.widths[i];
	}
	uint64_t grid = static_cast<uint64_t>(l.sm_count) * 8;
	if (grid > l.n_chunks) {
		grid = l.n_chunks;
	}
	b2a_filter_kernel<<<static_cast<uint32_t>(grid), kFilterThreads, 0, static_cast<cudaStream_t>(stream)>>>(
	    l.out, l.d_chunk_col, l.d_chunk_rows, l.n_chunks, fp, l.d_sel, l.d_count, l.d_total);
	return static_cast<int>(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------
// sum(a*b) over selected rows, int64 x int64 -> int128 (DuckDB: DECIMAL(15,2)*DECIMAL(15,2) summed as HUGEINT)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kFilterThreads)
    b2a_sum_product_kernel(const uint8_t *__restrict__ out, const uint64_t *__restrict__ chunk_col,
                           const uint32_t *__restrict__ chunk_rows, const uint32_t *__restrict__ sel,
                           const uint32_t *__restrict__ count, uint64_t n_chunks, unsigned long long *result) {
	__int128 acc = 0;
	for (uint64_t chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
		const int64_t *a = reinterpret_cast<const int64_t *>(out + chunk_col[4 * chunk]);
		const int64_t *b = reinterpret_cast<const int64_t *>(out + chunk_col[4 * chunk + 2]);
		const uint32_t n = sel ? count[chunk] : chunk_rows[chunk];
		const uint32_t *csel = sel ? sel + chunk * B2A_VECTOR_SIZE : nullptr;
		for (uint32_t k = threadIdx.x; k < n; k += kFilterThreads) {
			const uint32_t r = csel ? csel[k] : k;
			acc += static_cast<__int128>(a[r]) * static_cast<__int128>(b[r]);
		}
	}
	uint64_t lo = static_cast<uint64_t>(acc), hi = static_cast<uint64_t>(acc >> 64);
	for (int o = 16; o > 0; o >>= 1) {
		uint64_t olo = __shfl_xor_sync(0xffffffffu, lo, o), ohi = __shfl_xor_sync(0xffffffffu, hi, o);
		uint64_t nlo = lo + olo;
		hi = hi + ohi + (nlo < lo ? 1 : 0);
		lo = nlo;
	}
	if ((threadIdx.x & 31) == 0 && (lo | hi)) {
		unsigned long long old = atomicAdd(result, static_cast<unsigned long long>(lo));
		unsigned long long carry = (old + lo) < old ? 1ull : 0ull;
		atomicAdd(result + 1, static_cast<unsigned long long>(hi) + carry);
	}
}

int LaunchSumProduct(const SumProductLaunch &l, void *stream) {
	if (l.n_chunks == 0) {
		return 0;
	}
	uint64_t grid = static_cast<uint64_t>(l.sm_count) * 8;
	if (grid > l.n_chunks) {
		grid = l.n_chunks;
	}
	b2a_sum_product_kernel<<<static_cast<uint32_t>(grid), kFilterThreads, 0, static_cast<cudaStream_t>(stream)>>>(
	    l.out, l.d_chunk_col, l.d_chunk_rows, l.d_sel, l.d_count, l.n_chunks, l.d_result);
	return static_cast<int>(cudaGetLastError());
}

} // namespace b2a
