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
	uint8_t same_column[8]; // predicate p reads the column of predicate p - 1: its values are in registers already
	uint32_t n_preds;
};

// chunk_col: per chunk, per predicate: {arena offset of the chunk's first value, arena offset of the chunk's first
// validity word or B2A_NONE_OFF}
//
// One CTA per 2 048-row vector, 8 rows per thread (row = k * 256 + thread, k = 0..7, so that a warp reads 32 adjacent
// values).  All values of a predicate column are requested before the first comparison (8 loads in flight per thread and
// column; no short-circuit between predicates — the bytes are cheap, a dependent second round of loads is not), every
// k gives one ballot, the 64 (k, warp) counts are scanned by warp 0, and the passing rows are written in ascending row
// order: three barriers per vector.
static constexpr int kFilterRowsPerThread = B2A_VECTOR_SIZE / kFilterThreads; // 8

// One run of predicates on ONE column (the host sorts the predicates by column): the vector's values are loaded once, in the
// column's own type — a 4-byte date is compared as 4 bytes, not widened to 64 bits first —, masked by the column's NULLs,
// and every predicate of the run is a 3-bit truth table over {v < c, v == c, v > c} (bits 0, 1, 2) looked up per value: no
// branch on the operator.  A constant outside the type's range decides the predicate for the whole vector.
// Returns the index of the first predicate behind the run.  FULL: the vector holds all 2 048 rows (no bounds checks).
template <typename T, bool FULL>
__device__ __forceinline__ uint32_t filter_run(const uint8_t *__restrict__ out, const uint64_t *cc, uint32_t rows,
                                               const FilterParams &fp, uint32_t p, uint32_t &pass) {
	const T *col = reinterpret_cast<const T *>(out + cc[2 * p]);
	const uint64_t voff = cc[2 * p + 1];
	T v[kFilterRowsPerThread];
#pragma unroll
	for (int k = 0; k < kFilterRowsPerThread; k++) {
		const uint32_t r = k * kFilterThreads + threadIdx.x;
		v[k] = (FULL || r < rows) ? __ldg(col + r) : T(0);
	}
	if (voff != B2A_NONE_OFF) {
		const unsigned long long *mask = reinterpret_cast<const unsigned long long *>(out + voff);
#pragma unroll
		for (int k = 0; k < kFilterRowsPerThread; k++) {
			const uint32_t r = k * kFilterThreads + threadIdx.x;
			if (FULL || r < rows) {
				const uint64_t w = __ldg(mask + (r >> 6));
				if (!((w >> (r & 63)) & 1)) {
					pass &= ~(1u << k);
				}
			}
		}
	}
	constexpr bool kSigned = T(-1) < T(0);
	constexpr int64_t kLo = !kSigned ? 0 : sizeof(T) == 8 ? INT64_MIN : -(int64_t(1) << (8 * (sizeof(T) & 7) - 1));
	constexpr int64_t kHi = sizeof(T) == 8 ? INT64_MAX
	                        : kSigned      ? (int64_t(1) << (8 * (sizeof(T) & 7) - 1)) - 1
	                                       : (int64_t(1) << (8 * (sizeof(T) & 7))) - 1;
	do {
		const uint32_t cmp = fp.preds[p].cmp;
		const uint32_t accept = cmp == B2A_CMP_EQ   ? 2u
		                        : cmp == B2A_CMP_NE ? 5u
		                        : cmp == B2A_CMP_LT ? 1u
		                        : cmp == B2A_CMP_LE ? 3u
		                        : cmp == B2A_CMP_GT ? 4u
		                                            : 6u;
		const int64_t c = fp.preds[p].constant;
		if (sizeof(T) < 8 && c > kHi) { // every value is below the constant
			pass = (accept & 1u) ? pass : 0u;
		} else if (sizeof(T) < 8 && c < kLo) { // every value is above it
			pass = (accept & 4u) ? pass : 0u;
		} else {
			const T ct = static_cast<T>(c);
#pragma unroll
			for (int k = 0; k < kFilterRowsPerThread; k++) {
				const uint32_t which = v[k] < ct ? 0u : (v[k] == ct ? 1u : 2u);
				pass &= ~((((accept >> which) & 1u) ^ 1u) << k);
			}
		}
		p++;
	} while (p < fp.n_preds && fp.same_column[p]);
	return p;
}

// bit k of the result: row k * 256 + thread of the vector is valid and passes every predicate.  cc = the vector's table
// entries {data offset, validity offset} per predicate (shared memory).
template <bool FULL>
__device__ __forceinline__ uint32_t filter_pass_bits_t(const uint8_t *__restrict__ out, const uint64_t *cc, uint32_t rows,
                                                       const FilterParams &fp) {
	uint32_t pass = 0xFFu; // bit k: row k * 256 + thread passes
	if (!FULL) {
		pass = 0;
#pragma unroll
		for (int k = 0; k < kFilterRowsPerThread; k++) {
			pass |= (k * kFilterThreads + threadIdx.x < rows ? 1u : 0u) << k;
		}
	}
	uint32_t p = 0;
	while (p < fp.n_preds) {
		switch (fp.widths[p]) { // 0x80 | width: unsigned column (UTINYINT / USMALLINT / UINTEGER)
		case 1:
			p = filter_run<int8_t, FULL>(out, cc, rows, fp, p, pass);
			break;
		case 2:
			p = filter_run<int16_t, FULL>(out, cc, rows, fp, p, pass);
			break;
		case 4:
			p = filter_run<int32_t, FULL>(out, cc, rows, fp, p, pass);
			break;
		case 0x81:
			p = filter_run<uint8_t, FULL>(out, cc, rows, fp, p, pass);
			break;
		case 0x82:
			p = filter_run<uint16_t, FULL>(out, cc, rows, fp, p, pass);
			break;
		case 0x84:
			p = filter_run<uint32_t, FULL>(out, cc, rows, fp, p, pass);
			break;
		default:
			p = filter_run<int64_t, FULL>(out, cc, rows, fp, p, pass);
			break;
		}
	}
	return pass;
}

__device__ __forceinline__ uint32_t filter_pass_bits(const uint8_t *__restrict__ out, const uint64_t *cc, uint32_t rows,
                                                     const FilterParams &fp) {
	return rows == B2A_VECTOR_SIZE ? filter_pass_bits_t<true>(out, cc, rows, fp) : filter_pass_bits_t<false>(out, cc, rows, fp);
}

__global__ void __launch_bounds__(kFilterThreads, 8)
    b2a_filter_kernel(const uint8_t *__restrict__ out, const uint64_t *__restrict__ chunk_col,
                      const uint32_t *__restrict__ chunk_rows, uint64_t n_chunks, FilterParams fp,
                      uint32_t *__restrict__ sel, uint32_t *__restrict__ count, unsigned long long *total) {
	constexpr int kWarps = kFilterThreads / 32;
	__shared__ uint32_t counts[kFilterRowsPerThread * kWarps]; // (k, warp) -> passing rows, then exclusive prefix
	__shared__ uint32_t chunk_total;
	// the vector's table entries ({data, validity} offset per predicate + its row count), fetched one vector ahead by the
	// first threads of the CTA: no vector starts with an exposed table -> data dependent-load pair
	__shared__ uint64_t s_tbl[2][16];
	__shared__ uint32_t s_rows[2];
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint32_t n_entries = 2 * fp.n_preds;
	if (blockIdx.x < n_chunks) {
		if (threadIdx.x < n_entries) {
			s_tbl[0][threadIdx.x] = __ldg(chunk_col + static_cast<uint64_t>(blockIdx.x) * n_entries + threadIdx.x);
		}
		if (threadIdx.x == 32) {
			s_rows[0] = __ldg(chunk_rows + blockIdx.x);
		}
	}
	__syncthreads();
	uint32_t cur = 0;
	for (uint64_t chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x, cur ^= 1u) {
		const uint64_t next = chunk + gridDim.x;
		if (next < n_chunks) { // (visible after this iteration's barriers; the buffer was last read two vectors ago)
			if (threadIdx.x < n_entries) {
				s_tbl[cur ^ 1u][threadIdx.x] = __ldg(chunk_col + next * n_entries + threadIdx.x);
			}
			if (threadIdx.x == 32) {
				s_rows[cur ^ 1u] = __ldg(chunk_rows + next);
			}
		}
		const uint32_t rows = s_rows[cur];
		const uint64_t *cc = s_tbl[cur];
		uint32_t *csel = sel + chunk * B2A_VECTOR_SIZE;
		const uint32_t pass = filter_pass_bits(out, cc, rows, fp); // bit k: row k * 256 + thread passes
		// ---- compaction: ascending row order = k-major, then warp, then lane
		uint32_t before[kFilterRowsPerThread];
#pragma unroll
		for (int k = 0; k < kFilterRowsPerThread; k++) {
			const uint32_t ballot = __ballot_sync(0xffffffffu, (pass >> k) & 1u);
			before[k] = __popc(ballot & ((1u << lane) - 1u));
			if (lane == 0) {
				counts[k * kWarps + warp] = __popc(ballot);
			}
		}
		__syncthreads();
		if (warp == 0) { // exclusive scan of the 64 counts: two per lane
			const uint32_t a = counts[2 * lane], b = counts[2 * lane + 1];
			uint32_t inc = a + b;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
				if (lane >= static_cast<uint32_t>(o)) {
					inc += n;
				}
			}
			counts[2 * lane] = inc - a - b;
			counts[2 * lane + 1] = inc - b;
			if (lane == 31) {
				chunk_total = inc;
			}
		}
		__syncthreads();
#pragma unroll
		for (int k = 0; k < kFilterRowsPerThread; k++) {
			if ((pass >> k) & 1u) {
				csel[counts[k * kWarps + warp] + before[k]] = k * kFilterThreads + threadIdx.x;
			}
		}
		if (threadIdx.x == 0) {
			const uint32_t n = chunk_total;
			count[chunk] = n;
			if (n) {
				atomicAdd(total, static_cast<unsigned long long>(n));
			}
		}
		__syncthreads(); // counts / chunk_total are reused by the next vector
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
		fp.same_column[i] = i > 0 && l.preds[i].column == l.preds[i - 1].column;
	}
	// one wave of resident CTAs striding over the vectors (a grid of 8 per SM at 6 resident left a third wave-tail)
	static int per_sm_cached[64] = {};
	int dev = 0;
	if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
		dev = 0;
	}
	if (per_sm_cached[dev] == 0) {
		int per_sm = 0;
		if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, b2a_filter_kernel, kFilterThreads, 0) != cudaSuccess ||
		    per_sm <= 0) {
			per_sm = 4;
		}
		per_sm_cached[dev] = per_sm;
	}
	uint64_t grid = static_cast<uint64_t>(l.sm_count) * static_cast<uint64_t>(per_sm_cached[dev]);
	if (grid > l.n_chunks) {
		grid = l.n_chunks;
	}
	b2a_filter_kernel<<<static_cast<uint32_t>(grid), kFilterThreads, 0, static_cast<cudaStream_t>(stream)>>>(
	    l.out, l.d_chunk_col, l.d_chunk_rows, l.n_chunks, fp, l.d_sel, l.d_count, l.d_total);
	return static_cast<int>(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------
// Filter + aggregate in one kernel (the Q6 shape): count and sum(a * b) over the rows that pass, with no selection vector in
// between — b2a_filter_kernel writes sel_t / counts that b2a_sum_product_kernel reads back, and the pair is two launches.
// Per vector: the predicate bits, then the two factors of the passing rows (NULL factors are skipped, as DuckDB's sum
// skips a NULL product); per CTA: one warp reduction and one pair of atomics at the very end.  Table per vector =
// n_preds {data, validity} pairs followed by the pairs of a and b.  One barrier per vector (the table prefetch).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kFilterThreads, 6)
    b2a_filter_sum_kernel(const uint8_t *__restrict__ out, const uint64_t *__restrict__ chunk_col,
                          const uint32_t *__restrict__ chunk_rows, uint64_t n_chunks, FilterParams fp,
                          unsigned long long *total, unsigned long long *result) {
	__shared__ uint64_t s_tbl[2][20];
	__shared__ uint32_t s_rows[2];
	const uint32_t n_entries = 2 * (fp.n_preds + 2);
	if (blockIdx.x < n_chunks) {
		if (threadIdx.x < n_entries) {
			s_tbl[0][threadIdx.x] = __ldg(chunk_col + static_cast<uint64_t>(blockIdx.x) * n_entries + threadIdx.x);
		}
		if (threadIdx.x == 32) {
			s_rows[0] = __ldg(chunk_rows + blockIdx.x);
		}
	}
	__syncthreads();
	__int128 acc = 0;
	uint32_t n_pass = 0;
	uint32_t cur = 0;
	for (uint64_t chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x, cur ^= 1u) {
		const uint64_t next = chunk + gridDim.x;
		if (next < n_chunks) {
			if (threadIdx.x < n_entries) {
				s_tbl[cur ^ 1u][threadIdx.x] = __ldg(chunk_col + next * n_entries + threadIdx.x);
			}
			if (threadIdx.x == 32) {
				s_rows[cur ^ 1u] = __ldg(chunk_rows + next);
			}
		}
		const uint32_t rows = s_rows[cur];
		const uint64_t *cc = s_tbl[cur];
		uint32_t pass = filter_pass_bits(out, cc, rows, fp);
		if (pass) {
			const uint64_t *ab = cc + 2 * fp.n_preds;
			const int64_t *a = reinterpret_cast<const int64_t *>(out + ab[0]);
			const int64_t *b = reinterpret_cast<const int64_t *>(out + ab[2]);
			const unsigned long long *va =
			    ab[1] != B2A_NONE_OFF ? reinterpret_cast<const unsigned long long *>(out + ab[1]) : nullptr;
			const unsigned long long *vb =
			    ab[3] != B2A_NONE_OFF ? reinterpret_cast<const unsigned long long *>(out + ab[3]) : nullptr;
#pragma unroll
			for (int k = 0; k < kFilterRowsPerThread; k++) {
				if ((pass >> k) & 1u) {
					const uint32_t r = k * kFilterThreads + threadIdx.x;
					const bool ok = (!va || ((__ldg(va + (r >> 6)) >> (r & 63)) & 1)) &&
					                (!vb || ((__ldg(vb + (r >> 6)) >> (r & 63)) & 1));
					if (ok) {
						acc += static_cast<__int128>(__ldg(a + r)) * static_cast<__int128>(__ldg(b + r));
					}
				}
			}
			n_pass += __popc(pass);
		}
		__syncthreads(); // the next vector's table entries are in place; this vector's are free to be overwritten
	}
	uint64_t lo = static_cast<uint64_t>(acc), hi = static_cast<uint64_t>(acc >> 64);
	for (int o = 16; o > 0; o >>= 1) {
		const uint64_t olo = __shfl_xor_sync(0xffffffffu, lo, o), ohi = __shfl_xor_sync(0xffffffffu, hi, o);
		const uint64_t nlo = lo + olo;
		hi = hi + ohi + (nlo < lo ? 1 : 0);
		lo = nlo;
		n_pass += __shfl_xor_sync(0xffffffffu, n_pass, o);
	}
	if ((threadIdx.x & 31) == 0) {
		if (n_pass) {
			atomicAdd(total, static_cast<unsigned long long>(n_pass));
		}
		if (lo | hi) {
			const unsigned long long old = atomicAdd(result, static_cast<unsigned long long>(lo));
			const unsigned long long carry = (old + lo) < old ? 1ull : 0ull;
			atomicAdd(result + 1, static_cast<unsigned long long>(hi) + carry);
		}
	}
}

int LaunchFilterSum(const FilterLaunch &l, unsigned long long *d_result, void *stream) {
	if (l.n_chunks == 0) {
		return 0;
	}
	FilterParams fp;
	fp.n_preds = l.n_preds;
	for (uint32_t i = 0; i < l.n_preds; i++) {
		fp.preds[i] = l.preds[i];
		fp.widths[i] = l.widths[i];
		fp.same_column[i] = i > 0 && l.preds[i].column == l.preds[i - 1].column;
	}
	static int per_sm_cached[64] = {};
	int dev = 0;
	if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
		dev = 0;
	}
	if (per_sm_cached[dev] == 0) {
		int per_sm = 0;
		if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, b2a_filter_sum_kernel, kFilterThreads, 0) != cudaSuccess ||
		    per_sm <= 0) {
			per_sm = 4;
		}
		per_sm_cached[dev] = per_sm;
	}
	uint64_t grid = static_cast<uint64_t>(l.sm_count) * static_cast<uint64_t>(per_sm_cached[dev]);
	if (grid > l.n_chunks) {
		grid = l.n_chunks;
	}
	b2a_filter_sum_kernel<<<static_cast<uint32_t>(grid), kFilterThreads, 0, static_cast<cudaStream_t>(stream)>>>(
	    l.out, l.d_chunk_col, l.d_chunk_rows, l.n_chunks, fp, l.d_total, d_result);
	return static_cast<int>(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------
// No-allocation variant of the filter launch.  The kernel's per-chunk tables (where every 2 048-row vector of every
// predicate column starts in the arena, and how many rows it holds) are scratch the ordinary path keeps in cudaMalloc'ed
// memory; here they live in ONE file-scope device array of a fixed size that every GPU loading the library holds until
// the process exits.  In host code `g_filter_scratch` is only a stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, once per device, with that device current.  Every call shares the array and overwrites the part
// it needs (nothing in it has to start at zero: both tables are written in full before the kernel reads them, on the
// same stream); two calls that overlap on the GPU — different streams, different scans, different host threads — race.
// ------------------------------------------------------------------------------------------------
static constexpr uint64_t kFilterScratchBytes = 4ull << 20; // 4 MiB: 49 932 vectors (102 M rows) under 5 predicates
static __device__ __align__(256) unsigned char g_filter_scratch[kFilterScratchBytes];

uint64_t FilterNoAllocScratchBytes() {
	return kFilterScratchBytes;
}

// the column table first (8-byte entries), the row counts behind it on a 256-byte boundary
static uint64_t FilterRowsOffset(uint64_t n_chunks, uint32_t n_preds) {
	return (n_chunks * 2 * n_preds * sizeof(uint64_t) + 255) & ~uint64_t(255);
}

uint64_t FilterNoAllocBytesNeeded(uint64_t n_chunks, uint32_t n_preds) {
	return FilterRowsOffset(n_chunks, n_preds) + n_chunks * sizeof(uint32_t);
}

static int FilterScratchAddress(unsigned char **out) {
	static unsigned char *addr[64] = {};
	int dev = 0;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) {
		return static_cast<int>(e);
	}
	if (dev < 0 || dev >= 64) {
		return static_cast<int>(cudaErrorInvalidDevice);
	}
	if (!addr[dev]) { // (a race between two host threads stores the same value twice)
		void *p = nullptr;
		e = cudaGetSymbolAddress(&p, g_filter_scratch);
		if (e != cudaSuccess) {
			return static_cast<int>(e);
		}
		addr[dev] = static_cast<unsigned char *>(p);
	}
	*out = addr[dev];
	return 0;
}

int LaunchFilterNoAlloc(const FilterLaunch &l, const uint64_t *h_chunk_col, const uint32_t *h_chunk_rows, void *stream) {
	if (l.n_chunks == 0) {
		return 0;
	}
	if (FilterNoAllocBytesNeeded(l.n_chunks, l.n_preds) > kFilterScratchBytes) {
		return -1;
	}
	unsigned char *scratch = nullptr;
	int rc = FilterScratchAddress(&scratch);
	if (rc != 0) {
		return rc;
	}
	cudaStream_t st = static_cast<cudaStream_t>(stream);
	const uint64_t col_bytes = l.n_chunks * 2 * l.n_preds * sizeof(uint64_t);
	const uint64_t rows_off = FilterRowsOffset(l.n_chunks, l.n_preds);
	// pageable host memory: the copies are staged before the calls return, so the caller's tables may go away
	cudaError_t e = cudaMemcpyAsync(scratch, h_chunk_col, col_bytes, cudaMemcpyHostToDevice, st);
	if (e != cudaSuccess) {
		return static_cast<int>(e);
	}
	e = cudaMemcpyAsync(scratch + rows_off, h_chunk_rows, l.n_chunks * sizeof(uint32_t), cudaMemcpyHostToDevice, st);
	if (e != cudaSuccess) {
		return static_cast<int>(e);
	}
	FilterLaunch k = l;
	k.d_chunk_col = reinterpret_cast<const uint64_t *>(scratch);
	k.d_chunk_rows = reinterpret_cast<const uint32_t *>(scratch + rows_off);
	return LaunchFilter(k, stream);
}

// ------------------------------------------------------------------------------------------------
// sum(a*b) over selected rows, int64 x int64 -> int128 (DuckDB: DECIMAL(15,2)*DECIMAL(15,2) summed as HUGEINT)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kFilterThreads)
    b2a_sum_product_kernel(const uint8_t *__restrict__ out, const uint64_t *__restrict__ chunk_col,
                           const uint32_t *__restrict__ chunk_rows, const uint32_t *__restrict__ sel,
                           const uint32_t *__restrict__ count, uint64_t n_chunks, unsigned long long *result) {
	// one WARP per 2 048-row vector: a selective filter leaves a few rows per lane, and eight vectors per CTA keep eight
	// independent {table entry -> count -> selection -> values} load chains in flight instead of one
	__int128 acc = 0;
	const uint32_t lane = threadIdx.x & 31;
	const uint64_t n_warps = static_cast<uint64_t>(gridDim.x) * (kFilterThreads / 32);
	for (uint64_t chunk = static_cast<uint64_t>(blockIdx.x) * (kFilterThreads / 32) + (threadIdx.x >> 5); chunk < n_chunks;
	     chunk += n_warps) {
		const int64_t *a = reinterpret_cast<const int64_t *>(out + __ldg(chunk_col + 4 * chunk));
		const int64_t *b = reinterpret_cast<const int64_t *>(out + __ldg(chunk_col + 4 * chunk + 2));
		const uint32_t n = sel ? count[chunk] : __ldg(chunk_rows + chunk);
		const uint32_t *csel = sel ? sel + chunk * B2A_VECTOR_SIZE : nullptr;
		for (uint32_t k = lane; k < n; k += 32) {
			const uint32_t r = csel ? csel[k] : k;
			acc += static_cast<__int128>(__ldg(a + r)) * static_cast<__int128>(__ldg(b + r));
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
