// sm_100a decode kernels: Arrow IPC buffers (resident in HBM as a byte image of the stream) ->
// DuckDB vectors.  ONE persistent launch walks a tile table covering every (batch, field, buffer)
// task of the scan; grid = 148 SMs x resident CTAs.  All integer/byte work, HBM-bound; no tensor cores.
//
// Each op restates a loop of duckdb/src/function/table/arrow_conversion.cpp (line refs at each op).
#include "decode.cuh"

#include <cuda_runtime.h>

#include <cstdlib>

namespace b2a {

static constexpr int kThreads = 256;

// ------------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 ldg_stream16(const void *p) {
	uint4 r;
	asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
	             : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
	             : "l"(p));
	return r;
}
__device__ __forceinline__ uint2 ldg_stream8(const void *p) {
	uint2 r;
	asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
	return r;
}
__device__ __forceinline__ void stg_stream16(void *p, uint4 v) {
	asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
	             : "memory");
}
__device__ __forceinline__ uint32_t bswap32(uint32_t v) {
	return __byte_perm(v, 0, 0x0123);
}
__device__ __forceinline__ uint64_t bswap64(uint64_t v) {
	uint32_t lo = static_cast<uint32_t>(v), hi = static_cast<uint32_t>(v >> 32);
	return (static_cast<uint64_t>(bswap32(lo)) << 32) | bswap32(hi);
}
__device__ __forceinline__ void raise(uint32_t *err, uint32_t code) {
	atomicMax(err, code);
}

// byte-swap every `width`-byte element inside a 16-byte vector (K8, big-endian streams)
__device__ __forceinline__ uint4 bswap_elems(uint4 v, int width) {
	switch (width) {
	case 2:
		v.x = __byte_perm(v.x, 0, 0x2301);
		v.y = __byte_perm(v.y, 0, 0x2301);
		v.z = __byte_perm(v.z, 0, 0x2301);
		v.w = __byte_perm(v.w, 0, 0x2301);
		break;
	case 4:
		v.x = bswap32(v.x);
		v.y = bswap32(v.y);
		v.z = bswap32(v.z);
		v.w = bswap32(v.w);
		break;
	case 8: {
		uint32_t a = bswap32(v.y), b = bswap32(v.x), c = bswap32(v.w), d = bswap32(v.z);
		v = make_uint4(a, b, c, d);
		break;
	}
	case 16: {
		uint32_t a = bswap32(v.w), b = bswap32(v.z), c = bswap32(v.y), d = bswap32(v.x);
		v = make_uint4(a, b, c, d);
		break;
	}
	default:
		break;
	}
	return v;
}

// effective validity of row r = AND over the task's mask chain (own bitmap + struct / fixed-size-list
// ancestors).  Reference: SetValidityMask + the parent-mask loops, arrow_conversion.cpp:1143-1151,245-253,313-323
__device__ __forceinline__ bool eff_valid(const b2a_task &t, const uint8_t *__restrict__ image, uint64_t r) {
	bool ok = true;
	for (uint32_t m = 0; m < t.n_masks; m++) {
		uint64_t rr = t.masks[m].div == 1 ? r : r / t.masks[m].div;
		uint8_t b = __ldg(image + t.masks[m].src + (rr >> 3));
		ok = ok && ((b >> (rr & 7)) & 1);
	}
	return ok;
}

// ------------------------------------------------------------------------------------------------
// K1: fixed-width copy (DirectConversion, :454-461 — a pointer alias on the CPU, a real copy here)
// src is 8-byte aligned (IPC buffers), dst 16-byte aligned (our arena), tile starts keep both.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void op_copy(const b2a_task &t, const uint8_t *__restrict__ image, uint8_t *__restrict__ out,
                                        uint64_t row0, uint64_t nrows) {
	const uint64_t w = t.src_width;
	const uint8_t *src = image + t.src + row0 * w;
	uint8_t *dst = out + t.dst + row0 * w;
	const uint64_t bytes = nrows * w;
	const uint64_t nvec = bytes >> 4;
	const bool swap = (t.flags & B2A_TASK_F_BSWAP) && w > 1;
	const bool aligned16 = ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
	constexpr int U = 4;
	if (aligned16) {
		uint64_t i = threadIdx.x;
		for (; i + (U - 1) * kThreads < nvec; i += U * kThreads) {
			uint4 v[U];
#pragma unroll
			for (int u = 0; u < U; u++) {
				v[u] = ldg_stream16(src + ((i + u * kThreads) << 4));
			}
#pragma unroll
			for (int u = 0; u < U; u++) {
				if (swap) {
					v[u] = bswap_elems(v[u], static_cast<int>(w));
				}
				stg_stream16(dst + ((i + u * kThreads) << 4), v[u]);
			}
		}
		for (; i < nvec; i += kThreads) {
			uint4 v = ldg_stream16(src + (i << 4));
			if (swap) {
				v = bswap_elems(v, static_cast<int>(w));
			}
			stg_stream16(dst + (i << 4), v);
		}
	} else {
		uint64_t i = threadIdx.x;
		for (; i + (U - 1) * kThreads < nvec; i += U * kThreads) {
			uint2 a[U], b[U];
#pragma unroll
			for (int u = 0; u < U; u++) {
				const uint8_t *p = src + ((i + u * kThreads) << 4);
				a[u] = ldg_stream8(p);
				b[u] = ldg_stream8(p + 8);
			}
#pragma unroll
			for (int u = 0; u < U; u++) {
				uint4 v = make_uint4(a[u].x, a[u].y, b[u].x, b[u].y);
				if (swap) {
					v = bswap_elems(v, static_cast<int>(w));
				}
				stg_stream16(dst + ((i + u * kThreads) << 4), v);
			}
		}
		for (; i < nvec; i += kThreads) {
			const uint8_t *p = src + (i << 4);
			uint2 a = ldg_stream8(p), b = ldg_stream8(p + 8);
			uint4 v = make_uint4(a.x, a.y, b.x, b.y);
			if (swap) {
				v = bswap_elems(v, static_cast<int>(w));
			}
			stg_stream16(dst + (i << 4), v);
		}
	}
	// tail (< 16 bytes): whole elements, byte by byte
	const uint64_t tail0 = nvec << 4;
	for (uint64_t e = tail0 / w + threadIdx.x; e < nrows; e += kThreads) {
		for (uint64_t k = 0; k < w; k++) {
			dst[e * w + k] = src[e * w + (swap ? (w - 1 - k) : k)];
		}
	}
}

// ------------------------------------------------------------------------------------------------
// K2: validity (GetValidityMask, :63-108).  One thread per 64-row word of the DuckDB ValidityMask.
// Bytes past ceil(rows/8) stay 0xFF (fresh mask), exactly what the reference's memcpy leaves behind.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t mask_word(const b2a_task &t, const uint8_t *__restrict__ image, uint32_t m,
                                              uint64_t word, uint64_t nbytes) {
	const uint8_t *p = image + t.masks[m].src;
	const uint64_t div = t.masks[m].div;
	if (div == 1) {
		uint64_t b0 = word << 3;
		if (b0 + 8 <= nbytes) {
			uint2 v = *reinterpret_cast<const uint2 *>(p + b0); // bitmap buffers are 8-byte aligned
			return (static_cast<uint64_t>(v.y) << 32) | v.x;
		}
		uint64_t r = ~uint64_t(0);
		for (uint64_t k = 0; k < 8 && b0 + k < nbytes; k++) {
			r = (r & ~(uint64_t(0xFF) << (8 * k))) | (static_cast<uint64_t>(p[b0 + k]) << (8 * k));
		}
		return r;
	}
	// fixed-size-list broadcast: row r reads bit r/div of the ancestor
	uint64_t r = ~uint64_t(0);
	uint64_t row = word << 6;
	for (int b = 0; b < 64; b++, row++) {
		if (row >= t.rows) {
			break;
		}
		uint64_t rr = row / div;
		if (!((p[rr >> 3] >> (rr & 7)) & 1)) {
			r &= ~(uint64_t(1) << b);
		}
	}
	return r;
}

__device__ __forceinline__ void op_validity(const b2a_task &t, const uint8_t *__restrict__ image,
                                            uint8_t *__restrict__ out, uint64_t row0, uint64_t nrows) {
	const bool add_null = t.flags & B2A_TASK_F_ADD_NULL;
	const uint64_t total_rows = t.rows + (add_null ? 1 : 0);
	const uint64_t nbytes = (t.rows + 7) >> 3;
	const uint64_t w0 = row0 >> 6;
	// [row0, row0+nrows) is a slice of total_rows (the dispatcher counts the extra NULL slot as a row)
	(void)total_rows;
	const uint64_t w1 = (row0 + nrows + 63) >> 6;
	uint64_t *dst = reinterpret_cast<uint64_t *>(out + t.dst);
	unsigned long long valid = 0;
	for (uint64_t w = w0 + threadIdx.x; w < w1; w += kThreads) {
		uint64_t v = ~uint64_t(0);
		for (uint32_t m = 0; m < t.n_masks; m++) {
			v &= mask_word(t, image, m, w, nbytes);
		}
		if (add_null && (t.rows >> 6) == w) {
			v &= ~(uint64_t(1) << (t.rows & 63)); // mask.SetInvalid(size), :105-106
		}
		dst[w] = v;
		uint64_t lo = w << 6;
		uint64_t counted = v;
		if (lo + 64 > t.rows) {
			counted = (t.rows > lo) ? (v & ((uint64_t(1) << (t.rows - lo)) - 1)) : 0;
		}
		valid += __popcll(counted);
	}
	// popcount: warp reduce, one atomic per warp
	for (int o = 16; o > 0; o >>= 1) {
		valid += __shfl_xor_sync(0xffffffffu, valid, o);
	}
	if ((threadIdx.x & 31) == 0 && valid && t.aux != B2A_NONE_OFF) {
		atomicAdd(reinterpret_cast<unsigned long long *>(out + t.aux), valid);
	}
}

// ------------------------------------------------------------------------------------------------
// K3: bool bit-unpack (:869-891): 16 rows per thread = 2 source bytes -> one 16-byte store
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void op_bool(const b2a_task &t, const uint8_t *__restrict__ image, uint8_t *__restrict__ out,
                                        uint64_t row0, uint64_t nrows) {
	const uint8_t *src = image + t.src + (row0 >> 3);
	uint8_t *dst = out + t.dst + row0;
	const uint64_t ngroups = (nrows + 15) >> 4;
	for (uint64_t g = threadIdx.x; g < ngroups; g += kThreads) {
		uint64_t r = g << 4;
		uint32_t bits = src[g * 2];
		if (r + 8 < nrows) {
			bits |= static_cast<uint32_t>(src[g * 2 + 1]) << 8;
		}
		uint32_t wds[4];
#pragma unroll
		for (int k = 0; k < 4; k++) {
			uint32_t nib = (bits >> (4 * k)) & 0xF;
			// spread 4 bits into 4 bytes of 0/1
			wds[k] = (nib & 1) | ((nib & 2) << 7) | ((nib & 4) << 14) | ((nib & 8) << 21);
		}
		if (r + 16 <= nrows) {
			stg_stream16(dst + r, make_uint4(wds[0], wds[1], wds[2], wds[3]));
		} else {
			for (uint64_t k = 0; r + k < nrows; k++) {
				dst[r + k] = (bits >> k) & 1;
			}
		}
	}
}

// ------------------------------------------------------------------------------------------------
// K4: offsets + characters -> string_t (SetVectorString, :409-423; string_t ctor
// duckdb/src/include/duckdb/common/types/string_type.hpp:43-64).  One row per thread:
//   len <= 12 : {u32 len, 12 inline bytes zero padded}
//   len  > 12 : {u32 len, 4 prefix bytes, u64 pointer into the character buffer}
// Character bytes are fetched with aligned 32-bit loads + funnel shifts (the window of one warp is a
// contiguous span of the char buffer, so the loads coalesce in L1).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 make_string_t(const uint8_t *__restrict__ chars, uint64_t o0, uint32_t len,
                                                uint64_t ptr) {
	const uint8_t *base = chars + o0;
	const uint32_t *wp = reinterpret_cast<const uint32_t *>(reinterpret_cast<uintptr_t>(base) & ~uintptr_t(3));
	const uint32_t shift = (static_cast<uint32_t>(reinterpret_cast<uintptr_t>(base)) & 3u) * 8u;
	const uint32_t need = len > 12 ? 4u : len;                    // bytes of character data to fetch
	const uint32_t nwords = (need + (shift >> 3) + 3u) >> 2;       // aligned words that cover them
	uint32_t w0 = 0, w1 = 0, w2 = 0, w3 = 0;
	if (nwords > 0) {
		w0 = __ldg(wp);
	}
	if (nwords > 1) {
		w1 = __ldg(wp + 1);
	}
	if (nwords > 2) {
		w2 = __ldg(wp + 2);
	}
	if (nwords > 3) {
		w3 = __ldg(wp + 3);
	}
	uint32_t b0 = __funnelshift_r(w0, w1, shift);
	uint32_t b1 = __funnelshift_r(w1, w2, shift);
	uint32_t b2 = __funnelshift_r(w2, w3, shift);
	if (len > 12) {
		return make_uint4(len, b0, static_cast<uint32_t>(ptr), static_cast<uint32_t>(ptr >> 32));
	}
	// zero padding past `len`
	uint32_t m0 = len >= 4 ? 0xFFFFFFFFu : ((1u << (8 * len)) - 1u);
	uint32_t m1 = len >= 8 ? 0xFFFFFFFFu : (len > 4 ? ((1u << (8 * (len - 4))) - 1u) : 0u);
	uint32_t m2 = len >= 12 ? 0xFFFFFFFFu : (len > 8 ? ((1u << (8 * (len - 8))) - 1u) : 0u);
	return make_uint4(len, b0 & m0, b1 & m1, b2 & m2);
}

template <typename OFF>
__device__ __forceinline__ OFF load_off(const uint8_t *p, uint64_t i, bool swap);
template <>
__device__ __forceinline__ uint32_t load_off<uint32_t>(const uint8_t *p, uint64_t i, bool swap) {
	uint32_t v = __ldg(reinterpret_cast<const uint32_t *>(p) + i);
	return swap ? bswap32(v) : v;
}
template <>
__device__ __forceinline__ uint64_t load_off<uint64_t>(const uint8_t *p, uint64_t i, bool swap) {
	uint2 v = __ldg(reinterpret_cast<const uint2 *>(p) + i);
	uint64_t r = (static_cast<uint64_t>(v.y) << 32) | v.x;
	return swap ? bswap64(r) : r;
}

// Rows are processed kStrUnroll at a time per thread with the loads of all rows issued before the first use
// (offsets, then characters): the op is a dependent chain offsets -> characters -> store, so memory-level
// parallelism has to come from independent rows.
static constexpr int kStrUnroll = 4;

struct StrFetch {
	uint32_t w0, w1, w2, w3; // aligned words covering the bytes that end up in the string_t
	uint32_t shift;          // bit offset of the first character inside w0
};

__device__ __forceinline__ StrFetch fetch_chars(const uint8_t *__restrict__ base, uint32_t len) {
	StrFetch f;
	const uint32_t *wp = reinterpret_cast<const uint32_t *>(reinterpret_cast<uintptr_t>(base) & ~uintptr_t(3));
	f.shift = (static_cast<uint32_t>(reinterpret_cast<uintptr_t>(base)) & 3u) * 8u;
	const uint32_t need = len > 12 ? 4u : len;
	const uint32_t nwords = (need + (f.shift >> 3) + 3u) >> 2;
	f.w0 = f.w1 = f.w2 = f.w3 = 0;
	if (nwords > 0) {
		f.w0 = __ldg(wp);
	}
	if (nwords > 1) {
		f.w1 = __ldg(wp + 1);
	}
	if (nwords > 2) {
		f.w2 = __ldg(wp + 2);
	}
	if (nwords > 3) {
		f.w3 = __ldg(wp + 3);
	}
	return f;
}

__device__ __forceinline__ uint4 assemble_string_t(const StrFetch &f, uint32_t len, uint64_t ptr) {
	uint32_t b0 = __funnelshift_r(f.w0, f.w1, f.shift);
	if (len > 12) {
		return make_uint4(len, b0, static_cast<uint32_t>(ptr), static_cast<uint32_t>(ptr >> 32));
	}
	uint32_t b1 = __funnelshift_r(f.w1, f.w2, f.shift);
	uint32_t b2 = __funnelshift_r(f.w2, f.w3, f.shift);
	uint32_t m0 = len >= 4 ? 0xFFFFFFFFu : ((1u << (8 * len)) - 1u);
	uint32_t m1 = len >= 8 ? 0xFFFFFFFFu : (len > 4 ? ((1u << (8 * (len - 4))) - 1u) : 0u);
	uint32_t m2 = len >= 12 ? 0xFFFFFFFFu : (len > 8 ? ((1u << (8 * (len - 8))) - 1u) : 0u);
	return make_uint4(len, b0 & m0, b1 & m1, b2 & m2);
}

template <typename OFF>
__device__ __forceinline__ void op_string(const b2a_task &t, const uint8_t *__restrict__ image,
                                          uint8_t *__restrict__ out, uint64_t row0, uint64_t nrows,
                                          uint64_t ptr_origin, uint32_t *err) {
	const uint8_t *offs = image + t.src;
	const uint8_t *chars = image + t.src2;
	const bool swap = t.flags & B2A_TASK_F_BSWAP;
	const bool masked = t.n_masks != 0;
	const uint64_t src2_len = t.src2_len;
	const uint64_t ptr_base = ptr_origin + t.src2;
	uint4 *dst = reinterpret_cast<uint4 *>(out + t.dst);
	const uint64_t end = row0 + nrows;
	for (uint64_t base = row0 + threadIdx.x; base < end; base += kStrUnroll * kThreads) {
		OFF o0[kStrUnroll], o1[kStrUnroll];
#pragma unroll
		for (int u = 0; u < kStrUnroll; u++) {
			const uint64_t i = base + u * kThreads;
			o0[u] = o1[u] = 0;
			if (i < end) {
				o0[u] = load_off<OFF>(offs, i, swap);
				o1[u] = load_off<OFF>(offs, i + 1, swap);
			}
		}
		StrFetch f[kStrUnroll];
		uint32_t len[kStrUnroll];
		bool live[kStrUnroll];
#pragma unroll
		for (int u = 0; u < kStrUnroll; u++) {
			const uint64_t i = base + u * kThreads;
			live[u] = false;
			len[u] = 0;
			f[u].w0 = f[u].w1 = f[u].w2 = f[u].w3 = f[u].shift = 0;
			if (i < end) {
				if (o1[u] < o0[u] || static_cast<uint64_t>(o1[u]) > src2_len) {
					raise(err, B2A_DEV_BAD_OFFSETS);
				} else if (!masked || eff_valid(t, image, i)) {
					const uint64_t l = static_cast<uint64_t>(o1[u]) - static_cast<uint64_t>(o0[u]);
					if (l > 0xFFFFFFFFull) {
						raise(err, B2A_DEV_STRING_4GB);
					} else {
						live[u] = true;
						len[u] = static_cast<uint32_t>(l);
						f[u] = fetch_chars(chars + o0[u], len[u]);
					}
				}
			}
		}
#pragma unroll
		for (int u = 0; u < kStrUnroll; u++) {
			const uint64_t i = base + u * kThreads;
			if (i < end) {
				uint4 v = make_uint4(0, 0, 0, 0);
				if (live[u]) {
					v = assemble_string_t(f[u], len[u], ptr_base + o0[u]);
				}
				stg_stream16(dst + i, v);
			}
		}
	}
}

// K4v: Utf8View (SetVectorStringView, :425-452; arrow_string_view_type.hpp:24-85)
__device__ __forceinline__ void op_string_view(const b2a_task &t, const uint8_t *__restrict__ image,
                                               uint8_t *__restrict__ out, uint64_t row0, uint64_t nrows,
                                               uint64_t ptr_origin, const uint64_t *__restrict__ variadic,
                                               uint32_t *err) {
	const uint4 *views = reinterpret_cast<const uint4 *>(image + t.src); // 16-byte views, 8-byte aligned buffer
	uint4 *dst = reinterpret_cast<uint4 *>(out + t.dst);
	for (uint64_t i = row0 + threadIdx.x; i < row0 + nrows; i += kThreads) {
		const uint2 *vp = reinterpret_cast<const uint2 *>(views + i);
		uint2 a = __ldg(vp), b = __ldg(vp + 1);
		uint4 v = make_uint4(0, 0, 0, 0);
		if (t.n_masks == 0 || eff_valid(t, image, i)) {
			uint32_t len = a.x;
			if (len <= 12) {
				uint32_t m0 = len >= 4 ? 0xFFFFFFFFu : ((1u << (8 * len)) - 1u);
				uint32_t m1 = len >= 8 ? 0xFFFFFFFFu : (len > 4 ? ((1u << (8 * (len - 4))) - 1u) : 0u);
				uint32_t m2 = len >= 12 ? 0xFFFFFFFFu : (len > 8 ? ((1u << (8 * (len - 8))) - 1u) : 0u);
				v = make_uint4(len, a.y & m0, b.x & m1, b.y & m2);
			} else {
				uint32_t buffer_index = b.x;
				uint32_t offset = b.y;
				if (buffer_index >= static_cast<uint32_t>(t.param)) {
					raise(err, B2A_DEV_VIEW_RANGE);
				} else {
					const uint64_t *vb = variadic + t.variadic0 + 2 * static_cast<uint64_t>(buffer_index);
					if (static_cast<uint64_t>(offset) + len > vb[1]) { // {image offset, length} of the data buffer
						raise(err, B2A_DEV_VIEW_RANGE);
					} else {
						uint64_t data_off = vb[0] + offset;
						v = make_string_t(image, data_off, len, ptr_origin + data_off);
					}
				}
			}
		}
		stg_stream16(dst + i, v);
	}
}

// ------------------------------------------------------------------------------------------------
// K6: list / map offsets -> list_entry_t{offset,length} (ConvertArrowListOffsetsTemplated, :139-165).
// The reference rebases every 2 048-row chunk to its own first offset (cur_offset starts at 0 per call);
// aux[k] receives the child row where chunk k starts, aux[n_chunks] the end.
// ------------------------------------------------------------------------------------------------
// Row of THIS list column at which the reference's conversion call that covers row i started.  For a top-level
// list that is the 2 048-row chunk start.  A list nested under other lists / fixed-size lists is converted once
// per chunk of the TOP-LEVEL column, over the child range that chunk maps to, so the start is
// g(k) = offs_{d-1}[...offs_1[k*2048]] for the chunk k that contains i (ArrowToDuckDBList :234-278 recursion).
// rebase table (in `variadic`): [top_rows, n_levels, (kind, value) x n_levels], kind 0/1 = list32/64 (value = image
// offset of the ancestor's offsets), kind 2 = fixed-size list (value = list size).
__device__ __forceinline__ uint64_t rebase_eval(const uint64_t *__restrict__ tab, const uint8_t *__restrict__ image,
                                                uint64_t k, bool swap) {
	uint64_t r = k * B2A_VECTOR_SIZE;
	if (r > tab[0]) {
		r = tab[0];
	}
	const uint64_t n = tab[1];
	for (uint64_t l = 0; l < n; l++) {
		const uint64_t kind = tab[2 + 2 * l], val = tab[3 + 2 * l];
		if (kind == 2) {
			r *= val;
		} else if (kind == 0) {
			r = load_off<uint32_t>(image + val, r, swap);
		} else {
			r = load_off<uint64_t>(image + val, r, swap);
		}
	}
	return r;
}
__device__ __forceinline__ uint64_t chunk_base_row(const b2a_task &t, const uint8_t *__restrict__ image,
                                                   const uint64_t *__restrict__ variadic, uint64_t i, bool swap) {
	if (t.src2_len == 0) {
		return i & ~uint64_t(B2A_VECTOR_SIZE - 1);
	}
	const uint64_t *tab = variadic + t.variadic0;
	uint64_t lo = 0, hi = (tab[0] + B2A_VECTOR_SIZE - 1) / B2A_VECTOR_SIZE; // chunks of the top-level column
	while (hi - lo > 1) { // largest k with g(k) <= i
		uint64_t mid = (lo + hi) >> 1;
		if (rebase_eval(tab, image, mid, swap) <= i) {
			lo = mid;
		} else {
			hi = mid;
		}
	}
	return rebase_eval(tab, image, lo, swap);
}

template <typename OFF>
__device__ __forceinline__ void op_list(const b2a_task &t, const uint8_t *__restrict__ image, uint8_t *__restrict__ out,
                                        uint64_t row0, uint64_t nrows, const uint64_t *__restrict__ variadic,
                                        uint32_t *err) {
	const uint8_t *offs = image + t.src;
	const bool swap = t.flags & B2A_TASK_F_BSWAP;
	uint4 *dst = reinterpret_cast<uint4 *>(out + t.dst);
	uint64_t *aux = reinterpret_cast<uint64_t *>(out + t.aux);
	for (uint64_t i = row0 + threadIdx.x; i < row0 + nrows; i += kThreads) {
		OFF o0 = load_off<OFF>(offs, i, swap);
		OFF o1 = load_off<OFF>(offs, i + 1, swap);
		OFF cb = load_off<OFF>(offs, chunk_base_row(t, image, variadic, i, swap), swap);
		if (o1 < o0 || o0 < cb) {
			raise(err, B2A_DEV_BAD_OFFSETS);
		}
		uint64_t rel = static_cast<uint64_t>(o0) - static_cast<uint64_t>(cb);
		uint64_t len = static_cast<uint64_t>(static_cast<OFF>(o1 - o0));
		stg_stream16(dst + i, make_uint4(static_cast<uint32_t>(rel), static_cast<uint32_t>(rel >> 32),
		                                 static_cast<uint32_t>(len), static_cast<uint32_t>(len >> 32)));
		if ((i & (B2A_VECTOR_SIZE - 1)) == 0) {
			aux[i / B2A_VECTOR_SIZE] = o0;
		}
		if (i + 1 == t.rows) {
			aux[(t.rows + B2A_VECTOR_SIZE - 1) / B2A_VECTOR_SIZE] = o1;
			if (static_cast<uint64_t>(o1) > static_cast<uint64_t>(t.param)) { // offsets are monotone: the last one bounds all
				raise(err, B2A_DEV_LIST_CHILD);
			}
		}
	}
}

// fixed-size binary (ArrowToDuckDBBlob FIXED_SIZE, :347-360).  Bug-compatible with the reference: its running
// `offset` advances only on VALID rows (the `continue` at :353-355 skips `offset += blob_len`), so within one
// conversion call row i reads element  call_start + (#valid rows in [call_start, i))  of the data buffer.
__device__ __forceinline__ uint64_t count_valid(const b2a_task &t, const uint8_t *__restrict__ image, uint64_t from,
                                                uint64_t to) {
	uint64_t n = 0;
	const uint64_t nbytes = (t.rows + 7) >> 3;
	for (uint64_t w = from >> 6; (w << 6) < to; w++) {
		uint64_t v = ~uint64_t(0);
		for (uint32_t m = 0; m < t.n_masks; m++) {
			v &= mask_word(t, image, m, w, nbytes);
		}
		const uint64_t lo = w << 6;
		if (from > lo) {
			v &= ~uint64_t(0) << (from - lo);
		}
		if (to < lo + 64) {
			v &= (uint64_t(1) << (to - lo)) - 1;
		}
		n += __popcll(v);
	}
	return n;
}

__device__ __forceinline__ void op_blob_fixed(const b2a_task &t, const uint8_t *__restrict__ image,
                                              uint8_t *__restrict__ out, uint64_t row0, uint64_t nrows,
                                              uint64_t ptr_origin, const uint64_t *__restrict__ variadic) {
	const uint8_t *chars = image + t.src;
	const uint64_t w = static_cast<uint64_t>(t.param);
	uint4 *dst = reinterpret_cast<uint4 *>(out + t.dst);
	for (uint64_t i = row0 + threadIdx.x; i < row0 + nrows; i += kThreads) {
		uint4 v = make_uint4(0, 0, 0, 0);
		if (t.n_masks == 0) {
			v = make_string_t(chars, i * w, static_cast<uint32_t>(w), ptr_origin + t.src + i * w);
		} else if (eff_valid(t, image, i)) {
			const uint64_t call_start = chunk_base_row(t, image, variadic, i, false);
			const uint64_t e = call_start + count_valid(t, image, call_start, i);
			v = make_string_t(chars, e * w, static_cast<uint32_t>(w), ptr_origin + t.src + e * w);
		}
		stg_stream16(dst + i, v);
	}
}

// ------------------------------------------------------------------------------------------------
// K7: scalar transforms.  One row per thread; src/dst are naturally aligned for their widths.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int64_t load_i64(const uint8_t *p, uint64_t i, bool swap) {
	return static_cast<int64_t>(load_off<uint64_t>(p, i, swap));
}
__device__ __forceinline__ int32_t load_i32(const uint8_t *p, uint64_t i, bool swap) {
	return static_cast<int32_t>(load_off<uint32_t>(p, i, swap));
}
// TryMultiplyOperator::Operation(int64_t,int64_t,int64_t&) == __builtin_mul_overflow
// (duckdb/src/function/scalar/operator/multiply.cpp:190-194)
__device__ __forceinline__ bool try_mul_i64(int64_t a, int64_t b, int64_t &r) {
	r = static_cast<int64_t>(static_cast<uint64_t>(a) * static_cast<uint64_t>(b));
	int64_t hi = __mul64hi(a, b);
	return hi == (r >> 63);
}
__device__ __forceinline__ void store_interval(uint8_t *dst, uint64_t i, int32_t months, int32_t days, int64_t micros) {
	stg_stream16(reinterpret_cast<uint4 *>(dst) + i,
	             make_uint4(static_cast<uint32_t>(months), static_cast<uint32_t>(days), static_cast<uint32_t>(micros),
	                        static_cast<uint32_t>(static_cast<uint64_t>(micros) >> 32)));
}

struct i128 {
	uint64_t lo;
	int64_t hi;
};
__device__ __forceinline__ i128 load_i128(const uint8_t *p, uint64_t i, bool swap) {
	const uint2 *q = reinterpret_cast<const uint2 *>(p) + 2 * i;
	uint2 a = __ldg(q), b = __ldg(q + 1);
	uint64_t w0 = (static_cast<uint64_t>(a.y) << 32) | a.x, w1 = (static_cast<uint64_t>(b.y) << 32) | b.x;
	i128 r;
	if (swap) {
		r.lo = bswap64(w1);
		r.hi = static_cast<int64_t>(bswap64(w0));
	} else {
		r.lo = w0;
		r.hi = static_cast<int64_t>(w1);
	}
	return r;
}

// SetSelectionVector* fast path (no NULL indices, little-endian, index width 1/2/4): 8 rows per thread per trip —
// 8-byte loads (IPC buffers are 8-byte aligned and tiles start on a multiple of 2 048 rows), two 16-byte sel_t stores.
template <int W, bool SIGNED>
__device__ __forceinline__ void dict_index_widen(const uint2 (&x)[W == 4 ? 4 : (W == 2 ? 2 : 1)], uint4 *d,
                                                 uint32_t limit, uint32_t *err) {
	uint32_t v[8];
	if (W == 1) {
#pragma unroll
		for (int k = 0; k < 4; k++) {
			const uint32_t a = (x[0].x >> (8 * k)) & 0xFFu, b = (x[0].y >> (8 * k)) & 0xFFu;
			v[k] = SIGNED ? static_cast<uint32_t>(static_cast<int32_t>(static_cast<int8_t>(a))) : a;
			v[4 + k] = SIGNED ? static_cast<uint32_t>(static_cast<int32_t>(static_cast<int8_t>(b))) : b;
		}
	} else if (W == 2) {
		const uint32_t w[4] = {x[0].x, x[0].y, x[W == 2 ? 1 : 0].x, x[W == 2 ? 1 : 0].y};
#pragma unroll
		for (int k = 0; k < 4; k++) {
			const uint32_t a = w[k] & 0xFFFFu, b = w[k] >> 16;
			v[2 * k] = SIGNED ? static_cast<uint32_t>(static_cast<int32_t>(static_cast<int16_t>(a))) : a;
			v[2 * k + 1] = SIGNED ? static_cast<uint32_t>(static_cast<int32_t>(static_cast<int16_t>(b))) : b;
		}
	} else {
#pragma unroll
		for (int k = 0; k < 4; k++) {
			v[2 * k] = x[W == 4 ? k : 0].x;
			v[2 * k + 1] = x[W == 4 ? k : 0].y;
		}
	}
	const uint32_t m = max(max(max(v[0], v[1]), max(v[2], v[3])), max(max(v[4], v[5]), max(v[6], v[7])));
	if (m >= limit) { // sign-extended negatives are >= limit as well
		raise(err, B2A_DEV_DICT_INDEX);
	}
	stg_stream16(d, make_uint4(v[0], v[1], v[2], v[3]));
	stg_stream16(d + 1, make_uint4(v[4], v[5], v[6], v[7]));
}

template <int W, bool SIGNED>
__device__ __forceinline__ void dict_index_fast(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, uint64_t row0,
                                                uint64_t nrows, uint32_t limit, uint32_t *err) {
	constexpr int L = W == 4 ? 4 : (W == 2 ? 2 : 1); // 8-byte loads per group of 8 rows
	constexpr int U = W == 4 ? 2 : 4;                // groups in flight per thread
	const uint64_t groups = nrows >> 3;
	const uint2 *s = reinterpret_cast<const uint2 *>(src + row0 * W);
	uint4 *d = reinterpret_cast<uint4 *>(dst + row0 * 4);
	uint64_t g = threadIdx.x;
	for (; g + (U - 1) * kThreads < groups; g += U * kThreads) {
		uint2 x[U][L];
#pragma unroll
		for (int u = 0; u < U; u++) {
#pragma unroll
			for (int k = 0; k < L; k++) {
				x[u][k] = ldg_stream8(s + (g + u * kThreads) * L + k);
			}
		}
#pragma unroll
		for (int u = 0; u < U; u++) {
			dict_index_widen<W, SIGNED>(x[u], d + 2 * (g + u * kThreads), limit, err);
		}
	}
	for (; g < groups; g += kThreads) {
		uint2 x[L];
#pragma unroll
		for (int k = 0; k < L; k++) {
			x[k] = ldg_stream8(s + g * L + k);
		}
		dict_index_widen<W, SIGNED>(x, d + 2 * g, limit, err);
	}
}

__device__ __forceinline__ void op_scalar(const b2a_task &t, const uint8_t *__restrict__ image, uint8_t *__restrict__ out,
                                          uint64_t row0, uint64_t nrows, uint32_t *err) {
	const uint8_t *src = image + t.src;
	uint8_t *dst = out + t.dst;
	const bool swap = t.flags & B2A_TASK_F_BSWAP;
	if (t.op == B2A_OP_DICT_INDEX && t.n_masks == 0 && !swap && t.src_width <= 4) {
		const bool sgn = t.flags & B2A_TASK_F_SIGNED;
		const uint32_t limit = t.param > 0xFFFFFFFFll ? 0xFFFFFFFFu : static_cast<uint32_t>(t.param);
		switch (t.src_width) {
		case 1:
			sgn ? dict_index_fast<1, true>(src, dst, row0, nrows, limit, err)
			    : dict_index_fast<1, false>(src, dst, row0, nrows, limit, err);
			break;
		case 2:
			sgn ? dict_index_fast<2, true>(src, dst, row0, nrows, limit, err)
			    : dict_index_fast<2, false>(src, dst, row0, nrows, limit, err);
			break;
		default:
			dict_index_fast<4, false>(src, dst, row0, nrows, limit, err);
			break;
		}
		const uint64_t done = nrows & ~7ull; // the last (< 8) rows of the batch take the row-wise path below
		row0 += done;
		nrows -= done;
	}
	for (uint64_t i = row0 + threadIdx.x; i < row0 + nrows; i += kThreads) {
		switch (t.op) {
		case B2A_OP_MUL_I32:   // TimeConversion<int32_t>, :463-478 (valid rows only)
		case B2A_OP_MUL_I64: { // TimeConversion<int64_t> / TimestampTZConversion, :497-511
			int64_t r = 0;
			if (t.n_masks == 0 || eff_valid(t, image, i)) {
				int64_t v = t.op == B2A_OP_MUL_I32 ? static_cast<int64_t>(load_i32(src, i, swap)) : load_i64(src, i, swap);
				if (!try_mul_i64(v, t.param, r)) {
					raise(err, B2A_DEV_MUL_OVERFLOW);
				}
			}
			reinterpret_cast<int64_t *>(dst)[i] = r;
			break;
		}
		case B2A_OP_DIV_I64: // :986-994, :1018-1026 — every row, truncating division
			reinterpret_cast<int64_t *>(dst)[i] = load_i64(src, i, swap) / t.param;
			break;
		case B2A_OP_DATE64: // :951-961
			reinterpret_cast<int32_t *>(dst)[i] = static_cast<int32_t>(load_i64(src, i, swap) / 86400000ll);
			break;
		case B2A_OP_INTERVAL_MUL: { // IntervalConversionUs, :513-525 — every row, NULL slots included
			int64_t r;
			if (!try_mul_i64(load_i64(src, i, swap), t.param, r)) {
				raise(err, B2A_DEV_MUL_OVERFLOW);
			}
			store_interval(dst, i, 0, 0, r);
			break;
		}
		case B2A_OP_INTERVAL_DIV: // :1052-1062
			store_interval(dst, i, 0, 0, load_i64(src, i, swap) / t.param);
			break;
		case B2A_OP_INTERVAL_MONTHS: // :527-537
			store_interval(dst, i, load_i32(src, i, swap), 0, 0);
			break;
		case B2A_OP_INTERVAL_MDN: { // :539-549, ArrowInterval{i32 months,i32 days,i64 nanos}
			int32_t months = load_i32(src, 4 * i, swap), days = load_i32(src, 4 * i + 1, swap);
			int64_t nanos = load_i64(src, 2 * i + 1, swap);
			store_interval(dst, i, months, days, nanos / 1000);
			break;
		}
		case B2A_OP_UUID: { // UUIDConversion, :480-495 (valid rows only)
			uint4 v = make_uint4(0, 0, 0, 0);
			if (t.n_masks == 0 || eff_valid(t, image, i)) {
				i128 s = load_i128(src, i, false);
				uint64_t lower = bswap64(static_cast<uint64_t>(s.hi));
				uint64_t upper = bswap64(s.lo) ^ (uint64_t(1) << 63);
				v = make_uint4(static_cast<uint32_t>(lower), static_cast<uint32_t>(lower >> 32),
				               static_cast<uint32_t>(upper), static_cast<uint32_t>(upper >> 32));
			}
			stg_stream16(reinterpret_cast<uint4 *>(dst) + i, v);
			break;
		}
		case B2A_OP_DECIMAL: { // ConvertDecimal, :768-844: TryCast src -> physical type, valid rows only
			i128 v = {0, 0};
			bool valid = t.n_masks == 0 || eff_valid(t, image, i);
			if (valid) {
				if (t.src_width == 4) {
					int32_t s = load_i32(src, i, swap);
					v.lo = static_cast<uint64_t>(static_cast<int64_t>(s));
					v.hi = s < 0 ? -1 : 0;
				} else if (t.src_width == 8) {
					int64_t s = load_i64(src, i, swap);
					v.lo = static_cast<uint64_t>(s);
					v.hi = s < 0 ? -1 : 0;
				} else {
					v = load_i128(src, i, swap);
				}
			}
			// range check of the narrowing cast (HugeintTryCastInteger, hugeint.cpp:559-582);
			// a failed cast leaves the slot unwritten in the reference (release build) — we store 0
			int64_t s64 = static_cast<int64_t>(v.lo);
			bool fits64 = (v.hi == 0 && s64 >= 0) || (v.hi == -1 && s64 < 0);
			switch (t.dst_width) {
			case 2:
				reinterpret_cast<int16_t *>(dst)[i] =
				    (valid && fits64 && s64 >= -32768 && s64 <= 32767) ? static_cast<int16_t>(s64) : 0;
				break;
			case 4:
				reinterpret_cast<int32_t *>(dst)[i] =
				    (valid && fits64 && s64 >= -2147483648ll && s64 <= 2147483647ll) ? static_cast<int32_t>(s64) : 0;
				break;
			case 8:
				reinterpret_cast<int64_t *>(dst)[i] = (valid && fits64) ? s64 : 0;
				break;
			default:
				stg_stream16(reinterpret_cast<uint4 *>(dst) + i,
				             make_uint4(static_cast<uint32_t>(v.lo), static_cast<uint32_t>(v.lo >> 32),
				                        static_cast<uint32_t>(static_cast<uint64_t>(v.hi)),
				                        static_cast<uint32_t>(static_cast<uint64_t>(v.hi) >> 32)));
				break;
			}
			break;
		}
		case B2A_OP_DICT_INDEX: { // SetSelectionVector*, :1227-1348: widen to sel_t, NULL -> dictionary length
			uint64_t idx;
			const bool sgn = t.flags & B2A_TASK_F_SIGNED;
			switch (t.src_width) {
			case 1:
				idx = sgn ? static_cast<uint64_t>(static_cast<int64_t>(static_cast<int8_t>(src[i]))) : src[i];
				break;
			case 2: {
				uint16_t v = reinterpret_cast<const uint16_t *>(src)[i];
				if (swap) {
					v = static_cast<uint16_t>((v >> 8) | (v << 8));
				}
				idx = sgn ? static_cast<uint64_t>(static_cast<int64_t>(static_cast<int16_t>(v))) : v;
				break;
			}
			case 4: {
				uint32_t v = load_off<uint32_t>(src, i, swap);
				idx = sgn ? static_cast<uint64_t>(static_cast<int64_t>(static_cast<int32_t>(v))) : v;
				break;
			}
			default:
				idx = load_off<uint64_t>(src, i, swap);
				break;
			}
			if (t.n_masks != 0 && !eff_valid(t, image, i)) {
				idx = static_cast<uint64_t>(t.param);
			} else if (t.src_width == 8 && static_cast<uint64_t>(t.param) > 0xFFFFFFFFull && idx > 0xFFFFFFFFull) {
				raise(err, B2A_DEV_INDEX_U32);
			} else if (idx >= static_cast<uint64_t>(t.param)) { // negative indices land here too (sign-extended)
				raise(err, B2A_DEV_DICT_INDEX);
			}
			reinterpret_cast<uint32_t *>(dst)[i] = static_cast<uint32_t>(idx);
			break;
		}
		default:
			break;
		}
	}
}

// ------------------------------------------------------------------------------------------------
// Run-end encoded: FindRunIndex + FlattenRunEnds (arrow_conversion.cpp:551-663).  One row per thread: binary search of
// the first run end that is greater than the row, then the DECODED value of that run (already in the arena: this op
// runs in the second launch) is copied; when the values can be NULL the result mask is written through a warp ballot.
// Rows past the last run end are left untouched, like the reference's loop leaves them.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int64_t ree_run_end(const uint8_t *re, uint32_t width, uint64_t i, bool swap) {
	if (width == 2) {
		uint16_t v = reinterpret_cast<const uint16_t *>(re)[i];
		if (swap) {
			v = static_cast<uint16_t>((v >> 8) | (v << 8));
		}
		return static_cast<int16_t>(v);
	}
	uint32_t v = reinterpret_cast<const uint32_t *>(re)[i];
	return static_cast<int32_t>(swap ? bswap32(v) : v);
}

__device__ __forceinline__ void op_ree(const b2a_task &t, const uint8_t *__restrict__ image, uint8_t *__restrict__ out,
                                       uint64_t row0, uint64_t nrows) {
	const uint8_t *re = image + t.src;
	const uint8_t *vals = out + t.src2;
	const uint64_t *vmask = t.src2_len == B2A_NONE_OFF ? nullptr : reinterpret_cast<const uint64_t *>(out + t.src2_len);
	uint8_t *dst = out + t.dst;
	uint32_t *omask = t.aux == B2A_NONE_OFF ? nullptr : reinterpret_cast<uint32_t *>(out + t.aux);
	const uint64_t R = static_cast<uint64_t>(t.param);
	const bool swap = t.flags & B2A_TASK_F_BSWAP;
	const uint32_t w = t.dst_width;
	const uint64_t end = row0 + nrows;
	// whole warps walk the tile so that the ballot always covers 32 consecutive rows
	for (uint64_t base = row0 + (threadIdx.x & ~31u); base < end; base += kThreads) {
		const uint64_t i = base + (threadIdx.x & 31u);
		bool valid = true;
		if (i < end) {
			uint64_t lo = 0, hi = R;
			while (lo < hi) {
				const uint64_t mid = (lo + hi) >> 1;
				if (static_cast<int64_t>(i) >= ree_run_end(re, t.src_width, mid, swap)) {
					lo = mid + 1;
				} else {
					hi = mid;
				}
			}
			if (lo < R) {
				valid = !vmask || ((vmask[lo >> 6] >> (lo & 63)) & 1ull);
				if (valid) {
					const uint8_t *s = vals + lo * w;
					uint8_t *d = dst + i * w;
					switch (w) {
					case 1:
						*d = *s;
						break;
					case 2:
						*reinterpret_cast<uint16_t *>(d) = *reinterpret_cast<const uint16_t *>(s);
						break;
					case 4:
						*reinterpret_cast<uint32_t *>(d) = *reinterpret_cast<const uint32_t *>(s);
						break;
					case 8:
						*reinterpret_cast<uint64_t *>(d) = *reinterpret_cast<const uint64_t *>(s);
						break;
					default:
						*reinterpret_cast<uint4 *>(d) = *reinterpret_cast<const uint4 *>(s);
						break;
					}
				}
			}
		}
		if (omask) {
			const uint32_t bits = __ballot_sync(0xFFFFFFFFu, valid);
			if ((threadIdx.x & 31u) == 0) {
				omask[base >> 5] = bits;
			}
		}
	}
}

// Sparse UNION (arrow_conversion.cpp:1173-1221).  Per row: tag = type_ids[row_idx] with row_idx the row WITHIN THE
// 2 048-ROW CHUNK — the reference indexes the type id buffer without the chunk offset (:1209) while the members are
// converted at it, so past the first chunk of a batch the tags repeat the batch's first 2 048; kept as is.  The member
// the tag names is boxed (GetValue) and stored back with SetValue: union row valid iff that member row is, tag child =
// tag, member m valid where tag == m and its own value is.  variadic[v0 .. v0+n) = arena offsets of the members'
// decoded masks (none = all valid), variadic[v0+n .. v0+2n) = arena offsets of the masks written here.
__device__ __forceinline__ void op_union(const b2a_task &t, const uint8_t *__restrict__ image, uint8_t *__restrict__ out,
                                         uint64_t row0, uint64_t nrows, const uint64_t *__restrict__ variadic,
                                         uint32_t *err) {
	const int8_t *type_ids = reinterpret_cast<const int8_t *>(image + t.src);
	uint8_t *tags = out + t.dst;
	uint32_t *umask = reinterpret_cast<uint32_t *>(out + t.aux);
	const uint32_t n = static_cast<uint32_t>(t.param);
	const uint64_t *offs = variadic + t.variadic0;
	const uint64_t end = row0 + nrows;
	const uint32_t lane = threadIdx.x & 31u;
	for (uint64_t base = row0 + (threadIdx.x & ~31u); base < end; base += kThreads) {
		const uint64_t i = base + lane;
		uint32_t tag = 0xFFFFFFFFu;
		if (i < end) {
			const int8_t raw_tag = type_ids[i % B2A_VECTOR_SIZE];
			tag = static_cast<uint8_t>(raw_tag);
			if (raw_tag < 0 || tag >= n) {
				raise(err, B2A_DEV_UNION_TAG | (tag << 8));
				tag = 0xFFFFFFFFu;
			} else {
				tags[i] = static_cast<uint8_t>(tag);
			}
		}
		bool valid = false;
		for (uint32_t m = 0; m < n; m++) {
			const uint64_t raw_off = __ldg(offs + m);
			const uint32_t raw = raw_off == B2A_NONE_OFF ? 0xFFFFFFFFu
			                                             : reinterpret_cast<const uint32_t *>(out + raw_off)[base >> 5];
			const uint32_t eq = __ballot_sync(0xFFFFFFFFu, tag == m);
			if (tag == m) {
				valid = (raw >> lane) & 1u;
			}
			if (lane == 0) {
				reinterpret_cast<uint32_t *>(out + __ldg(offs + n + m))[base >> 5] = eq & raw;
			}
		}
		const uint32_t bits = __ballot_sync(0xFFFFFFFFu, valid);
		if (lane == 0) {
			umask[base >> 5] = bits;
		}
	}
}

// List views (ConvertArrowListViewOffsetsTemplated, arrow_conversion.cpp:166-203).  Per 2 048-row chunk (= one conversion
// call of the reference): lowest = min(offsets[0], offsets[i] of the non-empty views), list_size = sum of the sizes;
// entry i = {offsets[i] <= lowest ? 0 : offsets[i] - lowest, sizes[i]}; aux[chunk] = {lowest, list_size} is the window of
// the child column this chunk's child vector covers.  One CTA per chunk: 8 rows per thread in registers, block-wide
// min / sum.  offsets[i] + sizes[i] beyond the child array fails like nanoarrow's FULL validation.
template <typename T>
__device__ __forceinline__ void op_listview(const b2a_task &t, const uint8_t *__restrict__ image, uint8_t *__restrict__ out,
                                            uint64_t row0, uint64_t nrows, uint32_t *err) {
	constexpr uint32_t kPer = B2A_VECTOR_SIZE / kThreads;
	__shared__ uint64_t s_min[kThreads / 32], s_sum[kThreads / 32];
	const T *offsets = reinterpret_cast<const T *>(image + t.src);
	const T *sizes = reinterpret_cast<const T *>(image + t.src2);
	ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(out + t.dst);
	ulonglong2 *aux = reinterpret_cast<ulonglong2 *>(out + t.aux);
	const bool swap = t.flags & B2A_TASK_F_BSWAP;
	const uint64_t child_len = static_cast<uint64_t>(t.param);
	for (uint64_t c0 = row0; c0 < row0 + nrows; c0 += B2A_VECTOR_SIZE) {
		const uint64_t cn = min(static_cast<uint64_t>(B2A_VECTOR_SIZE), row0 + nrows - c0);
		uint64_t o[kPer], z[kPer];
		T first = offsets[c0];
		if (swap) {
			first = sizeof(T) == 4 ? static_cast<T>(bswap32(static_cast<uint32_t>(first))) : static_cast<T>(bswap64(first));
		}
		uint64_t lo = first, sum = 0;
#pragma unroll
		for (uint32_t k = 0; k < kPer; k++) {
			const uint64_t r = threadIdx.x + k * kThreads;
			o[k] = 0;
			z[k] = 0;
			if (r < cn) {
				T ov = offsets[c0 + r], zv = sizes[c0 + r];
				if (swap) {
					ov = sizeof(T) == 4 ? static_cast<T>(bswap32(static_cast<uint32_t>(ov))) : static_cast<T>(bswap64(ov));
					zv = sizeof(T) == 4 ? static_cast<T>(bswap32(static_cast<uint32_t>(zv))) : static_cast<T>(bswap64(zv));
				}
				o[k] = ov;
				z[k] = zv;
				if (o[k] > child_len || z[k] > child_len - o[k]) {
					raise(err, B2A_DEV_BAD_OFFSETS);
				}
				if (z[k] != 0 && o[k] < lo) {
					lo = o[k];
				}
				sum += z[k];
			}
		}
#pragma unroll
		for (int d = 16; d > 0; d >>= 1) {
			const uint64_t ol = __shfl_xor_sync(0xFFFFFFFFu, lo, d), os = __shfl_xor_sync(0xFFFFFFFFu, sum, d);
			lo = ol < lo ? ol : lo;
			sum += os;
		}
		__syncthreads(); // the previous chunk's readers are done with s_min / s_sum
		if ((threadIdx.x & 31u) == 0) {
			s_min[threadIdx.x >> 5] = lo;
			s_sum[threadIdx.x >> 5] = sum;
		}
		__syncthreads();
		lo = s_min[0];
		sum = s_sum[0];
#pragma unroll
		for (uint32_t w = 1; w < kThreads / 32; w++) {
			lo = s_min[w] < lo ? s_min[w] : lo;
			sum += s_sum[w];
		}
#pragma unroll
		for (uint32_t k = 0; k < kPer; k++) {
			const uint64_t r = threadIdx.x + k * kThreads;
			if (r < cn) {
				dst[c0 + r] = make_ulonglong2(o[k] <= lo ? 0 : o[k] - lo, z[k]);
			}
		}
		if (threadIdx.x == 0) {
			aux[c0 / B2A_VECTOR_SIZE] = make_ulonglong2(lo, sum);
		}
	}
}

// arrow.bool8 (ArrowBool8::ArrowToDuck, arrow_type_extension.cpp:344-350): result[i] = (bool)source[i], i.e. 1 for
// every non-zero int8.  Eight rows per 64-bit word when both sides are 8-byte aligned.
__device__ __forceinline__ void op_bool8(const b2a_task &t, const uint8_t *__restrict__ image, uint8_t *__restrict__ out,
                                         uint64_t row0, uint64_t nrows) {
	const uint8_t *src = image + t.src + row0;
	uint8_t *dst = out + t.dst + row0;
	uint64_t done = 0;
	if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 7u) == 0) {
		const uint64_t words = nrows / 8;
		for (uint64_t w = threadIdx.x; w < words; w += kThreads) {
			const uint64_t v = __ldg(reinterpret_cast<const uint64_t *>(src) + w);
			const uint64_t nz = (((v & 0x7F7F7F7F7F7F7F7Full) + 0x7F7F7F7F7F7F7F7Full) | v) >> 7;
			reinterpret_cast<uint64_t *>(dst)[w] = nz & 0x0101010101010101ull;
		}
		done = words * 8;
	}
	for (uint64_t i = done + threadIdx.x; i < nrows; i += kThreads) {
		dst[i] = src[i] != 0;
	}
}

// ------------------------------------------------------------------------------------------------
// the one kernel: persistent CTAs stride over the tile table
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
	const uint32_t sa = static_cast<uint32_t>(__cvta_generic_to_shared(smem));
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
	asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

static constexpr uint32_t kTaskVecs = sizeof(b2a_task) / 16;
static_assert(sizeof(b2a_task) % 16 == 0, "b2a_task is copied in 16-byte pieces");

// Persistent CTAs stride over the tile table.  The descriptor of the NEXT tile is fetched with cp.async while the
// current one is processed (its task index was loaded one iteration earlier), so no tile starts with an exposed
// tile_task -> task dependent load pair.
// kFull = false is the lean instantiation for scans whose tasks are all COPY / VALIDITY / BOOL / STRING32 / STRING64
// (TPC-H lineitem, most flat tables): without the rarely used ops in its switch the register allocation of the hot
// paths is not disturbed by them (adding ONE op to the switch cost 3.5 %, see b2a_ree_kernel).
template <bool kFull>
__global__ void __launch_bounds__(kThreads)
    b2a_decode_kernel(const b2a_task *__restrict__ tasks, const uint32_t *__restrict__ tile_task, uint32_t tile_begin,
                      uint32_t tile_end,
                      const uint8_t *__restrict__ image, uint8_t *__restrict__ out, uint64_t ptr_origin,
                      uint32_t *err, const uint64_t *__restrict__ variadic) {
	__shared__ b2a_task s_task[2];
	uint32_t tile = tile_begin + blockIdx.x;
	if (tile >= tile_end) {
		return;
	}
	const uint32_t stride = gridDim.x;
	const bool loader = threadIdx.x < kTaskVecs;
	uint32_t ti_next = 0;
	if (loader) {
		const uint32_t ti = __ldg(tile_task + tile);
		reinterpret_cast<uint4 *>(&s_task[0])[threadIdx.x] = __ldg(reinterpret_cast<const uint4 *>(tasks + ti) + threadIdx.x);
		if (tile + stride < tile_end) {
			ti_next = __ldg(tile_task + tile + stride);
		}
	}
	__syncthreads();
	for (uint32_t buf = 0;; buf ^= 1) {
		const uint32_t next = tile + stride;
		if (loader && next < tile_end) {
			cp_async16(reinterpret_cast<uint4 *>(&s_task[buf ^ 1]) + threadIdx.x,
			           reinterpret_cast<const uint4 *>(tasks + ti_next) + threadIdx.x);
			if (next + stride < tile_end) {
				ti_next = __ldg(tile_task + next + stride);
			}
		}
		const b2a_task &t = s_task[buf];
		const uint64_t row0 = static_cast<uint64_t>(tile - t.tile0) * t.tile_rows;
		uint64_t limit = t.rows;
		if (t.op == B2A_OP_VALIDITY && (t.flags & B2A_TASK_F_ADD_NULL)) {
			limit += 1;
		}
		uint64_t nrows = limit - row0;
		if (nrows > t.tile_rows) {
			nrows = t.tile_rows;
		}
		switch (t.op) {
		case B2A_OP_COPY:
			op_copy(t, image, out, row0, nrows);
			break;
		case B2A_OP_VALIDITY:
			op_validity(t, image, out, row0, nrows);
			break;
		case B2A_OP_BOOL:
			op_bool(t, image, out, row0, nrows);
			break;
		case B2A_OP_STRING32:
			op_string<uint32_t>(t, image, out, row0, nrows, ptr_origin, err);
			break;
		case B2A_OP_STRING64:
			op_string<uint64_t>(t, image, out, row0, nrows, ptr_origin, err);
			break;
		case B2A_OP_STRING_VIEW:
			if (kFull) {
				op_string_view(t, image, out, row0, nrows, ptr_origin, variadic, err);
			}
			break;
		case B2A_OP_BLOB_FIXED:
			if (kFull) {
				op_blob_fixed(t, image, out, row0, nrows, ptr_origin, variadic);
			}
			break;
		case B2A_OP_LIST32:
			if (kFull) {
				op_list<uint32_t>(t, image, out, row0, nrows, variadic, err);
			}
			break;
		case B2A_OP_LIST64:
			if (kFull) {
				op_list<uint64_t>(t, image, out, row0, nrows, variadic, err);
			}
			break;
		default:
			if (kFull) {
				op_scalar(t, image, out, row0, nrows, err);
			}
			break;
		}
		if (next >= tile_end) {
			break;
		}
		cp_async_wait_all();
		__syncthreads(); // next descriptor has landed; everyone is done with the current one
		tile = next;
	}
}

int LaunchDecode(const DecodeLaunch &l, void *stream) {
	if (l.n_tiles == 0) {
		return 0;
	}
	// persistent grid: SM count x resident CTAs per SM (queried once per instantiation)
	static int ctas_per_sm[2] = {0, 0};
	const int v = l.lean ? 0 : 1;
	if (ctas_per_sm[v] == 0) {
		cudaError_t qe = l.lean ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm[v], b2a_decode_kernel<false>,
		                                                                        kThreads, 0)
		                        : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm[v], b2a_decode_kernel<true>,
		                                                                        kThreads, 0);
		if (qe != cudaSuccess || ctas_per_sm[v] <= 0) {
			ctas_per_sm[v] = 4;
		}
	}
	static int ctas_override = -1;
	if (ctas_override < 0) {
		const char *e = getenv("B2A_CTAS_PER_SM"); // tuning knob for experiments
		ctas_override = e ? atoi(e) : 0;
	}
	uint32_t grid = static_cast<uint32_t>(l.sm_count) *
	                static_cast<uint32_t>(ctas_override > 0 ? ctas_override : ctas_per_sm[v]);
	if (grid > l.n_tiles) {
		grid = l.n_tiles;
	}
	if (l.lean) {
		b2a_decode_kernel<false><<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
		    l.d_tasks, l.d_tile_task, l.tile_begin, l.tile_begin + l.n_tiles, l.image_origin, l.out, l.string_ptr_origin,
		    l.d_err, l.d_variadic);
	} else {
		b2a_decode_kernel<true><<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
		    l.d_tasks, l.d_tile_task, l.tile_begin, l.tile_begin + l.n_tiles, l.image_origin, l.out, l.string_ptr_origin,
		    l.d_err, l.d_variadic);
	}
	return static_cast<int>(cudaGetLastError());
}

// Second-launch kernel: the tiles of tasks that read DECODED vectors (run-end encoded gathers, sparse unions).  Kept out of
// b2a_decode_kernel on purpose: with the extra op in its switch the hot kernel lost 3.5 % (register allocation).
__global__ void __launch_bounds__(kThreads)
    b2a_ree_kernel(const b2a_task *__restrict__ tasks, const uint32_t *__restrict__ tile_task, uint32_t tile_begin,
                   uint32_t tile_end, const uint8_t *__restrict__ image, uint8_t *__restrict__ out,
                   const uint64_t *__restrict__ variadic, uint32_t *err) {
	__shared__ b2a_task s_task;
	for (uint32_t tile = tile_begin + blockIdx.x; tile < tile_end; tile += gridDim.x) {
		__syncthreads();
		if (threadIdx.x < kTaskVecs) {
			reinterpret_cast<uint4 *>(&s_task)[threadIdx.x] =
			    __ldg(reinterpret_cast<const uint4 *>(tasks + __ldg(tile_task + tile)) + threadIdx.x);
		}
		__syncthreads();
		const b2a_task &t = s_task;
		const uint64_t row0 = static_cast<uint64_t>(tile - t.tile0) * t.tile_rows;
		uint64_t nrows = t.rows - row0;
		if (nrows > t.tile_rows) {
			nrows = t.tile_rows;
		}
		if (t.op == B2A_OP_REE) {
			op_ree(t, image, out, row0, nrows);
		} else if (t.op == B2A_OP_UNION) {
			op_union(t, image, out, row0, nrows, variadic, err);
		} else if (t.op == B2A_OP_BOOL8) {
			op_bool8(t, image, out, row0, nrows);
		} else if (t.op == B2A_OP_LISTVIEW32) {
			op_listview<uint32_t>(t, image, out, row0, nrows, err);
		} else if (t.op == B2A_OP_LISTVIEW64) {
			op_listview<uint64_t>(t, image, out, row0, nrows, err);
		}
	}
}

int LaunchPhase2(const DecodeLaunch &l, void *stream) {
	if (l.n_tiles == 0) {
		return 0;
	}
	uint32_t grid = static_cast<uint32_t>(l.sm_count) * 4;
	if (grid > l.n_tiles) {
		grid = l.n_tiles;
	}
	b2a_ree_kernel<<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
	    l.d_tasks, l.d_tile_task, l.tile_begin, l.tile_begin + l.n_tiles, l.image_origin, l.out, l.d_variadic, l.d_err);
	return static_cast<int>(cudaGetLastError());
}

} // namespace b2a
