// b2a_cast_kernel: the column casts of a multi-file scan (global column type != the file's column type), evaluated
// over whole decoded record batches.  One launch walks a tile table (2 048 rows per tile = one DuckDB vector), one
// row per thread.  NULL rows are not converted (their data slot is zeroed), like DuckDB's UnaryExecutor.
//
// Semantics follow DuckDB's cast operators:
//   integer -> integer      NumericTryCast / TryCastWithOverflowCheck   (numeric_cast.hpp: range check on narrowing)
//   integer -> FLOAT/DOUBLE static_cast
//   FLOAT/DOUBLE -> integer TryCastWithOverflowCheckFloat               (numeric_cast.hpp:73-84: finite, min <= v < max, nearbyint)
//   FLOAT -> DOUBLE exact; DOUBLE -> FLOAT overflow to +-inf is an error (numeric_cast.hpp)
//   integer -> VARCHAR      NumericHelper::FormatSigned                 (decimal digits, '-' sign)
//   FLOAT/DOUBLE -> VARCHAR StringCast::Operation<double> = fmt "{}"    (shortest round-trip digits; fixed notation for
//                           -4 <= exp10 < 16 with a trailing ".0" for integers, else d.ddde[+-]XX; nan / inf / -inf)
//   VARCHAR -> DOUBLE       TryCast::Operation<string_t,double> = TryDoubleCast over fast_float, restated in
//                           cast_parse_double.cuh (grammar, Clinger, Eisel-Lemire, the big-decimal fallback, inf / nan)
//   VARCHAR -> BIGINT       TryCast::Operation<string_t,int64_t>: optional sign, digits, optional fraction that rounds
//                           half away from zero (IntegerCastLoop: decimals are rounded)
//   DATE -> TIMESTAMP       days * 86 400 000 000 us, +-infinity preserved
//   VARCHAR -> DATE / TIMESTAMP   Date::TryConvertDate / Timestamp::TryConvertTimestamp, restated in cast_parse_temporal.cuh
//   VARCHAR -> DECIMAL      TryDecimalStringCast, restated in cast_parse_decimal.cuh
//   BOOLEAN / DATE / TIME / TIMESTAMP / DECIMAL / HUGEINT -> VARCHAR   StringCast::Operation, restated in cast_format.cuh
//   integer / FLOAT / DOUBLE -> DECIMAL, integer -> HUGEINT, DECIMAL -> DECIMAL / integer / DOUBLE / FLOAT   cast_decimal.cuh
#include "cast.cuh"
#include "cast_parse.cuh"
#include "cast_parse_double.cuh"
#include "cast_format.cuh"
#include "cast_decimal.cuh"
#include "cast_parse_temporal.cuh"
#include "cast_parse_decimal.cuh"

#include "b2arrow.h"

#include <cuda_runtime.h>

namespace b2a {

#define NONE_OFF 0xFFFFFFFFFFFFFFFFull

__device__ __forceinline__ bool is_signed_int(int t) {
	return t == B2A_T_TINYINT || t == B2A_T_SMALLINT || t == B2A_T_INTEGER || t == B2A_T_BIGINT;
}
__device__ __forceinline__ bool is_unsigned_int(int t) {
	return t == B2A_T_UTINYINT || t == B2A_T_USMALLINT || t == B2A_T_UINTEGER || t == B2A_T_UBIGINT;
}

__device__ __forceinline__ void raise(uint32_t *d_err, uint32_t code) {
	atomicMax(d_err, code);
}

// ---- loads / stores of little-endian integers of 1/2/4/8 bytes
__device__ __forceinline__ int64_t load_signed(const uint8_t *p, int w) {
	switch (w) {
	case 1:
		return *reinterpret_cast<const int8_t *>(p);
	case 2:
		return *reinterpret_cast<const int16_t *>(p);
	case 4:
		return *reinterpret_cast<const int32_t *>(p);
	default:
		return *reinterpret_cast<const int64_t *>(p);
	}
}
__device__ __forceinline__ uint64_t load_unsigned(const uint8_t *p, int w) {
	switch (w) {
	case 1:
		return *p;
	case 2:
		return *reinterpret_cast<const uint16_t *>(p);
	case 4:
		return *reinterpret_cast<const uint32_t *>(p);
	default:
		return *reinterpret_cast<const uint64_t *>(p);
	}
}
__device__ __forceinline__ void store_int(uint8_t *p, int w, uint64_t v) {
	switch (w) {
	case 1:
		*p = static_cast<uint8_t>(v);
		break;
	case 2:
		*reinterpret_cast<uint16_t *>(p) = static_cast<uint16_t>(v);
		break;
	case 4:
		*reinterpret_cast<uint32_t *>(p) = static_cast<uint32_t>(v);
		break;
	default:
		*reinterpret_cast<uint64_t *>(p) = v;
	}
}

// range of an integer destination type as doubles [lo, hi) and as integers
__device__ __forceinline__ bool int_fits(bool src_signed, int64_t sv, uint64_t uv, int to, int w) {
	if (is_signed_int(to)) {
		const int64_t hi = w == 8 ? INT64_MAX : (1ll << (8 * w - 1)) - 1, lo = -hi - 1;
		if (src_signed) {
			return sv >= lo && sv <= hi;
		}
		return uv <= static_cast<uint64_t>(hi);
	}
	const uint64_t hi = w == 8 ? UINT64_MAX : (1ull << (8 * w)) - 1;
	if (src_signed) {
		return sv >= 0 && static_cast<uint64_t>(sv) <= hi;
	}
	return uv <= hi;
}

__device__ __forceinline__ __int128 load_i128(const uint8_t *p, int w) { // a signed integer of 2 / 4 / 8 / 16 bytes
	if (w == 16) {
		const uint64_t lo = *reinterpret_cast<const uint64_t *>(p);
		const int64_t hi = *reinterpret_cast<const int64_t *>(p + 8);
		return (static_cast<__int128>(hi) << 64) | static_cast<__int128>(lo);
	}
	return load_signed(p, w);
}
__device__ __forceinline__ void store_i128(uint8_t *p, int w, __int128 v) {
	if (w == 16) {
		*reinterpret_cast<uint64_t *>(p) = static_cast<uint64_t>(v);
		*reinterpret_cast<int64_t *>(p + 8) = static_cast<int64_t>(v >> 64);
	} else {
		store_int(p, w, static_cast<uint64_t>(v));
	}
}

// ---- VARCHAR results: DuckDB string_t (len, 12 inlined bytes | 4 prefix bytes + pointer)
__device__ __forceinline__ void store_string(uint8_t *dst, uint8_t *heap, uint64_t heap_addr, const char *s, uint32_t n) {
	uint32_t w[4] = {n, 0, 0, 0};
	uint8_t *b = reinterpret_cast<uint8_t *>(w);
	if (n <= 12) {
		for (uint32_t i = 0; i < n; i++) {
			b[4 + i] = static_cast<uint8_t>(s[i]);
		}
	} else {
		for (uint32_t i = 0; i < 4; i++) {
			b[4 + i] = static_cast<uint8_t>(s[i]);
		}
		for (uint32_t i = 0; i < n; i++) {
			heap[i] = static_cast<uint8_t>(s[i]);
		}
		w[2] = static_cast<uint32_t>(heap_addr);
		w[3] = static_cast<uint32_t>(heap_addr >> 32);
	}
	*reinterpret_cast<uint4 *>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
}

__device__ __forceinline__ uint32_t format_u64(uint64_t v, bool neg, char *out) {
	char tmp[20];
	int n = 0;
	do {
		tmp[n++] = static_cast<char>('0' + v % 10);
		v /= 10;
	} while (v);
	uint32_t k = 0;
	if (neg) {
		out[k++] = '-';
	}
	while (n) {
		out[k++] = tmp[--n];
	}
	return k;
}

// ---- shortest round-trip decimal digits of a finite positive double (Steele & White / dragon4 with a fixed-size
// bignum): exact, table-free.  Produces digits[0..n) and the decimal exponent of the first digit.
struct Big {
	static const int W = 36; // 36 * 32 = 1 152 bits >= 2^1074 * 10^k headroom
	uint32_t w[W];
	int n; // used words
	__device__ void set(uint64_t v) {
		for (int i = 0; i < W; i++) {
			w[i] = 0;
		}
		w[0] = static_cast<uint32_t>(v);
		w[1] = static_cast<uint32_t>(v >> 32);
		n = w[1] ? 2 : (w[0] ? 1 : 0);
	}
	__device__ void mul_small(uint32_t m) {
		uint64_t carry = 0;
		for (int i = 0; i < n; i++) {
			uint64_t t = static_cast<uint64_t>(w[i]) * m + carry;
			w[i] = static_cast<uint32_t>(t);
			carry = t >> 32;
		}
		if (carry && n < W) {
			w[n++] = static_cast<uint32_t>(carry);
		}
	}
	__device__ void shl(int bits) {
		const int ws = bits >> 5, bs = bits & 31;
		if (n == 0) {
			return;
		}
		int top = n + ws + 1;
		if (top > W) {
			top = W;
		}
		for (int i = top - 1; i >= 0; i--) {
			uint32_t lo = 0, hi = 0;
			const int j = i - ws;
			if (j >= 0 && j < n) {
				hi = w[j];
			}
			if (j - 1 >= 0 && j - 1 < n) {
				lo = w[j - 1];
			}
			uint32_t v = bs ? ((hi << bs) | (lo >> (32 - bs))) : hi;
			w[i] = v;
		}
		n = top;
		while (n > 0 && w[n - 1] == 0) {
			n--;
		}
	}
	__device__ void mul_pow10(int k) {
		// 10^k = 5^k * 2^k
		for (int i = 0; i + 13 <= k; i += 13) {
			mul_small(1220703125u); // 5^13
		}
		for (int i = 0; i < k % 13; i++) {
			mul_small(5u);
		}
		shl(k);
	}
	__device__ int cmp(const Big &o) const {
		if (n != o.n) {
			return n < o.n ? -1 : 1;
		}
		for (int i = n - 1; i >= 0; i--) {
			if (w[i] != o.w[i]) {
				return w[i] < o.w[i] ? -1 : 1;
			}
		}
		return 0;
	}
	__device__ void sub(const Big &o) { // this -= o, this >= o
		int64_t borrow = 0;
		for (int i = 0; i < n; i++) {
			int64_t t = static_cast<int64_t>(w[i]) - (i < o.n ? o.w[i] : 0) - borrow;
			borrow = t < 0;
			w[i] = static_cast<uint32_t>(t);
		}
		while (n > 0 && w[n - 1] == 0) {
			n--;
		}
	}
	__device__ void add(const Big &o) {
		uint64_t carry = 0;
		const int m = n > o.n ? n : o.n;
		for (int i = 0; i < m; i++) {
			uint64_t t = static_cast<uint64_t>(i < n ? w[i] : 0) + (i < o.n ? o.w[i] : 0) + carry;
			w[i] = static_cast<uint32_t>(t);
			carry = t >> 32;
		}
		n = m;
		if (carry && n < W) {
			w[n++] = static_cast<uint32_t>(carry);
		}
	}
};

// returns the number of digits; *exp10 = decimal exponent such that value = 0.d1d2... * 10^(*exp10)
__device__ __noinline__ int shortest_digits(uint64_t mant, int exp2, bool boundary_low_closer, bool even, char *digits, int *exp10) {
	// value = mant * 2^exp2.  R/S = value, M+ / M- = half-gaps to the neighbours (all scaled by the same S)
	Big R, S, Mp, Mm;
	if (exp2 >= 0) {
		if (!boundary_low_closer) {
			R.set(mant);
			R.shl(exp2 + 1);
			S.set(2);
			Mp.set(1);
			Mp.shl(exp2);
			Mm.set(1);
			Mm.shl(exp2);
		} else {
			R.set(mant);
			R.shl(exp2 + 2);
			S.set(4);
			Mp.set(1);
			Mp.shl(exp2 + 1);
			Mm.set(1);
			Mm.shl(exp2);
		}
	} else {
		if (!boundary_low_closer) {
			R.set(mant);
			R.shl(1);
			S.set(1);
			S.shl(-exp2 + 1);
			Mp.set(1);
			Mm.set(1);
		} else {
			R.set(mant);
			R.shl(2);
			S.set(1);
			S.shl(-exp2 + 2);
			Mp.set(2);
			Mm.set(1);
		}
	}
	// estimate k = ceil(log10(value)); fix up below
	const int bits = 64 - __clzll(mant) + exp2; // value in [2^(bits-1), 2^bits)
	int k = static_cast<int>(ceil((bits - 1) * 0.30102999566398120 - 1e-10));
	if (k >= 0) {
		S.mul_pow10(k);
	} else {
		R.mul_pow10(-k);
		Mp.mul_pow10(-k);
		Mm.mul_pow10(-k);
	}
	// fixup: ensure R + M+ < S (or <= when even)  <=>  first digit lands in [1, 9]
	{
		Big t = R;
		t.add(Mp);
		const int c = t.cmp(S);
		if (even ? c >= 0 : c > 0) {
			k++;
		} else {
			R.mul_small(10);
			Mp.mul_small(10);
			Mm.mul_small(10);
		}
	}
	*exp10 = k;
	int n = 0;
	while (true) {
		// digit = floor(R / S) (0..9) by repeated subtraction
		int d = 0;
		while (R.cmp(S) >= 0) {
			R.sub(S);
			d++;
		}
		const int cl = R.cmp(Mm);
		const bool low = even ? cl <= 0 : cl < 0;
		Big t = R;
		t.add(Mp);
		const int ch = t.cmp(S);
		const bool high = even ? ch >= 0 : ch > 0;
		if (!low && !high) {
			digits[n++] = static_cast<char>('0' + d);
			R.mul_small(10);
			Mp.mul_small(10);
			Mm.mul_small(10);
			if (n >= 17) {
				// cannot happen for doubles (17 digits always suffice); guard against an endless loop
				break;
			}
			continue;
		}
		if (low && !high) {
			digits[n++] = static_cast<char>('0' + d);
		} else if (high && !low) {
			digits[n++] = static_cast<char>('0' + d + 1);
		} else {
			// both: round to the closer one (2R vs S)
			Big r2 = R;
			r2.shl(1);
			const int c2 = r2.cmp(S);
			digits[n++] = static_cast<char>('0' + (c2 < 0 ? d : d + 1));
		}
		break;
	}
	// a final digit of 10 carries
	for (int i = n - 1; i >= 0 && digits[i] > '9'; i--) {
		digits[i] = '0';
		if (i == 0) {
			// shift right, leading 1
			for (int j = n; j > 0; j--) {
				digits[j] = digits[j - 1];
			}
			digits[0] = '1';
			n++;
			(*exp10)++;
		} else {
			digits[i - 1]++;
		}
	}
	while (n > 1 && digits[n - 1] == '0') {
		n--;
	}
	return n;
}

// fmt "{}" of a double (or of a float widened the way fmt does it: shortest digits of the FLOAT value)
__device__ __noinline__ uint32_t format_double(double v, bool is_float, char *out) {
	uint32_t k = 0;
	if (isnan(v)) {
		out[0] = 'n';
		out[1] = 'a';
		out[2] = 'n';
		return 3;
	}
	if (signbit(v)) {
		out[k++] = '-';
		v = -v;
	}
	if (isinf(v)) {
		out[k++] = 'i';
		out[k++] = 'n';
		out[k++] = 'f';
		return k;
	}
	if (v == 0.0) {
		out[k++] = '0';
		out[k++] = '.';
		out[k++] = '0';
		return k;
	}
	char digits[20];
	int e10, n;
	if (is_float) {
		const uint32_t bits = __float_as_uint(static_cast<float>(v));
		const uint32_t frac = bits & 0x7FFFFFu, be = (bits >> 23) & 0xFF;
		const uint64_t mant = be ? (frac | 0x800000u) : frac;
		const int exp2 = be ? static_cast<int>(be) - 150 : -149;
		n = shortest_digits(mant, exp2, frac == 0 && be > 1, (mant & 1) == 0, digits, &e10);
	} else {
		const uint64_t bits = static_cast<uint64_t>(__double_as_longlong(v));
		const uint64_t frac = bits & 0xFFFFFFFFFFFFFull, be = (bits >> 52) & 0x7FF;
		const uint64_t mant = be ? (frac | (1ull << 52)) : frac;
		const int exp2 = be ? static_cast<int>(be) - 1075 : -1074;
		n = shortest_digits(mant, exp2, frac == 0 && be > 1, (mant & 1) == 0, digits, &e10);
	}
	// value = 0.d1..dn * 10^e10  =>  scientific exponent = e10 - 1
	const int sci = e10 - 1;
	if (sci >= -4 && sci < 16) {
		if (e10 <= 0) {
			out[k++] = '0';
			out[k++] = '.';
			for (int i = 0; i < -e10; i++) {
				out[k++] = '0';
			}
			for (int i = 0; i < n; i++) {
				out[k++] = digits[i];
			}
		} else {
			for (int i = 0; i < e10; i++) {
				out[k++] = i < n ? digits[i] : '0';
			}
			out[k++] = '.';
			if (n > e10) {
				for (int i = e10; i < n; i++) {
					out[k++] = digits[i];
				}
			} else {
				out[k++] = '0';
			}
		}
		return k;
	}
	out[k++] = digits[0];
	if (n > 1) {
		out[k++] = '.';
		for (int i = 1; i < n; i++) {
			out[k++] = digits[i];
		}
	}
	out[k++] = 'e';
	int e = sci;
	out[k++] = e < 0 ? '-' : '+';
	if (e < 0) {
		e = -e;
	}
	if (e >= 100) {
		out[k++] = static_cast<char>('0' + e / 100);
		e %= 100;
		out[k++] = static_cast<char>('0' + e / 10);
		out[k++] = static_cast<char>('0' + e % 10);
	} else {
		out[k++] = static_cast<char>('0' + e / 10);
		out[k++] = static_cast<char>('0' + e % 10);
	}
	return k;
}

// ---- VARCHAR sources
__device__ __forceinline__ const char *string_chars(const uint8_t *rec, int64_t ptr_delta, uint32_t *len) {
	const uint32_t n = *reinterpret_cast<const uint32_t *>(rec);
	*len = n;
	if (n <= 12) {
		return reinterpret_cast<const char *>(rec + 4);
	}
	const uint64_t p = *reinterpret_cast<const uint64_t *>(rec + 8);
	return reinterpret_cast<const char *>(static_cast<uintptr_t>(static_cast<int64_t>(p) + ptr_delta));
}

// VARCHAR -> DOUBLE: DuckDB's TryDoubleCast over its vendored fast_float (grammar with '_' separators, Clinger's fast path,
// Eisel-Lemire, inf / nan), restated in cast_parse_double.cuh and pinned on the CPU against the compiled reference
// (tests/test_cast_parse.py).  B2A_CAST_DEV_PRECISION only for the one input shape on which the reference reads past the
// end of its string (see the header).
__device__ __noinline__ uint32_t parse_double(const char *s, uint32_t n, double *out) {
	return castd::varchar_to_double(s, n, out);
}
static_assert(castd::NOT_A_NUMBER == B2A_CAST_DEV_PARSE && castd::NEEDS_MORE_PRECISION == B2A_CAST_DEV_PRECISION,
              "cast error codes");

// VARCHAR -> BIGINT: DuckDB's TrySimpleIntegerCast<int64_t>, restated in cast_parse.cuh (pinned on the CPU against the compiled
// reference: tests/test_cast_parse.py).  Result codes are B2A_CAST_DEV_RANGE / B2A_CAST_DEV_PARSE.
__device__ __noinline__ uint32_t parse_int64(const char *s, uint32_t n, int64_t *out) {
	return castp::varchar_to_bigint(s, n, out);
}
static_assert(castp::OUT_OF_RANGE == B2A_CAST_DEV_RANGE && castp::NOT_A_NUMBER == B2A_CAST_DEV_PARSE, "cast error codes");
static_assert(castt::OUT_OF_RANGE == B2A_CAST_DEV_RANGE && castt::BAD_FORMAT == B2A_CAST_DEV_PARSE, "cast error codes");

// casts that cannot fail and do not need the validity mask (a NULL row's data slot is unspecified in DuckDB):
// integer widening, integer -> FLOAT/DOUBLE, FLOAT -> DOUBLE.  8 rows in flight per thread.
__device__ __forceinline__ bool never_fails(int from, int to, int sw, int dw) {
	const bool ss = is_signed_int(from), su = is_unsigned_int(from);
	if (ss || su) {
		if (to == B2A_T_DOUBLE || to == B2A_T_FLOAT) {
			return true;
		}
		if (is_signed_int(to)) {
			return ss ? dw >= sw : dw > sw;
		}
		return is_unsigned_int(to) && su && dw >= sw;
	}
	return from == B2A_T_FLOAT && to == B2A_T_DOUBLE;
}

__device__ __forceinline__ uint64_t widen_bits(int from, int to, uint64_t raw, int sw) {
	if (from == B2A_T_FLOAT) {
		return static_cast<uint64_t>(__double_as_longlong(static_cast<double>(__uint_as_float(static_cast<uint32_t>(raw)))));
	}
	const bool ss = is_signed_int(from);
	int64_t sv = static_cast<int64_t>(raw);
	if (ss && sw < 8) {
		const int sh = 64 - 8 * sw;
		sv = (sv << sh) >> sh;
	}
	if (to == B2A_T_DOUBLE) {
		return static_cast<uint64_t>(__double_as_longlong(ss ? static_cast<double>(sv) : static_cast<double>(raw)));
	}
	if (to == B2A_T_FLOAT) {
		return __float_as_uint(ss ? static_cast<float>(sv) : static_cast<float>(raw));
	}
	return static_cast<uint64_t>(sv);
}

__device__ __forceinline__ void widen_tile(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, uint32_t n, int from,
                                           int to, int sw, int dw) {
	constexpr int U = 8;
	uint32_t r = threadIdx.x;
	for (; r + (U - 1) * 256 < n; r += U * 256) {
		uint64_t raw[U];
#pragma unroll
		for (int u = 0; u < U; u++) {
			raw[u] = load_unsigned(src + static_cast<uint64_t>(r + u * 256) * sw, sw);
		}
#pragma unroll
		for (int u = 0; u < U; u++) {
			store_int(dst + static_cast<uint64_t>(r + u * 256) * dw, dw, widen_bits(from, to, raw[u], sw));
		}
	}
	for (; r < n; r += 256) {
		store_int(dst + static_cast<uint64_t>(r) * dw, dw, widen_bits(from, to, load_unsigned(src + static_cast<uint64_t>(r) * sw, sw), sw));
	}
}

__global__ void __launch_bounds__(256, 4) b2a_cast_kernel(CastLaunch l) {
	for (uint32_t tile = blockIdx.x; tile < l.n_tiles; tile += gridDim.x) {
		const b2a_cast_task t = l.d_tasks[l.d_tile_task[tile]];
		const uint64_t row0 = static_cast<uint64_t>(tile - t.tile0) * B2A_VECTOR_SIZE;
		const uint64_t left = t.rows - row0;
		const uint32_t n = left < B2A_VECTOR_SIZE ? static_cast<uint32_t>(left) : B2A_VECTOR_SIZE;
		const uint64_t *mask =
		    t.validity == NONE_OFF ? nullptr : reinterpret_cast<const uint64_t *>(l.scan_arena + t.validity);
		const int from = t.from, to = t.to, sw = t.src_width, dw = t.dst_width;
		if (never_fails(from, to, sw, dw)) {
			widen_tile(l.scan_arena + t.src + row0 * sw, l.final_arena + t.dst + row0 * dw, n, from, to, sw, dw);
			continue;
		}
		for (uint32_t r = threadIdx.x; r < n; r += blockDim.x) {
			const uint64_t row = row0 + r;
			uint8_t *dst = l.final_arena + t.dst + row * dw;
			const bool valid = !mask || ((mask[row >> 6] >> (row & 63)) & 1ull);
			if (!valid) {
				if (dw == 16) {
					*reinterpret_cast<uint4 *>(dst) = make_uint4(0, 0, 0, 0);
				} else {
					store_int(dst, dw, 0);
				}
				continue;
			}
			const uint8_t *src = l.scan_arena + t.src + row * sw;
			const bool s_sint = is_signed_int(from), s_uint = is_unsigned_int(from);
			if (s_sint || s_uint) {
				const int64_t sv = s_sint ? load_signed(src, sw) : 0;
				const uint64_t uv = s_uint ? load_unsigned(src, sw) : 0;
				if (is_signed_int(to) || is_unsigned_int(to)) {
					if (!int_fits(s_sint, sv, uv, to, dw)) {
						raise(l.d_err, B2A_CAST_DEV_RANGE);
						continue;
					}
					store_int(dst, dw, s_sint ? static_cast<uint64_t>(sv) : uv);
				} else if (to == B2A_T_HUGEINT) {
					store_i128(dst, 16, s_sint ? static_cast<__int128>(sv) : static_cast<__int128>(uv));
				} else if (to == B2A_T_DECIMAL) {
					__int128 r;
					if (!castx::int_to_decimal(s_sint ? static_cast<__int128>(sv) : static_cast<__int128>(uv),
					                           t.dst_decimal_width, t.dst_decimal_scale, &r)) {
						raise(l.d_err, B2A_CAST_DEV_RANGE);
						continue;
					}
					store_i128(dst, dw, r);
				} else if (to == B2A_T_DOUBLE) {
					*reinterpret_cast<double *>(dst) = s_sint ? static_cast<double>(sv) : static_cast<double>(uv);
				} else if (to == B2A_T_FLOAT) {
					*reinterpret_cast<float *>(dst) = s_sint ? static_cast<float>(sv) : static_cast<float>(uv);
				} else if (to == B2A_T_VARCHAR) {
					char buf[24];
					const uint32_t k = s_sint ? format_u64(sv < 0 ? 0 - static_cast<uint64_t>(sv) : static_cast<uint64_t>(sv),
					                                       sv < 0, buf)
					                          : format_u64(uv, false, buf);
					const uint64_t h = t.heap + row * t.heap_stride;
					store_string(dst, l.final_arena + h, l.dst_ptr_origin + h, buf, k);
				}
			} else if (from == B2A_T_FLOAT || from == B2A_T_DOUBLE) {
				const double v = from == B2A_T_FLOAT ? static_cast<double>(*reinterpret_cast<const float *>(src))
				                                     : *reinterpret_cast<const double *>(src);
				if (to == B2A_T_DOUBLE) {
					*reinterpret_cast<double *>(dst) = v;
				} else if (to == B2A_T_FLOAT) {
					const float f = static_cast<float>(v);
					if (isfinite(v) && !isfinite(f)) {
						raise(l.d_err, B2A_CAST_DEV_RANGE);
						continue;
					}
					*reinterpret_cast<float *>(dst) = f;
				} else if (to == B2A_T_DECIMAL) {
					__int128 r;
					if (!castx::real_to_decimal(v, from == B2A_T_FLOAT, t.dst_decimal_width, t.dst_decimal_scale, &r)) {
						raise(l.d_err, B2A_CAST_DEV_RANGE);
						continue;
					}
					store_i128(dst, dw, r);
				} else if (to == B2A_T_VARCHAR) {
					char buf[32];
					const uint32_t k = format_double(v, from == B2A_T_FLOAT, buf);
					const uint64_t h = t.heap + row * t.heap_stride;
					store_string(dst, l.final_arena + h, l.dst_ptr_origin + h, buf, k);
				} else {
					// TryCastWithOverflowCheckFloat: finite, lo <= v < hi, nearbyint (round half to even)
					double lo, hi;
					if (is_signed_int(to)) {
						hi = ldexp(1.0, 8 * dw - 1);
						lo = -hi;
					} else {
						hi = ldexp(1.0, 8 * dw);
						lo = 0.0;
					}
					if (!isfinite(v) || !(v >= lo && v < hi)) {
						raise(l.d_err, B2A_CAST_DEV_RANGE);
						continue;
					}
					const double rv = rint(v);
					if (is_signed_int(to)) {
						store_int(dst, dw, static_cast<uint64_t>(static_cast<int64_t>(rv)));
					} else {
						store_int(dst, dw, static_cast<uint64_t>(rv));
					}
				}
			} else if (from == B2A_T_VARCHAR) {
				uint32_t len;
				const char *s = string_chars(src, l.src_ptr_delta, &len);
				if (to == B2A_T_DOUBLE) {
					double d = 0;
					const uint32_t e = parse_double(s, len, &d);
					if (e) {
						raise(l.d_err, e);
						continue;
					}
					*reinterpret_cast<double *>(dst) = d;
				} else if (to == B2A_T_DECIMAL) {
					__int128 r;
					if (!castq::varchar_to_decimal(s, len, t.dst_decimal_width, t.dst_decimal_scale, &r)) {
						raise(l.d_err, B2A_CAST_DEV_PARSE);
						continue;
					}
					store_i128(dst, dw, r);
				} else if (to == B2A_T_DATE) {
					int32_t d = 0;
					const uint32_t e = castt::varchar_to_date(s, len, &d);
					if (e) {
						raise(l.d_err, e);
						continue;
					}
					*reinterpret_cast<int32_t *>(dst) = d;
				} else if (to == B2A_T_TIMESTAMP) {
					int64_t v = 0;
					const uint32_t e = castt::varchar_to_timestamp(s, len, &v);
					if (e) {
						raise(l.d_err, e);
						continue;
					}
					*reinterpret_cast<int64_t *>(dst) = v;
				} else {
					int64_t v = 0;
					const uint32_t e = parse_int64(s, len, &v);
					if (e) {
						raise(l.d_err, e);
						continue;
					}
					*reinterpret_cast<int64_t *>(dst) = v;
				}
			} else if (from == B2A_T_DECIMAL && to != B2A_T_VARCHAR) { // cast_decimal.cuh
				const __int128 v = load_i128(src, sw);
				if (to == B2A_T_DOUBLE) {
					*reinterpret_cast<double *>(dst) = castx::decimal_to_double(v, sw, t.decimal_scale);
				} else if (to == B2A_T_FLOAT) {
					*reinterpret_cast<float *>(dst) = castx::decimal_to_float(v, sw, t.decimal_scale);
				} else if (is_signed_int(to) || is_unsigned_int(to)) {
					const __int128 one = 1;
					const bool sg = is_signed_int(to);
					__int128 r;
					if (!castx::decimal_to_int(v, t.decimal_scale, sg ? -(one << (8 * dw - 1)) : 0,
					                           sg ? (one << (8 * dw - 1)) - 1 : (one << (8 * dw)) - 1, &r)) {
						raise(l.d_err, B2A_CAST_DEV_RANGE);
						continue;
					}
					store_int(dst, dw, static_cast<uint64_t>(r));
				} else {
					__int128 r;
					if (!castx::decimal_to_decimal(v, t.decimal_width, t.decimal_scale, t.dst_decimal_width,
					                               t.dst_decimal_scale, &r)) {
						raise(l.d_err, B2A_CAST_DEV_RANGE);
						continue;
					}
					store_i128(dst, dw, r);
				}
			} else if (to == B2A_T_VARCHAR) { // BOOLEAN / DATE / TIME / TIMESTAMP / DECIMAL / HUGEINT (cast_format.cuh)
				char buf[castf::kMaxChars];
				uint32_t k = 0;
				if (from == B2A_T_BOOLEAN) {
					k = castf::format_bool(*src != 0, buf);
				} else if (from == B2A_T_DATE) {
					k = castf::format_date(*reinterpret_cast<const int32_t *>(src), buf);
				} else if (from == B2A_T_TIME) {
					k = castf::format_time(*reinterpret_cast<const int64_t *>(src), buf);
				} else if (from == B2A_T_TIMESTAMP) {
					k = castf::format_timestamp(*reinterpret_cast<const int64_t *>(src), buf);
					if (k == 0) {
						raise(l.d_err, B2A_CAST_DEV_RANGE);
						continue;
					}
				} else { // DECIMAL in its physical width, HUGEINT
					k = castf::format_decimal(load_i128(src, sw), t.decimal_width, t.decimal_scale, buf);
				}
				const uint64_t h = t.heap + row * t.heap_stride;
				store_string(dst, l.final_arena + h, l.dst_ptr_origin + h, buf, k);
			} else if (from == B2A_T_DATE && to == B2A_T_TIMESTAMP) {
				const int32_t d = *reinterpret_cast<const int32_t *>(src);
				int64_t us;
				if (d == INT32_MAX) {
					us = INT64_MAX;
				} else if (d == -INT32_MAX) {
					us = -INT64_MAX;
				} else {
					us = static_cast<int64_t>(d) * 86400000000ll;
					if (us / 86400000000ll != d) {
						raise(l.d_err, B2A_CAST_DEV_RANGE);
						continue;
					}
				}
				*reinterpret_cast<int64_t *>(dst) = us;
			}
		}
	}
}

int LaunchCast(const CastLaunch &l, void *stream) {
	if (l.n_tiles == 0) {
		return 0;
	}
	uint32_t grid = static_cast<uint32_t>(l.sm_count) * 8;
	if (grid > l.n_tiles) {
		grid = l.n_tiles;
	}
	b2a_cast_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(l);
	return static_cast<int>(cudaGetLastError());
}

} // namespace b2a
