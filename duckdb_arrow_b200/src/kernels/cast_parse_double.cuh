// VARCHAR -> DOUBLE as DuckDB casts it: TryDoubleCast (duckdb/src/include/duckdb/common/operator/double_cast_operator.hpp:
// 17-52, non-strict) over its vendored fast_float (third_party/fast_float/fast_float/fast_float.h: parse_number_string
// :507-735 — digits with '_' separators, fraction, exponent; from_chars :2462-2504; compute_float :1598-1693, the
// Eisel-Lemire algorithm; parse_infnan :2400-2441).  Restated here:
//   * the grammar, with the reference's bookkeeping of digit counts (an '_' of the integer part counts as a digit there);
//   * Clinger's fast path (|exponent| <= 22, mantissa <= 2^53: one exactly rounded IEEE operation);
//   * Eisel-Lemire: the 64-bit decimal mantissa times a 128-bit approximation of 5^q (cast_pow5_table.h), rounded to
//     nearest-even with the reference's corner cases (subnormals, the halfway test for small powers);
//   * more than 19 significant digits: the truncated mantissa w and w + 1 must round to the same double.
// NOT restated: the big-decimal fallback (parse_long_mantissa) the reference takes when Eisel-Lemire cannot decide or
// when w and w + 1 disagree — a handful of inputs per billion random ones; those report NEEDS_MORE_PRECISION and the
// caller raises the NotImplemented error it raised for every slow-path input before.
// `__host__ __device__`: pinned on the CPU against the compiled reference's TRY_CAST(s AS DOUBLE) (tests/test_cast_parse.py).
#pragma once
#include "cast_pow5_table.h"

#include <stddef.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define CASTD_HD __host__ __device__
#else
#define CASTD_HD
#endif

namespace b2a {
namespace castd {

enum Result : uint32_t { OK = 0, OUT_OF_RANGE = 1, NOT_A_NUMBER = 2, NEEDS_MORE_PRECISION = 3 };

static const uint64_t h_pow5[2 * 651] = B2A_POW5_TABLE;
#if defined(__CUDACC__)
static __device__ const uint64_t d_pow5[2 * 651] = B2A_POW5_TABLE;
#endif

CASTD_HD inline uint64_t pow5_word(int index) {
#if defined(__CUDA_ARCH__)
	return d_pow5[index];
#else
	return h_pow5[index];
#endif
}

CASTD_HD inline bool is_digit(char c) {
	return c >= '0' && c <= '9';
}
CASTD_HD inline bool is_space(char c) {
	return c == ' ' || (c >= '\t' && c <= '\r');
}
CASTD_HD inline char lower(char c) {
	return (c >= 'A' && c <= 'Z') ? static_cast<char>(c - 'A' + 'a') : c;
}

struct U128 {
	uint64_t low, high;
};
CASTD_HD inline U128 mul64(uint64_t a, uint64_t b) {
#if defined(__CUDA_ARCH__)
	return U128 {a * b, __umul64hi(a, b)};
#else
	const unsigned __int128 p = static_cast<unsigned __int128>(a) * b;
	return U128 {static_cast<uint64_t>(p), static_cast<uint64_t>(p >> 64)};
#endif
}
CASTD_HD inline int clz64(uint64_t v) { // v != 0
#if defined(__CUDA_ARCH__)
	return __clzll(static_cast<long long>(v));
#else
	return __builtin_clzll(v);
#endif
}

struct Adjusted {
	uint64_t mantissa;
	int32_t power2; // < 0: the algorithm cannot decide
};
CASTD_HD inline bool same(const Adjusted &a, const Adjusted &b) {
	return a.mantissa == b.mantissa && a.power2 == b.power2;
}

// w * 5^q to `precision` bits: one 64 x 64 product with the high word of the table entry, a second one with the low word
// only when the first cannot decide the bits that matter (fast_float.h:1546-1567)
CASTD_HD inline U128 product_approximation(int64_t q, uint64_t w, int precision) {
	const int index = 2 * static_cast<int>(q - B2A_POW5_SMALLEST);
	U128 first = mul64(w, pow5_word(index));
	const uint64_t mask = precision < 64 ? (~uint64_t(0) >> precision) : ~uint64_t(0);
	if ((first.high & mask) == mask) {
		const U128 second = mul64(w, pow5_word(index + 1));
		first.low += second.high;
		if (second.high > first.low) {
			first.high++;
		}
	}
	return first;
}

// w * 10^q -> {mantissa (52 explicit bits), biased exponent}; binary64: mantissa_explicit_bits 52, minimum_exponent -1023,
// infinite_power 0x7FF, powers of ten -342 .. 308, round-to-even window q in [-4, 23]
CASTD_HD inline Adjusted compute_float(int64_t q, uint64_t w) {
	Adjusted a {0, 0};
	if (w == 0 || q < -342) {
		return a;
	}
	if (q > 308) {
		a.power2 = 0x7FF;
		return a;
	}
	const int lz = clz64(w);
	w <<= lz;
	const U128 product = product_approximation(q, w, 52 + 3);
	if (product.low == ~uint64_t(0)) {
		if (!(q >= -27 && q <= 55)) {
			a.power2 = -1;
			return a;
		}
	}
	const int upperbit = static_cast<int>(product.high >> 63);
	a.mantissa = product.high >> (upperbit + 64 - 52 - 3);
	const int power = static_cast<int>(((152170 + 65536) * q) >> 16) + 63; // floor(log2(5^q)) + q + 63
	a.power2 = power + upperbit - lz + 1023;
	if (a.power2 <= 0) { // subnormal
		if (-a.power2 + 1 >= 64) {
			a.power2 = 0;
			a.mantissa = 0;
			return a;
		}
		a.mantissa >>= -a.power2 + 1;
		a.mantissa += (a.mantissa & 1);
		a.mantissa >>= 1;
		a.power2 = (a.mantissa < (uint64_t(1) << 52)) ? 0 : 1;
		return a;
	}
	// exactly half way between two doubles can only happen for small powers of five: round to even there
	if (product.low <= 1 && q >= -4 && q <= 23 && (a.mantissa & 3) == 1) {
		if ((a.mantissa << (upperbit + 64 - 52 - 3)) == product.high) {
			a.mantissa &= ~uint64_t(1);
		}
	}
	a.mantissa += (a.mantissa & 1);
	a.mantissa >>= 1;
	if (a.mantissa >= (uint64_t(2) << 52)) {
		a.mantissa = uint64_t(1) << 52;
		a.power2++;
	}
	a.mantissa &= ~(uint64_t(1) << 52);
	if (a.power2 >= 0x7FF) {
		a.power2 = 0x7FF;
		a.mantissa = 0;
	}
	return a;
}

CASTD_HD inline double to_double(bool negative, const Adjusted &a) {
	uint64_t word = a.mantissa | (static_cast<uint64_t>(a.power2) << 52);
	if (negative) {
		word |= uint64_t(1) << 63;
	}
	double d;
	memcpy(&d, &word, sizeof(d));
	return d;
}

// 10^0 .. 10^22: the powers of ten a double holds exactly (Clinger's fast path)
CASTD_HD inline double exact_pow10(int e) {
	double p = 1.0;
	for (int i = 0; i < e; i++) {
		p *= 10.0; // every intermediate is exact up to 10^22
	}
	return p;
}

// the '_' after a digit: skipped when a digit follows, the whole number is invalid otherwise
#define CASTD_SKIP_UNDERSCORE(p, end, counter)                                                                         \
	if ((p) != (end) && *(p) == '_') {                                                                                 \
		++(p);                                                                                                         \
		++(counter);                                                                                                   \
		if ((p) == (end) || !is_digit(*(p))) {                                                                         \
			return false;                                                                                              \
		}                                                                                                              \
	}

struct Parsed {
	int64_t exponent = 0;
	uint64_t mantissa = 0;
	const char *lastmatch = nullptr;
	bool negative = false;
	bool too_many_digits = false;
	bool reference_reads_past_end = false; // see parse_number
};

CASTD_HD inline bool eight_digits(const char *p) {
	for (int i = 0; i < 8; i++) {
		if (!is_digit(p[i])) {
			return false;
		}
	}
	return true;
}

// parse_number_string, fmt = general, non-strict, '.' (fast_float.h:507-735); false = "not valid" (inf / nan is tried next)
CASTD_HD inline bool parse_number(const char *p, const char *pend, Parsed &out) {
	out.negative = *p == '-';
	if (*p == '-') {
		++p;
		if (p == pend || (!is_digit(*p) && *p != '.')) {
			return false;
		}
	}
	const char *const start_digits = p;
	uint64_t i = 0;
	int64_t unused = 0;
	while (p != pend && is_digit(*p)) {
		i = 10 * i + static_cast<uint64_t>(*p - '0');
		++p;
		CASTD_SKIP_UNDERSCORE(p, pend, unused)
	}
	const char *const end_of_integer_part = p;
	int64_t digit_count = end_of_integer_part - start_digits; // the reference counts the '_' of the integer part too
	int64_t exponent = 0;
	if (p != pend && *p == '.') {
		++p;
		// the reference swallows up to two blocks of eight plain digits before its digit loop: an '_' right after such
		// a block is not skipped (it ends the number)
		if (pend - p >= 8 && eight_digits(p)) {
			for (int k = 0; k < 8; k++) {
				i = i * 10 + static_cast<uint64_t>(p[k] - '0');
			}
			p += 8;
			if (pend - p >= 8 && eight_digits(p)) {
				for (int k = 0; k < 8; k++) {
					i = i * 10 + static_cast<uint64_t>(p[k] - '0');
				}
				p += 8;
			}
		}
		int64_t skipped = 0;
		while (p != pend && is_digit(*p)) {
			i = i * 10 + static_cast<uint64_t>(*p - '0');
			++p;
			CASTD_SKIP_UNDERSCORE(p, pend, skipped)
		}
		exponent = (end_of_integer_part + 1 - p) + skipped;
		digit_count -= exponent;
	}
	if (digit_count == 0) {
		return false;
	}
	int64_t exp_number = 0;
	if (p != pend && (*p == 'e' || *p == 'E')) {
		const char *const location_of_e = p;
		++p;
		bool neg_exp = false;
		if (p != pend && *p == '-') {
			neg_exp = true;
			++p;
		} else if (p != pend && *p == '+') {
			++p;
		}
		if (p == pend || !is_digit(*p)) {
			p = location_of_e; // the 'e' is not part of the number
		} else {
			while (p != pend && is_digit(*p)) {
				if (exp_number < 0x10000) {
					exp_number = 10 * exp_number + (*p - '0');
				}
				++p;
				CASTD_SKIP_UNDERSCORE(p, pend, unused)
			}
			if (neg_exp) {
				exp_number = -exp_number;
			}
			exponent += exp_number;
		}
	}
	out.lastmatch = p;
	if (digit_count > 19) {
		const char *start = start_digits;
		while (start != pend && (*start == '0' || *start == '.')) {
			if (*start == '0') {
				digit_count--;
			}
			start++;
		}
		if (digit_count > 19) {
			// start again and stop at 19 digits: the mantissa is truncated, the exponent makes up for it
			out.too_many_digits = true;
			i = 0;
			p = start_digits;
			const uint64_t nineteen = 1000000000000000000ull;
			while (i < nineteen && p != pend && is_digit(*p)) {
				i = i * 10 + static_cast<uint64_t>(*p - '0');
				++p;
				CASTD_SKIP_UNDERSCORE(p, pend, unused)
			}
			if (i >= nineteen) {
				exponent = (end_of_integer_part - p) + exp_number;
			} else {
				if (p == pend || *p != '.') {
					// '_' separators made the integer part look longer than 19 digits and there is no fraction: the
					// reference steps over a character that is not there (fast_float.h:700) — not restated
					out.reference_reads_past_end = true;
					return true;
				}
				p++;
				const char *const first_after_period = p;
				int64_t skipped = 0;
				while (i < nineteen && p != pend && is_digit(*p)) {
					i = i * 10 + static_cast<uint64_t>(*p - '0');
					++p;
					CASTD_SKIP_UNDERSCORE(p, pend, skipped)
				}
				exponent = (first_after_period - p) + exp_number + skipped;
			}
		}
	}
	out.exponent = exponent;
	out.mantissa = i;
	return true;
}

CASTD_HD inline bool match_nocase(const char *p, const char *word, int n) {
	for (int i = 0; i < n; i++) {
		if (lower(p[i]) != word[i]) {
			return false;
		}
	}
	return true;
}

// parse_infnan (fast_float.h:2400-2441): [-]nan, [-]nan(chars), [-]inf, [-]infinity; returns the end of the match or nullptr
CASTD_HD inline const char *parse_infnan(const char *first, const char *last, double *value) {
	bool minus = false;
	if (*first == '-') {
		minus = true;
		++first;
	}
	if (last - first < 3) {
		return nullptr;
	}
	if (match_nocase(first, "nan", 3)) {
		first += 3;
		const char *end = first;
		const uint64_t bits = minus ? 0xFFF8000000000000ull : 0x7FF8000000000000ull;
		memcpy(value, &bits, sizeof(bits));
		if (first != last && *first == '(') {
			for (const char *ptr = first + 1; ptr != last; ++ptr) {
				if (*ptr == ')') {
					end = ptr + 1;
					break;
				}
				const char c = *ptr;
				if (!((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || is_digit(c) || c == '_')) {
					break;
				}
			}
		}
		return end;
	}
	if (match_nocase(first, "inf", 3)) {
		const uint64_t bits = minus ? 0xFFF0000000000000ull : 0x7FF0000000000000ull;
		memcpy(value, &bits, sizeof(bits));
		if (last - first >= 8 && match_nocase(first + 3, "inity", 5)) {
			return first + 8;
		}
		return first + 3;
	}
	return nullptr;
}

// TryDoubleCast<double>(buf, len, result, strict = false)
CASTD_HD inline uint32_t varchar_to_double(const char *buf, size_t len, double *out) {
	while (len > 0 && is_space(*buf)) {
		buf++;
		len--;
	}
	if (len == 0) {
		return NOT_A_NUMBER;
	}
	if (*buf == '+') {
		buf++;
		len--;
		if (len == 0) {
			return NOT_A_NUMBER;
		}
	}
	const char *const end = buf + len;
	Parsed pns;
	double value = 0;
	const char *matched;
	bool special = false;
	if (!parse_number(buf, end, pns)) {
		matched = parse_infnan(buf, end, &value);
		if (!matched) {
			return NOT_A_NUMBER;
		}
		special = true;
	} else {
		matched = pns.lastmatch;
	}
	while (matched < end && is_space(*matched)) {
		matched++;
	}
	if (matched != end) {
		return NOT_A_NUMBER;
	}
	if (special) {
		*out = value; // inf / nan
		return OK;
	}
	if (pns.reference_reads_past_end) {
		return NEEDS_MORE_PRECISION;
	}
	if (pns.exponent >= -22 && pns.exponent <= 22 && pns.mantissa <= (uint64_t(2) << 52) && !pns.too_many_digits) {
		value = static_cast<double>(pns.mantissa);
		if (pns.exponent < 0) {
			value = value / exact_pow10(static_cast<int>(-pns.exponent));
		} else {
			value = value * exact_pow10(static_cast<int>(pns.exponent));
		}
		*out = pns.negative ? -value : value;
		return OK;
	}
	Adjusted am = compute_float(pns.exponent, pns.mantissa);
	if (pns.too_many_digits && am.power2 >= 0) {
		if (!same(am, compute_float(pns.exponent, pns.mantissa + 1))) {
			am.power2 = -1;
		}
	}
	if (am.power2 < 0) {
		return NEEDS_MORE_PRECISION; // the reference's parse_long_mantissa
	}
	*out = to_double(pns.negative, am);
	return OK;
}

#undef CASTD_SKIP_UNDERSCORE

} // namespace castd
} // namespace b2a
