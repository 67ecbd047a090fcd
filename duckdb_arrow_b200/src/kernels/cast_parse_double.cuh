// VARCHAR -> DOUBLE as DuckDB casts it: TryDoubleCast (duckdb/src/include/duckdb/common/operator/double_cast_operator.hpp:
// 17-52, non-strict) over its vendored fast_float (third_party/fast_float/fast_float/fast_float.h: parse_number_string
// :507-735 — digits with '_' separators, fraction, exponent; from_chars :2462-2504; compute_float :1598-1693, the
// Eisel-Lemire algorithm; parse_infnan :2400-2441).  Restated here:
//   * the grammar, with the reference's bookkeeping of digit counts (an '_' of the integer part counts as a digit there);
//   * Clinger's fast path (|exponent| <= 22, mantissa <= 2^53: one exactly rounded IEEE operation);
//   * Eisel-Lemire: the 64-bit decimal mantissa times a 128-bit approximation of 5^q (cast_pow5_table.h), rounded to
//     nearest-even with the reference's corner cases (subnormals, the halfway test for small powers);
//   * more than 19 significant digits: the truncated mantissa w and w + 1 must round to the same double.
//   * the fallback when Eisel-Lemire cannot decide or w and w + 1 disagree (parse_long_mantissa: up to 768 decimal digits
//     shifted by powers of two, exact) — with the reference's quirk that this second parse stops at an '_'.
// NOT restated: an integer part that '_' separators make longer than 19 characters while it holds fewer than 19 digits,
// at the very end of the input — the reference reads past the end of the string there (undefined); NEEDS_MORE_PRECISION.
// `__host__ __device__`: pinned on the CPU against the compiled reference's TRY_CAST(s AS DOUBLE) (tests/test_cast_parse.py).
#pragma once
#include "cast_decimal_shift_table.h"
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
				if (p == pend) {
					// '_' separators made the integer part look longer than 19 digits and the input ends there: the
					// reference steps over a character that is not there and reads on (fast_float.h:700) — not restated
					out.reference_reads_past_end = true;
					return true;
				}
				// (the reference steps over this character whatever it is: the '.', or the 'e' / space after an integer part
				// that '_' separators made look long — restated as it is)
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

// ---- the fallback: "simple decimal conversion" (fast_float.h:1921-2380, parse_long_mantissa) -------------------------------
// Up to 768 significant decimal digits held one per byte and shifted by powers of two until the value lies in [1/2, 1);
// 53 more doublings leave the mantissa in front of the decimal point.  Exact however long the input is (digits past 768
// only set `truncated`, which breaks ties upwards).  Runs for the handful of inputs Eisel-Lemire cannot decide.
static const uint16_t h_shift_table[65] = B2A_SHIFT_TABLE;
static const uint8_t h_shift_pow5[0x051C] = B2A_SHIFT_POW5_DIGITS;
#if defined(__CUDACC__)
static __device__ const uint16_t d_shift_table[65] = B2A_SHIFT_TABLE;
static __device__ const uint8_t d_shift_pow5[0x051C] = B2A_SHIFT_POW5_DIGITS;
#endif
CASTD_HD inline uint32_t shift_table(uint32_t i) {
#if defined(__CUDA_ARCH__)
	return d_shift_table[i];
#else
	return h_shift_table[i];
#endif
}
CASTD_HD inline uint8_t shift_pow5(uint32_t i) {
#if defined(__CUDA_ARCH__)
	return d_shift_pow5[i];
#else
	return h_shift_pow5[i];
#endif
}

static constexpr uint32_t kMaxDigits = 768;
static constexpr int32_t kDecimalPointRange = 2047;

struct BigDecimal {
	uint32_t num_digits = 0;
	int32_t decimal_point = 0; // the value is 0.d0 d1 d2 ... x 10^decimal_point
	bool truncated = false;
	uint8_t digits[kMaxDigits];
};

// parse_decimal: the digits of an input parse_number accepted.  The reference's version knows nothing of '_' separators:
// it stops at the first one, and so does this (the value is then that of the digits before it).
CASTD_HD inline void parse_big_decimal(const char *p, const char *pend, BigDecimal &d) {
	if (*p == '-') {
		++p;
	}
	while (p != pend && *p == '0') {
		++p;
	}
	while (p != pend && is_digit(*p)) {
		if (d.num_digits < kMaxDigits) {
			d.digits[d.num_digits] = static_cast<uint8_t>(*p - '0');
		}
		d.num_digits++;
		++p;
	}
	if (p != pend && *p == '.') {
		++p;
		const char *const first_after_period = p;
		if (d.num_digits == 0) {
			while (p != pend && *p == '0') {
				++p;
			}
		}
		while (p != pend && is_digit(*p)) {
			if (d.num_digits < kMaxDigits) {
				d.digits[d.num_digits] = static_cast<uint8_t>(*p - '0');
			}
			d.num_digits++;
			++p;
		}
		d.decimal_point = static_cast<int32_t>(first_after_period - p);
	}
	if (d.num_digits > 0) { // trailing zeros are not significant digits
		const char *back = p - 1;
		int32_t trailing_zeros = 0;
		while (*back == '0' || *back == '.') {
			if (*back == '0') {
				trailing_zeros++;
			}
			--back;
		}
		d.decimal_point += static_cast<int32_t>(d.num_digits);
		d.num_digits -= static_cast<uint32_t>(trailing_zeros);
	}
	if (d.num_digits > kMaxDigits) {
		d.truncated = true;
		d.num_digits = kMaxDigits;
	}
	if (p != pend && (*p == 'e' || *p == 'E')) {
		++p;
		bool neg_exp = false;
		if (p != pend && *p == '-') {
			neg_exp = true;
			++p;
		} else if (p != pend && *p == '+') {
			++p;
		}
		int32_t exp_number = 0;
		while (p != pend && is_digit(*p)) {
			if (exp_number < 0x10000) {
				exp_number = 10 * exp_number + (*p - '0');
			}
			++p;
		}
		d.decimal_point += neg_exp ? -exp_number : exp_number;
	}
	for (uint32_t i = d.num_digits; i < 19; i++) {
		d.digits[i] = 0;
	}
}

CASTD_HD inline void trim(BigDecimal &h) {
	while (h.num_digits > 0 && h.digits[h.num_digits - 1] == 0) {
		h.num_digits--;
	}
}

// how many digits h * 2^shift has more than h: len(2^shift), or one fewer when h's leading digits are below 5^shift's
CASTD_HD inline uint32_t new_digits_of_left_shift(const BigDecimal &h, uint32_t shift) {
	shift &= 63;
	const uint32_t x_a = shift_table(shift), x_b = shift_table(shift + 1);
	const uint32_t num_new_digits = x_a >> 11;
	const uint32_t pow5_a = x_a & 0x7FF, pow5_b = x_b & 0x7FF;
	const uint32_t n = pow5_b - pow5_a;
	for (uint32_t i = 0; i < n; i++) {
		if (i >= h.num_digits) {
			return num_new_digits - 1;
		}
		const uint8_t five = shift_pow5(pow5_a + i);
		if (h.digits[i] == five) {
			continue;
		}
		return h.digits[i] < five ? num_new_digits - 1 : num_new_digits;
	}
	return num_new_digits;
}

// h *= 2^shift
CASTD_HD inline void big_left_shift(BigDecimal &h, uint32_t shift) {
	if (h.num_digits == 0) {
		return;
	}
	const uint32_t num_new_digits = new_digits_of_left_shift(h, shift);
	int32_t read_index = static_cast<int32_t>(h.num_digits - 1);
	uint32_t write_index = h.num_digits - 1 + num_new_digits;
	uint64_t n = 0;
	while (read_index >= 0) {
		n += static_cast<uint64_t>(h.digits[read_index]) << shift;
		const uint64_t quotient = n / 10, remainder = n - 10 * quotient;
		if (write_index < kMaxDigits) {
			h.digits[write_index] = static_cast<uint8_t>(remainder);
		} else if (remainder > 0) {
			h.truncated = true;
		}
		n = quotient;
		write_index--;
		read_index--;
	}
	while (n > 0) {
		const uint64_t quotient = n / 10, remainder = n - 10 * quotient;
		if (write_index < kMaxDigits) {
			h.digits[write_index] = static_cast<uint8_t>(remainder);
		} else if (remainder > 0) {
			h.truncated = true;
		}
		n = quotient;
		write_index--;
	}
	h.num_digits += num_new_digits;
	if (h.num_digits > kMaxDigits) {
		h.num_digits = kMaxDigits;
	}
	h.decimal_point += static_cast<int32_t>(num_new_digits);
	trim(h);
}

// h /= 2^shift
CASTD_HD inline void big_right_shift(BigDecimal &h, uint32_t shift) {
	uint32_t read_index = 0, write_index = 0;
	uint64_t n = 0;
	while ((n >> shift) == 0) {
		if (read_index < h.num_digits) {
			n = 10 * n + h.digits[read_index++];
		} else if (n == 0) {
			return;
		} else {
			while ((n >> shift) == 0) {
				n = 10 * n;
				read_index++;
			}
			break;
		}
	}
	h.decimal_point -= static_cast<int32_t>(read_index - 1);
	if (h.decimal_point < -kDecimalPointRange) { // zero
		h.num_digits = 0;
		h.decimal_point = 0;
		h.truncated = false;
		return;
	}
	const uint64_t mask = (uint64_t(1) << shift) - 1;
	while (read_index < h.num_digits) {
		const uint8_t new_digit = static_cast<uint8_t>(n >> shift);
		n = 10 * (n & mask) + h.digits[read_index++];
		h.digits[write_index++] = new_digit;
	}
	while (n > 0) {
		const uint8_t new_digit = static_cast<uint8_t>(n >> shift);
		n = 10 * (n & mask);
		if (write_index < kMaxDigits) {
			h.digits[write_index++] = new_digit;
		} else if (new_digit > 0) {
			h.truncated = true;
		}
	}
	h.num_digits = write_index;
	trim(h);
}

// the integer in front of the decimal point, ties to even (`truncated` counts as "more digits follow")
CASTD_HD inline uint64_t big_round(const BigDecimal &h) {
	if (h.num_digits == 0 || h.decimal_point < 0) {
		return 0;
	}
	if (h.decimal_point > 18) {
		return UINT64_MAX;
	}
	const uint32_t dp = static_cast<uint32_t>(h.decimal_point);
	uint64_t n = 0;
	for (uint32_t i = 0; i < dp; i++) {
		n = 10 * n + (i < h.num_digits ? h.digits[i] : 0);
	}
	bool round_up = false;
	if (dp < h.num_digits) {
		round_up = h.digits[dp] >= 5;
		if (h.digits[dp] == 5 && dp + 1 == h.num_digits) {
			round_up = h.truncated || (dp > 0 && (h.digits[dp - 1] & 1));
		}
	}
	return round_up ? n + 1 : n;
}

CASTD_HD inline uint32_t shift_for_power(uint32_t n) { // the largest shift that cannot overflow 10^n digits: floor(n * log2(10))
	const uint8_t powers[19] = {0, 3, 6, 9, 13, 16, 19, 23, 26, 29, 33, 36, 39, 43, 46, 49, 53, 56, 59};
	return n < 19 ? powers[n] : 60;
}

// compute_float(decimal &), binary64
CASTD_HD inline Adjusted big_decimal_to_double(BigDecimal &d) {
	Adjusted a {0, 0};
	if (d.num_digits == 0 || d.decimal_point < -324) {
		return a;
	}
	if (d.decimal_point >= 310) {
		a.power2 = 0x7FF;
		return a;
	}
	int32_t exp2 = 0;
	while (d.decimal_point > 0) {
		const uint32_t shift = shift_for_power(static_cast<uint32_t>(d.decimal_point));
		big_right_shift(d, shift);
		if (d.decimal_point < -kDecimalPointRange) {
			return a;
		}
		exp2 += static_cast<int32_t>(shift);
	}
	while (d.decimal_point <= 0) { // up into [1/2, 1)
		uint32_t shift;
		if (d.decimal_point == 0) {
			if (d.digits[0] >= 5) {
				break;
			}
			shift = d.digits[0] < 2 ? 2 : 1;
		} else {
			shift = shift_for_power(static_cast<uint32_t>(-d.decimal_point));
		}
		big_left_shift(d, shift);
		if (d.decimal_point > kDecimalPointRange) {
			a.power2 = 0x7FF;
			return a;
		}
		exp2 -= static_cast<int32_t>(shift);
	}
	exp2--; // [1/2, 1) -> [1, 2)
	const int32_t minimum_exponent = -1023;
	while (minimum_exponent + 1 > exp2) { // subnormal: give up mantissa bits
		uint32_t n = static_cast<uint32_t>((minimum_exponent + 1) - exp2);
		if (n > 60) {
			n = 60;
		}
		big_right_shift(d, n);
		exp2 += static_cast<int32_t>(n);
	}
	if (exp2 - minimum_exponent >= 0x7FF) {
		a.power2 = 0x7FF;
		return a;
	}
	big_left_shift(d, 53);
	uint64_t mantissa = big_round(d);
	if (mantissa >= (uint64_t(1) << 53)) { // rounded up to the next power of two
		big_right_shift(d, 1);
		exp2 += 1;
		mantissa = big_round(d);
		if (exp2 - minimum_exponent >= 0x7FF) {
			a.power2 = 0x7FF;
			a.mantissa = 0;
			return a;
		}
	}
	a.power2 = exp2 - minimum_exponent;
	if (mantissa < (uint64_t(1) << 52)) {
		a.power2--;
	}
	a.mantissa = mantissa & ((uint64_t(1) << 52) - 1);
	return a;
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
	if (am.power2 < 0) { // cannot be decided in 128 bits: every digit, exactly (the reference's parse_long_mantissa)
		BigDecimal d;
		parse_big_decimal(buf, end, d);
		am = big_decimal_to_double(d);
	}
	*out = to_double(pns.negative, am);
	return OK;
}

#undef CASTD_SKIP_UNDERSCORE

} // namespace castd
} // namespace b2a
