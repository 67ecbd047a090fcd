// VARCHAR -> BIGINT as DuckDB casts it: TrySimpleIntegerCast<int64_t>, non-strict
// (duckdb/src/include/duckdb/common/operator/integer_cast_operator.hpp:20-458).  Two passes like the reference: the
// plain integer grammar first (digits, '_' between digits, 0x / 0b literals, "1." — no fraction digits, no exponent),
// and when that fails the decimal grammar: fraction digits (the value rounds half away from zero on its FIRST fraction
// digit, with the reference's quirk that a fraction that shrinks to exactly 10 does not round) and an exponent.
// `__host__ __device__` so that the CPU suite pins the very same code against the compiled reference
// (tests/test_cast_parse.py); the product calls it from b2a_cast_kernel only.
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define CASTP_HD __host__ __device__
#else
#define CASTP_HD
#endif

namespace b2a {
namespace castp {

CASTP_HD inline bool is_digit(char c) {
	return c >= '0' && c <= '9';
}
CASTP_HD inline bool is_space(char c) { // StringUtil::CharacterIsSpace
	return c == ' ' || (c >= '\t' && c <= '\r');
}
CASTP_HD inline bool is_hex(char c) {
	return is_digit(c) || (c >= 'a' && c <= 'f') || (c >= 'A' && c <= 'F');
}

static constexpr int64_t kMax = INT64_MAX, kMin = INT64_MIN;

// TryMultiplyOperator / TryAddOperator / TrySubtractOperator on int64
CASTP_HD inline bool try_mul10(int64_t &v) {
	if (v > kMax / 10 || v < kMin / 10) {
		return false;
	}
	v *= 10;
	return true;
}
CASTP_HD inline bool try_add(int64_t a, int64_t b, int64_t &r) {
	if ((b > 0 && a > kMax - b) || (b < 0 && a < kMin - b)) {
		return false;
	}
	r = a + b;
	return true;
}
CASTP_HD inline bool try_sub(int64_t a, int64_t b, int64_t &r) {
	if ((b < 0 && a > kMax + b) || (b > 0 && a < kMin + b)) {
		return false;
	}
	r = a - b;
	return true;
}

// IntegerCastData<int64_t> / IntegerDecimalCastData<int64_t>
struct State {
	int64_t result = 0;
	int64_t decimal = 0;
	uint16_t decimal_digits = 0;
	bool overflow = false; // an arithmetic check failed somewhere (the caller reports "out of range" instead of "not a number")
};

// IntegerCastOperation::HandleDigit
template <bool NEGATIVE>
CASTP_HD inline bool handle_digit(State &st, uint8_t digit) {
	if (NEGATIVE) {
		if (st.result < (kMin + digit) / 10) {
			st.overflow = true;
			return false;
		}
		st.result = st.result * 10 - digit;
	} else {
		if (st.result > (kMax - digit) / 10) {
			st.overflow = true;
			return false;
		}
		st.result = st.result * 10 + digit;
	}
	return true;
}

// IntegerDecimalCastOperation::Finalize (the plain operation's Finalize does nothing)
template <bool NEGATIVE>
CASTP_HD inline bool finalize_decimal(State &st) {
	int64_t tmp = st.result;
	while (st.decimal > 10) {
		st.decimal /= 10;
		st.decimal_digits--;
	}
	bool ok = true;
	if (st.decimal_digits == 1 && st.decimal >= 5) {
		ok = NEGATIVE ? try_sub(tmp, 1, tmp) : try_add(tmp, 1, tmp);
		st.overflow = st.overflow || !ok;
	}
	st.result = tmp;
	return ok;
}

// IntegerDecimalCastOperation::HandleExponent
template <bool NEGATIVE>
CASTP_HD inline bool handle_exponent(State &st, int16_t exponent) {
	int16_t e = exponent;
	if (e < 0) {
		while (st.result != 0 && e++ < 0) {
			st.decimal = st.result % 10;
			st.result /= 10;
		}
		if (st.decimal < 0) {
			st.decimal = -st.decimal;
		}
		st.decimal_digits = 1;
		return finalize_decimal<NEGATIVE>(st);
	}
	while (st.result != 0 && e-- > 0) {
		if (!try_mul10(st.result)) {
			st.overflow = true;
			return false;
		}
	}
	if (st.decimal == 0) {
		return finalize_decimal<NEGATIVE>(st);
	}
	e = static_cast<int16_t>(exponent - st.decimal_digits);
	int64_t remainder = 0;
	if (e < 0) {
		if (static_cast<uint16_t>(-e) <= 19) { // NumericLimits<int64_t>::Digits()
			int64_t power = 1;
			while (e++ < 0) {
				power *= 10;
			}
			remainder = st.decimal % power;
			st.decimal /= power;
		} else {
			st.decimal = 0;
		}
	} else {
		while (e-- > 0) {
			if (!try_mul10(st.decimal)) {
				st.overflow = true;
				return false;
			}
		}
	}
	st.decimal_digits = static_cast<uint16_t>(st.decimal_digits - exponent);
	const bool ok = NEGATIVE ? try_sub(st.result, st.decimal, st.result) : try_add(st.result, st.decimal, st.result);
	if (!ok) {
		st.overflow = true;
		return false;
	}
	st.decimal = remainder;
	return finalize_decimal<NEGATIVE>(st);
}

// the exponent itself: IntegerCastLoop<IntegerCastData<int16_t>, NEGATIVE, false, IntegerCastOperation>
template <bool NEGATIVE>
CASTP_HD inline bool exponent_loop(const char *buf, size_t len, int16_t &out) {
	const size_t start_pos = NEGATIVE ? 1 : (buf[0] == '+' ? 1 : 0);
	size_t pos = start_pos;
	int32_t r = 0;
	while (pos < len) {
		if (!is_digit(buf[pos])) {
			if (buf[pos] == '.') {
				const bool number_before_period = pos > start_pos;
				pos++;
				if (pos < len && is_digit(buf[pos])) {
					return false; // HandleDecimal of the plain operation
				}
				if (!number_before_period) {
					return false;
				}
				if (pos >= len) {
					break;
				}
			}
			if (is_space(buf[pos])) {
				while (++pos < len) {
					if (!is_space(buf[pos])) {
						return false;
					}
				}
				break;
			}
			return false;
		}
		const int32_t digit = buf[pos++] - '0';
		if (NEGATIVE) {
			if (r < (INT16_MIN + digit) / 10) {
				return false;
			}
			r = r * 10 - digit;
		} else {
			if (r > (INT16_MAX - digit) / 10) {
				return false;
			}
			r = r * 10 + digit;
		}
		if (pos != len && buf[pos] == '_') {
			pos++;
			if (pos == len || !is_digit(buf[pos])) {
				return false;
			}
		}
	}
	out = static_cast<int16_t>(r);
	return pos > start_pos;
}

// IntegerCastLoop<T, NEGATIVE, ALLOW_EXPONENT = DECIMAL, OP = DECIMAL ? IntegerDecimalCastOperation : IntegerCastOperation>
template <bool NEGATIVE, bool DECIMAL>
CASTP_HD inline bool integer_loop(const char *buf, size_t len, State &st) {
	const size_t start_pos = NEGATIVE ? 1 : (buf[0] == '+' ? 1 : 0);
	size_t pos = start_pos;
	while (pos < len) {
		if (!is_digit(buf[pos])) {
			if (buf[pos] == '.') {
				const bool number_before_period = pos > start_pos;
				pos++;
				const size_t start_digit = pos;
				while (pos < len) {
					if (!is_digit(buf[pos])) {
						break;
					}
					if (!DECIMAL) {
						return false; // the plain operation has no HandleDecimal
					}
					const uint8_t digit = static_cast<uint8_t>(buf[pos] - '0');
					if (!(st.decimal > (kMax - digit) / 10)) { // (more fraction digits than fit are ignored)
						st.decimal_digits++;
						st.decimal = st.decimal * 10 + digit;
					}
					pos++;
					if (pos != len && buf[pos] == '_') {
						pos++;
						if (pos == len || !is_digit(buf[pos])) {
							return false;
						}
					}
				}
				if (!(number_before_period || pos > start_digit)) {
					return false; // "1." and ".1" are numbers, "." is not
				}
				if (pos >= len) {
					break;
				}
			}
			if (is_space(buf[pos])) {
				while (++pos < len) {
					if (!is_space(buf[pos])) {
						return false;
					}
				}
				break;
			}
			if (DECIMAL && (buf[pos] == 'e' || buf[pos] == 'E')) {
				if (pos == start_pos) {
					return false;
				}
				pos++;
				if (pos >= len) {
					return false;
				}
				int16_t exponent = 0;
				const bool ok = buf[pos] == '-' ? exponent_loop<true>(buf + pos, len - pos, exponent)
				                                : exponent_loop<false>(buf + pos, len - pos, exponent);
				if (!ok) {
					return false;
				}
				return handle_exponent<NEGATIVE>(st, exponent);
			}
			return false;
		}
		if (!handle_digit<NEGATIVE>(st, static_cast<uint8_t>(buf[pos++] - '0'))) {
			return false;
		}
		if (pos != len && buf[pos] == '_') {
			pos++;
			if (pos == len || !is_digit(buf[pos])) {
				return false;
			}
		}
	}
	if (DECIMAL && !finalize_decimal<NEGATIVE>(st)) {
		return false;
	}
	return pos > start_pos;
}

// IntegerHexCastLoop / IntegerBinaryCastLoop (buf[0] is the 'x' / 'b')
template <bool DECIMAL>
CASTP_HD inline bool radix_loop(const char *buf, size_t len, State &st, int radix) {
	size_t pos = 1;
	while (pos < len) {
		char c = buf[pos];
		if (c >= 'A' && c <= 'Z') {
			c = static_cast<char>(c - 'A' + 'a');
		}
		const bool ok_char = radix == 16 ? is_hex(c) : (c == '0' || c == '1');
		if (!ok_char) {
			return false;
		}
		const int64_t digit = c >= 'a' ? c - 'a' + 10 : c - '0';
		pos++;
		if (pos != len && buf[pos] == '_') {
			pos++;
			const bool next_ok = pos != len && (radix == 16 ? is_hex(buf[pos]) : (buf[pos] == '0' || buf[pos] == '1'));
			if (!next_ok) {
				return false;
			}
		}
		if (st.result > (kMax - digit) / radix) {
			st.overflow = true;
			return false;
		}
		st.result = st.result * radix + digit;
	}
	if (DECIMAL && !finalize_decimal<false>(st)) {
		return false;
	}
	return pos > 1;
}

// TryIntegerCast<.., IS_SIGNED = true, ALLOW_EXPONENT = DECIMAL, ..>, non-strict
template <bool DECIMAL>
CASTP_HD inline bool try_integer_cast(const char *buf, size_t len, State &st) {
	while (len > 0 && is_space(*buf)) {
		buf++;
		len--;
	}
	if (len == 0) {
		return false;
	}
	st = State();
	if (*buf == '-') {
		return integer_loop<true, DECIMAL>(buf, len, st);
	}
	if (len > 1 && *buf == '0') {
		if (buf[1] == 'x' || buf[1] == 'X') {
			return radix_loop<DECIMAL>(buf + 1, len - 1, st, 16);
		}
		if (buf[1] == 'b' || buf[1] == 'B') {
			return radix_loop<DECIMAL>(buf + 1, len - 1, st, 2);
		}
	}
	return integer_loop<false, DECIMAL>(buf, len, st);
}

enum Result : uint32_t { OK = 0, OUT_OF_RANGE = 1, NOT_A_NUMBER = 2 };

// TrySimpleIntegerCast<int64_t>: the plain grammar, then the decimal / exponent grammar
CASTP_HD inline uint32_t varchar_to_bigint(const char *s, size_t n, int64_t *out) {
	State st;
	if (try_integer_cast<false>(s, n, st)) {
		*out = st.result;
		return OK;
	}
	const bool overflow_first = st.overflow;
	if (try_integer_cast<true>(s, n, st)) {
		*out = st.result;
		return OK;
	}
	return (st.overflow || overflow_first) ? OUT_OF_RANGE : NOT_A_NUMBER;
}

} // namespace castp
} // namespace b2a
