// VARCHAR -> DECIMAL(w,s) as DuckDB casts it: TryDecimalStringCast = the integer cast loop (IntegerCastLoop with exponents,
// signed, non-strict — the control flow restated in cast_parse.cuh) over DecimalCastOperation
// (duckdb/src/include/duckdb/common/operator/decimal_cast_operators.hpp:483-660).  The reference works in the DECIMAL's
// physical type and two of its limits depend on it — how many digits a value may collect before further fraction digits are
// dropped (4 / 9 / 18 / 38) and the overflow guard of HandleDigit — so the restatement is a template over that type too.
//   digits in front of the point: at most width - scale of them (leading zeros are free), else the cast fails
//   fraction digits: kept (all of them, so that an exponent can still use them) until the type's digit capacity is reached;
//                    the first digit past `scale` decides the rounding when there is no exponent
//   exponent: shifts the digits; what falls off on the right rounds half away from zero
//   the result must stay below 10^width in magnitude
// `__host__ __device__`: pinned on the CPU against TRY_CAST of the compiled reference (tests/test_cast_parse.py).
#pragma once
#include "cast_parse.cuh"

namespace b2a {
namespace castq {

template <class T>
struct Store; // the physical types of DECIMAL
template <>
struct Store<int16_t> {
	static constexpr int max_digits = 4;
	CASTP_HD static int16_t max() {
		return INT16_MAX;
	}
	CASTP_HD static int16_t min() {
		return INT16_MIN;
	}
};
template <>
struct Store<int32_t> {
	static constexpr int max_digits = 9;
	CASTP_HD static int32_t max() {
		return INT32_MAX;
	}
	CASTP_HD static int32_t min() {
		return INT32_MIN;
	}
};
template <>
struct Store<int64_t> {
	static constexpr int max_digits = 18;
	CASTP_HD static int64_t max() {
		return INT64_MAX;
	}
	CASTP_HD static int64_t min() {
		return INT64_MIN;
	}
};
template <>
struct Store<__int128> {
	static constexpr int max_digits = 38;
	CASTP_HD static __int128 max() {
		return static_cast<__int128>((~static_cast<unsigned __int128>(0)) >> 1);
	}
	CASTP_HD static __int128 min() {
		return -max() - 1;
	}
};

enum Exponent : uint8_t { NONE = 0, POSITIVE = 1, NEGATIVE_EXP = 2 };

template <class T>
struct State { // DecimalCastData<T>
	T result = 0;
	uint8_t width = 0, scale = 0;
	uint8_t digit_count = 0, decimal_count = 0;
	bool round_set = false, should_round = false;
	uint8_t excessive_decimals = 0;
	uint8_t exponent_type = NONE;
	T limit = 0;
};

template <class T, bool NEG>
CASTP_HD inline bool handle_digit(State<T> &st, uint8_t digit) {
	if (st.result == 0 && digit == 0) {
		return true;
	}
	if (st.digit_count == st.width - st.scale) {
		return false;
	}
	st.digit_count++;
	if (NEG) {
		if (st.result < Store<T>::min() / 10) {
			return false;
		}
		st.result = static_cast<T>(st.result * 10 - digit);
	} else {
		if (st.result > Store<T>::max() / 10) {
			return false;
		}
		st.result = static_cast<T>(st.result * 10 + digit);
	}
	return true;
}

template <class T, bool NEG>
CASTP_HD inline void round_up(State<T> &st) {
	st.result = static_cast<T>(NEG ? st.result - 1 : st.result + 1);
}

template <class T, bool NEG>
CASTP_HD inline bool handle_decimal(State<T> &st, uint8_t digit) { // ALLOW_EXPONENT = true
	if (st.decimal_count == st.scale && !st.round_set) {
		st.round_set = true;
		st.should_round = digit >= 5;
	}
	if (st.digit_count + st.decimal_count >= Store<T>::max_digits) {
		return true;
	}
	st.decimal_count++;
	st.result = static_cast<T>(NEG ? st.result * 10 - digit : st.result * 10 + digit);
	return true;
}

template <class T, bool NEG>
CASTP_HD inline bool finalize(State<T> &st) {
	if (st.exponent_type != POSITIVE && st.decimal_count > st.scale) {
		st.excessive_decimals = static_cast<uint8_t>(st.decimal_count - st.scale);
	}
	if (st.excessive_decimals) { // TruncateExcessiveDecimals
		bool up = false;
		for (uint32_t i = 0; i < st.excessive_decimals; i++) {
			const T mod = static_cast<T>(st.result % 10);
			up = NEG ? mod <= -5 : mod >= 5;
			st.result = static_cast<T>(st.result / 10);
		}
		if (st.exponent_type == POSITIVE && up) {
			round_up<T, NEG>(st);
		}
		st.decimal_count = st.scale;
	}
	if (st.exponent_type == NONE && st.round_set && st.should_round) {
		round_up<T, NEG>(st);
	}
	for (uint8_t i = st.decimal_count; i < st.scale; i++) {
		st.result = static_cast<T>(st.result * 10);
	}
	return NEG ? st.result > -st.limit : st.result < st.limit;
}

template <class T, bool NEG>
CASTP_HD inline bool handle_exponent(State<T> &st, int32_t exponent) {
	const int32_t decimal_excess = st.decimal_count > st.scale ? st.decimal_count - st.scale : 0;
	if (exponent > 0) {
		st.exponent_type = POSITIVE;
		if (decimal_excess > exponent) {
			st.excessive_decimals = static_cast<uint8_t>(decimal_excess - exponent);
			exponent = 0;
		} else {
			exponent -= decimal_excess;
		}
	} else if (exponent < 0) {
		st.exponent_type = NEGATIVE_EXP;
	}
	if (!finalize<T, NEG>(st)) {
		return false;
	}
	if (exponent < 0) {
		bool up = false;
		for (int64_t i = 0; i < -static_cast<int64_t>(exponent); i++) {
			const T mod = static_cast<T>(st.result % 10);
			up = NEG ? mod <= -5 : mod >= 5;
			st.result = static_cast<T>(st.result / 10);
			if (st.result == 0) {
				break;
			}
		}
		if (up) {
			round_up<T, NEG>(st);
		}
		return true;
	}
	for (int32_t i = 0; i < exponent; i++) {
		if (!handle_digit<T, NEG>(st, 0)) {
			return false;
		}
	}
	return true;
}

// IntegerCastLoop<DecimalCastData<T>, NEGATIVE, ALLOW_EXPONENT = true, DecimalCastOperation>, non-strict
template <class T, bool NEG>
CASTP_HD inline bool loop(const char *buf, size_t len, State<T> &st) {
	using castp::is_digit;
	using castp::is_space;
	const size_t start_pos = NEG ? 1 : (buf[0] == '+' ? 1 : 0);
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
					if (!handle_decimal<T, NEG>(st, static_cast<uint8_t>(buf[pos] - '0'))) {
						return false;
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
			if (buf[pos] == 'e' || buf[pos] == 'E') {
				if (pos == start_pos) {
					return false;
				}
				pos++;
				if (pos >= len) {
					return false;
				}
				int16_t exponent = 0;
				const bool ok = buf[pos] == '-' ? castp::exponent_loop<true>(buf + pos, len - pos, exponent)
				                                : castp::exponent_loop<false>(buf + pos, len - pos, exponent);
				if (!ok) {
					return false;
				}
				return handle_exponent<T, NEG>(st, exponent);
			}
			return false;
		}
		if (!handle_digit<T, NEG>(st, static_cast<uint8_t>(buf[pos++] - '0'))) {
			return false;
		}
		if (pos != len && buf[pos] == '_') {
			pos++;
			if (pos == len || !is_digit(buf[pos])) {
				return false;
			}
		}
	}
	if (!finalize<T, NEG>(st)) {
		return false;
	}
	return pos > start_pos;
}

// TryDecimalStringCast<T>: false = "Could not convert string ... to DECIMAL(w,s)"
template <class T>
CASTP_HD inline bool varchar_to_decimal_as(const char *buf, size_t len, uint32_t width, uint32_t scale, __int128 *out) {
	while (len > 0 && castp::is_space(*buf)) {
		buf++;
		len--;
	}
	if (len == 0) {
		return false;
	}
	State<T> st;
	st.width = static_cast<uint8_t>(width);
	st.scale = static_cast<uint8_t>(scale);
	T limit = 1;
	for (uint32_t i = 0; i < width; i++) {
		limit = static_cast<T>(limit * 10);
	}
	st.limit = limit;
	bool ok;
	if (*buf == '-') {
		ok = loop<T, true>(buf, len, st);
	} else if (len > 1 && *buf == '0' && (buf[1] == 'x' || buf[1] == 'X' || buf[1] == 'b' || buf[1] == 'B')) {
		ok = false; // the hex / binary loops of the integer cast: DecimalCastOperation accepts no such digit
	} else {
		ok = loop<T, false>(buf, len, st);
	}
	*out = st.result;
	return ok;
}

CASTP_HD inline bool varchar_to_decimal(const char *buf, size_t len, uint32_t width, uint32_t scale, __int128 *out) {
	if (width <= 4) {
		return varchar_to_decimal_as<int16_t>(buf, len, width, scale, out);
	}
	if (width <= 9) {
		return varchar_to_decimal_as<int32_t>(buf, len, width, scale, out);
	}
	if (width <= 18) {
		return varchar_to_decimal_as<int64_t>(buf, len, width, scale, out);
	}
	return varchar_to_decimal_as<__int128>(buf, len, width, scale, out);
}

} // namespace castq
} // namespace b2a
