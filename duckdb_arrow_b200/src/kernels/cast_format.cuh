// BOOLEAN / DATE / TIME / TIMESTAMP / DECIMAL / HUGEINT -> VARCHAR as DuckDB casts them (StringCast::Operation,
// duckdb/src/common/operator/string_cast.cpp:16-23, 86-187; DateToStringCast / TimeToStringCast / DecimalToString,
// duckdb/src/include/duckdb/common/types/cast_helpers.hpp:109-171, 194-329; cast_helpers.cpp:212-315):
//   BOOLEAN    "true" / "false"
//   DATE       YYYY-MM-DD, the year padded to four digits, years <= 0 as (1 - year) with " (BC)" appended; +-infinity
//   TIME       HH:MM:SS[.ffffff], trailing zeros of the microseconds dropped
//   TIMESTAMP  the date (with its " (BC)"), a space, the time; +-infinity
//   DECIMAL    [-]major.minor with `scale` minor digits; no major digits at all when width == scale (".500"); scale 0 = integer
//   HUGEINT    the integer
// The calendar arithmetic is not the reference's table walk (Date::Convert) but the closed form over 400-year eras — the
// same proleptic Gregorian calendar, pinned over the whole int32 range.
// `__host__ __device__`: pinned on the CPU against CAST(x AS VARCHAR) of the compiled reference (tests/test_cast_parse.py).
#pragma once
#include <stddef.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define CASTF_HD __host__ __device__
#else
#define CASTF_HD
#endif

namespace b2a {
namespace castf {

static constexpr uint32_t kMaxChars = 48; // the longest: DECIMAL(38, s) = 40, TIMESTAMP = 33

CASTF_HD inline uint32_t put(char *out, const char *s, uint32_t n) {
	for (uint32_t i = 0; i < n; i++) {
		out[i] = s[i];
	}
	return n;
}

CASTF_HD inline uint32_t format_bool(bool v, char *out) {
	return v ? put(out, "true", 4) : put(out, "false", 5);
}

CASTF_HD inline void two_digits(char *out, int32_t v) {
	out[0] = static_cast<char>('0' + v / 10);
	out[1] = static_cast<char>('0' + v % 10);
}

// days since 1970-01-01 -> proleptic Gregorian year / month / day (year 0 = 1 BC)
CASTF_HD inline void civil_from_days(int64_t days, int64_t &year, int32_t &month, int32_t &day) {
	const int64_t z = days + 719468; // days since 0000-03-01
	const int64_t era = (z >= 0 ? z : z - 146096) / 146097;
	const int64_t doe = z - era * 146097;                                       // [0, 146096]
	const int64_t yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365; // [0, 399]
	const int64_t doy = doe - (365 * yoe + yoe / 4 - yoe / 100);               // [0, 365], March-based
	const int64_t mp = (5 * doy + 2) / 153;                                     // [0, 11]
	day = static_cast<int32_t>(doy - (153 * mp + 2) / 5 + 1);
	month = static_cast<int32_t>(mp < 10 ? mp + 3 : mp - 9);
	year = yoe + era * 400 + (month <= 2 ? 1 : 0);
}

// DateToStringCast::Length + Format
CASTF_HD inline uint32_t format_ymd(int64_t year, int32_t month, int32_t day, char *out) {
	bool add_bc = false;
	if (year <= 0) {
		year = -year + 1;
		add_bc = true;
	}
	uint32_t year_length = 4;
	year_length += year >= 10000;
	year_length += year >= 100000;
	year_length += year >= 1000000;
	year_length += year >= 10000000;
	for (uint32_t i = year_length; i > 0; i--) {
		out[i - 1] = static_cast<char>('0' + year % 10);
		year /= 10;
	}
	uint32_t k = year_length;
	out[k++] = '-';
	two_digits(out + k, month);
	k += 2;
	out[k++] = '-';
	two_digits(out + k, day);
	k += 2;
	if (add_bc) {
		k += put(out + k, " (BC)", 5);
	}
	return k;
}

CASTF_HD inline uint32_t format_date(int32_t days, char *out) {
	if (days == INT32_MAX) {
		return put(out, "infinity", 8);
	}
	if (days == -INT32_MAX) {
		return put(out, "-infinity", 9);
	}
	int64_t year;
	int32_t month, day;
	civil_from_days(days, year, month, day);
	return format_ymd(year, month, day, out);
}

// Time::Convert + TimeToStringCast: microseconds of the day (0 .. 86 400 000 000: 24:00:00 exists)
CASTF_HD inline uint32_t format_time(int64_t micros, char *out) {
	const int32_t hour = static_cast<int32_t>(micros / 3600000000ll);
	micros -= static_cast<int64_t>(hour) * 3600000000ll;
	const int32_t minute = static_cast<int32_t>(micros / 60000000ll);
	micros -= static_cast<int64_t>(minute) * 60000000ll;
	const int32_t second = static_cast<int32_t>(micros / 1000000ll);
	int32_t us = static_cast<int32_t>(micros - static_cast<int64_t>(second) * 1000000ll);
	two_digits(out, hour);
	out[2] = ':';
	two_digits(out + 3, minute);
	out[5] = ':';
	two_digits(out + 6, second);
	if (us == 0) {
		return 8;
	}
	out[8] = '.';
	uint32_t n = 6;
	while (us % 10 == 0) { // "90000" -> ".09"
		us /= 10;
		n--;
	}
	for (uint32_t i = n; i > 0; i--) {
		out[8 + i] = static_cast<char>('0' + us % 10);
		us /= 10;
	}
	return 9 + n;
}

// StringFromTimestamp<false>: Timestamp::Convert = floor division by the microseconds of a day; 0 = ConversionException
CASTF_HD inline uint32_t format_timestamp(int64_t micros, char *out) {
	if (micros == INT64_MAX) {
		return put(out, "infinity", 8);
	}
	if (micros == -INT64_MAX) {
		return put(out, "-infinity", 9);
	}
	const int64_t per_day = 86400000000ll;
	int64_t days = micros / per_day;
	int64_t rest = micros - days * per_day;
	if (rest < 0) {
		days--;
		rest += per_day;
	}
	if (days < -106751991) {
		return 0; // days x 86 400 000 000 leaves int64: "Date out of range in timestamp conversion" (Timestamp::Convert)
	}
	int64_t year;
	int32_t month, day;
	civil_from_days(days, year, month, day);
	uint32_t k = format_ymd(year, month, day, out);
	out[k++] = ' ';
	return k + format_time(rest, out + k);
}

// DecimalToString::FormatDecimal for every storage width (the value arrives sign-extended to 128 bits); scale 0 with
// width 0 formats a plain HUGEINT (NumericHelper::FormatSigned<hugeint_t>)
CASTF_HD inline uint32_t format_decimal(__int128 value, uint32_t width, uint32_t scale, char *out) {
	const bool negative = value < 0;
	unsigned __int128 v = negative ? -static_cast<unsigned __int128>(value) : static_cast<unsigned __int128>(value);
	char digits[40]; // least significant first
	uint32_t n = 0;
	while (v != 0) {
		// 19 decimal digits at a time: one 128-bit division per 19 digits instead of per digit
		const unsigned __int128 q = v / 10000000000000000000ull;
		uint64_t r = static_cast<uint64_t>(v - q * 10000000000000000000ull);
		v = q;
		for (int i = 0; i < 19 && (r != 0 || v != 0); i++) {
			digits[n++] = static_cast<char>('0' + r % 10);
			r /= 10;
		}
	}
	uint32_t k = 0;
	if (negative) {
		out[k++] = '-';
	}
	if (scale == 0) {
		if (n == 0) {
			out[k++] = '0';
		}
		while (n > 0) {
			out[k++] = digits[--n];
		}
		return k;
	}
	if (width > scale) { // the digits in front of the point, at least "0"
		if (n <= scale) {
			out[k++] = '0';
		}
		while (n > scale) {
			out[k++] = digits[--n];
		}
	}
	out[k++] = '.';
	for (uint32_t i = scale; i > 0; i--) {
		out[k++] = i <= n ? digits[i - 1] : '0';
	}
	return k;
}

} // namespace castf
} // namespace b2a
