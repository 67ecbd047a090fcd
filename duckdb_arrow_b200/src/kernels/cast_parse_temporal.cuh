// VARCHAR -> DATE / TIMESTAMP as DuckDB casts them, non-strict (what CAST and the multi-file mapper use):
//   Date::TryConvertDate          duckdb/src/common/types/date.cpp:182-331 — [-]Y..Y sep M[M] sep D[D] [" (BC)"], sep one of
//                                 '-' '/' '\' ' ' (the same twice), "infinity" / "-infinity" / "epoch", trailing text allowed
//                                 unless it starts with a digit; Date::IsValid's range -5877641-06-25 .. 5881580-07-10
//   Timestamp::TryConvertTimestamp   duckdb/src/common/types/timestamp.cpp:64-190, 255-325 — the date, ' ' or 'T', a time
//                                 H..H:M[M][:S[S][.f..f]] (Time::TryConvertInternal, time.cpp:17-127: up to nine hour digits,
//                                 at most one day in total), then "Z", a UTC offset +-HH[[:]MM[:SS]] or " UTC"; any other time
//                                 zone name is the reference's "timestamp with time zone needs ICU" conversion error
// The day number is the closed form over 400-year eras, not the reference's table walk — same calendar.
// `__host__ __device__`: pinned on the CPU against TRY_CAST of the compiled reference (tests/test_cast_parse.py).
#pragma once
#include <stddef.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define CASTT_HD __host__ __device__
#else
#define CASTT_HD
#endif

namespace b2a {
namespace castt {

enum Result : uint32_t { OK = 0, OUT_OF_RANGE = 1, BAD_FORMAT = 2 };

CASTT_HD inline bool is_digit(char c) {
	return c >= '0' && c <= '9';
}
CASTT_HD inline bool is_space(char c) {
	return c == ' ' || (c >= '\t' && c <= '\r');
}
CASTT_HD inline char lower(char c) {
	return (c >= 'A' && c <= 'Z') ? static_cast<char>(c - 'A' + 'a') : c;
}
CASTT_HD inline bool is_alpha(char c) {
	return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z');
}

// Date::ParseDoubleDigit
CASTT_HD inline bool two_digits(const char *buf, size_t len, size_t &pos, int32_t &out) {
	if (pos < len && is_digit(buf[pos])) {
		out = buf[pos++] - '0';
		if (pos < len && is_digit(buf[pos])) {
			out = (buf[pos++] - '0') + out * 10;
		}
		return true;
	}
	return false;
}

// Date::TryConvertDateSpecial
CASTT_HD inline bool special_word(const char *buf, size_t len, size_t &pos, const char *word, size_t word_len) {
	if (len - pos < word_len) {
		return false;
	}
	for (size_t i = 0; i < word_len; i++) {
		if (lower(buf[pos + i]) != word[i]) {
			return false;
		}
	}
	pos += word_len;
	return true;
}

CASTT_HD inline bool is_leap(int32_t year) {
	return year % 4 == 0 && (year % 100 != 0 || year % 400 == 0);
}

// Date::IsValid + Date::TryFromDate
CASTT_HD inline bool days_from_civil(int32_t year, int32_t month, int32_t day, int32_t *out) {
	if (month < 1 || month > 12 || day < 1) {
		return false;
	}
	if (year < -5877641 || (year == -5877641 && (month < 6 || (month == 6 && day < 25)))) {
		return false;
	}
	if (year > 5881580 || (year == 5881580 && (month > 7 || (month == 7 && day > 10)))) {
		return false;
	}
	const int32_t month_days = month == 2 ? (is_leap(year) ? 29 : 28) : (month == 4 || month == 6 || month == 9 || month == 11) ? 30 : 31;
	if (day > month_days) {
		return false;
	}
	const int64_t y = static_cast<int64_t>(year) - (month <= 2 ? 1 : 0);
	const int64_t era = (y >= 0 ? y : y - 399) / 400;
	const int64_t yoe = y - era * 400;
	const int64_t doy = (153 * (month > 2 ? month - 3 : month + 9) + 2) / 5 + day - 1;
	const int64_t doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
	*out = static_cast<int32_t>(era * 146097 + doe - 719468);
	return true;
}

// Date::TryConvertDate(strict = false); `pos` = where the date ended, `special` = infinity / epoch
CASTT_HD inline uint32_t parse_date(const char *buf, size_t len, size_t &pos, int32_t *result, bool &special) {
	special = false;
	pos = 0;
	if (len == 0) {
		return BAD_FORMAT;
	}
	int32_t day = 0, month = -1, year = 0;
	bool yearneg = false;
	while (pos < len && is_space(buf[pos])) {
		pos++;
	}
	if (pos >= len) {
		return BAD_FORMAT;
	}
	if (buf[pos] == '-') {
		yearneg = true;
		pos++;
		if (pos >= len) {
			return BAD_FORMAT;
		}
	}
	if (!is_digit(buf[pos])) {
		if (special_word(buf, len, pos, "infinity", 8)) {
			*result = yearneg ? -INT32_MAX : INT32_MAX;
		} else if (special_word(buf, len, pos, "epoch", 5)) {
			*result = 0;
		} else {
			return BAD_FORMAT;
		}
		while (pos < len && is_space(buf[pos])) {
			pos++;
		}
		special = true;
		return pos == len ? OK : BAD_FORMAT;
	}
	for (; pos < len && is_digit(buf[pos]); pos++) {
		if (year >= 100000000) {
			return OUT_OF_RANGE;
		}
		year = (buf[pos] - '0') + year * 10;
	}
	if (yearneg) {
		year = -year;
	}
	if (pos >= len) {
		return BAD_FORMAT;
	}
	const char sep = buf[pos++];
	if (sep != ' ' && sep != '-' && sep != '/' && sep != '\\') {
		return BAD_FORMAT;
	}
	if (!two_digits(buf, len, pos, month)) {
		return BAD_FORMAT;
	}
	if (pos >= len) {
		return BAD_FORMAT;
	}
	if (buf[pos++] != sep) {
		return BAD_FORMAT;
	}
	if (pos >= len) {
		return BAD_FORMAT;
	}
	if (!two_digits(buf, len, pos, day)) {
		return BAD_FORMAT;
	}
	if (len - pos >= 5 && is_space(buf[pos]) && buf[pos + 1] == '(' && lower(buf[pos + 2]) == 'b' && lower(buf[pos + 3]) == 'c' &&
	    buf[pos + 4] == ')') {
		if (yearneg || year == 0) {
			return BAD_FORMAT;
		}
		year = -year + 1;
		pos += 5;
	}
	if (pos < len && is_digit(buf[pos])) {
		return BAD_FORMAT;
	}
	return days_from_civil(year, month, day, result) ? OK : OUT_OF_RANGE;
}

// TryCast::Operation<string_t, date_t>, non-strict
CASTT_HD inline uint32_t varchar_to_date(const char *buf, size_t len, int32_t *out) {
	size_t pos;
	bool special;
	return parse_date(buf, len, pos, out, special);
}

// Time::TryConvertInternal(strict = false, no nanoseconds): microseconds since midnight (may exceed a day: the caller checks)
CASTT_HD inline bool parse_time(const char *buf, size_t len, size_t &pos, int64_t *out) {
	int32_t hour = 0, min = -1, sec = -1, micros = 0;
	pos = 0;
	if (len == 0) {
		return false;
	}
	while (pos < len && is_space(buf[pos])) {
		pos++;
	}
	if (pos >= len || !is_digit(buf[pos])) {
		return false;
	}
	for (int32_t digits = 9; pos < len && is_digit(buf[pos]); ++pos) {
		if (digits-- > 0) {
			hour = hour * 10 + (buf[pos] - '0');
		} else {
			return false;
		}
	}
	if (pos >= len) {
		return false;
	}
	if (buf[pos++] != ':') {
		return false;
	}
	if (pos == len) {
		min = 0;
	} else {
		if (!two_digits(buf, len, pos, min)) {
			return false;
		}
		if (min < 0 || min >= 60) {
			return false;
		}
	}
	if (pos == len) {
		sec = 0;
	} else {
		if (buf[pos++] != ':') {
			return false;
		}
		if (pos == len) {
			sec = 0;
		} else {
			if (!two_digits(buf, len, pos, sec)) {
				return false;
			}
			if (sec < 0 || sec >= 60) {
				return false;
			}
		}
	}
	if (pos < len && buf[pos] == '.') {
		pos++;
		int32_t mult = 100000;
		for (; pos < len && is_digit(buf[pos]); pos++, mult /= 10) {
			if (mult > 0) {
				micros += (buf[pos] - '0') * mult;
			}
		}
	}
	int64_t r = hour;
	r = r * 60 + min;
	r = r * 60 + sec;
	r = r * 1000000 + micros;
	*out = r;
	return true;
}

// Timestamp::TryParseUTCOffset
CASTT_HD inline bool parse_utc_offset(const char *str, size_t &pos, size_t len, int &hh, int &mm, int &ss) {
	mm = 0;
	ss = 0;
	size_t cur = pos;
	if (cur + 3 > len) {
		return false;
	}
	const char sign = str[cur];
	if (sign != '+' && sign != '-') {
		return false;
	}
	cur++;
	if (!is_digit(str[cur]) || !is_digit(str[cur + 1])) {
		return false;
	}
	hh = (str[cur] - '0') * 10 + (str[cur + 1] - '0');
	if (sign == '-') {
		hh = -hh;
	}
	cur += 2;
	if (cur >= len) {
		pos = cur;
		return true;
	}
	const bool colons = str[cur] == ':';
	if (colons) {
		cur++;
	}
	if (cur + 2 > len || !is_digit(str[cur]) || !is_digit(str[cur + 1])) {
		pos = cur;
		return true;
	}
	mm = (str[cur] - '0') * 10 + (str[cur + 1] - '0');
	if (sign == '-') {
		mm = -mm;
	}
	cur += 2;
	if (cur >= len || !colons || str[cur] != ':') {
		pos = cur;
		return true;
	}
	cur++;
	if (cur + 2 > len || !is_digit(str[cur]) || !is_digit(str[cur + 1])) {
		pos = cur;
		return true;
	}
	ss = (str[cur] - '0') * 10 + (str[cur + 1] - '0');
	if (sign == '-') {
		ss = -ss;
	}
	pos = cur + 2;
	return true;
}

// Timestamp::TryFromDatetime: days x 86 400 000 000 + microseconds without leaving int64, and not one of the infinities
CASTT_HD inline bool timestamp_of(int32_t days, int64_t micros, int64_t *out) {
	const int64_t per_day = 86400000000ll;
	const int64_t d = days;
	if (d > INT64_MAX / per_day || d < INT64_MIN / per_day) {
		return false;
	}
	const int64_t base = d * per_day;
	if ((micros > 0 && base > INT64_MAX - micros) || (micros < 0 && base < INT64_MIN - micros)) {
		return false;
	}
	const int64_t r = base + micros;
	if (r == INT64_MAX || r == -INT64_MAX) {
		return false;
	}
	*out = r;
	return true;
}

// TryCast::Operation<string_t, timestamp_t>: Timestamp::TryConvertTimestamp over TryConvertTimestampTZ
CASTT_HD inline uint32_t varchar_to_timestamp(const char *str, size_t len, int64_t *out) {
	size_t pos;
	int32_t date;
	bool special;
	const uint32_t rc = parse_date(str, len, pos, &date, special);
	if (rc != OK) {
		return rc;
	}
	if (pos == len) {
		if (date == INT32_MAX) {
			*out = INT64_MAX;
			return OK;
		}
		if (date == -INT32_MAX) {
			*out = -INT64_MAX;
			return OK;
		}
		return timestamp_of(date, 0, out) ? OK : OUT_OF_RANGE;
	}
	if (str[pos] == ' ' || str[pos] == 'T') {
		pos++;
	}
	size_t time_pos = 0;
	int64_t micros;
	if (!parse_time(str + pos, len - pos, time_pos, &micros)) {
		return BAD_FORMAT;
	}
	if (micros > 86400000000ll) {
		return OUT_OF_RANGE;
	}
	pos += time_pos;
	int64_t result;
	if (!timestamp_of(date, micros, &result)) {
		return OUT_OF_RANGE;
	}
	if (pos < len) {
		int hh, mm, ss;
		if (str[pos] == 'Z') {
			pos++;
		} else if (parse_utc_offset(str, pos, len, hh, mm, ss)) {
			const int64_t delta = hh * 3600000000ll + mm * 60000000ll + ss * 1000000ll;
			if ((delta < 0 && result > INT64_MAX + delta) || (delta > 0 && result < INT64_MIN + delta)) {
				return OUT_OF_RANGE;
			}
			result -= delta;
		} else {
			if (str[pos++] != ' ') {
				return BAD_FORMAT; // ERROR_NON_UTC_TIMEZONE
			}
			const size_t tz = pos;
			for (; pos < len && (is_alpha(str[pos]) || is_digit(str[pos]) || str[pos] == '_' || str[pos] == '/' || str[pos] == '+' ||
			                     str[pos] == '-');
			     ++pos) {
			}
			const size_t tz_len = pos - tz;
			while (pos < len && is_space(str[pos])) {
				pos++;
			}
			if (pos < len) {
				return BAD_FORMAT;
			}
			if (tz_len != 0 && !(tz_len == 3 && lower(str[tz]) == 'u' && lower(str[tz + 1]) == 't' && lower(str[tz + 2]) == 'c')) {
				return BAD_FORMAT; // a time zone name: needs ICU in the reference too
			}
			*out = result;
			return OK;
		}
		while (pos < len && is_space(str[pos])) {
			pos++;
		}
		if (pos < len) {
			return BAD_FORMAT;
		}
	}
	*out = result;
	return OK;
}

} // namespace castt
} // namespace b2a
