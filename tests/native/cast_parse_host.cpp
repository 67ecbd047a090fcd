// CPU build of the device VARCHAR -> BIGINT / DOUBLE casts (src/kernels/cast_parse.cuh and cast_parse_double.cuh are
// __host__ __device__) so that they can be pinned against the compiled reference's own casts (SELECT TRY_CAST(s AS BIGINT /
// DOUBLE)) without a GPU.  Test infrastructure only.
#include "kernels/cast_parse.cuh"
#include "kernels/cast_parse_double.cuh"
#include "kernels/cast_format.cuh"
#include "kernels/cast_decimal.cuh"
#include "kernels/cast_parse_temporal.cuh"
#include "kernels/cast_parse_decimal.cuh"

extern "C" unsigned b2a_test_varchar_to_bigint(const char *s, size_t n, long long *out) {
	int64_t v = 0;
	const uint32_t rc = b2a::castp::varchar_to_bigint(s, n, &v);
	*out = v;
	return rc;
}

// bits of the double (so that -0.0, the NaN sign and every last bit are compared); rc as castd::Result
extern "C" unsigned b2a_test_varchar_to_double(const char *s, size_t n, unsigned long long *bits) {
	double v = 0;
	const uint32_t rc = b2a::castd::varchar_to_double(s, n, &v);
	uint64_t b;
	memcpy(&b, &v, sizeof(b));
	*bits = b;
	return rc;
}

// kind: 0 BOOLEAN, 1 DATE, 2 TIME, 3 TIMESTAMP, 4 DECIMAL / HUGEINT (value = low / high 64 bits, width, scale); -> length
extern "C" unsigned b2a_test_format(int kind, unsigned long long low, long long high, unsigned width, unsigned scale, char *out) {
	using namespace b2a::castf;
	switch (kind) {
	case 0:
		return format_bool(low != 0, out);
	case 1:
		return format_date(static_cast<int32_t>(static_cast<int64_t>(low)), out);
	case 2:
		return format_time(static_cast<int64_t>(low), out);
	case 3:
		return format_timestamp(static_cast<int64_t>(low), out);
	default:
		return format_decimal((static_cast<__int128>(high) << 64) | static_cast<__int128>(low), width, scale, out);
	}
}

static __int128 wide(unsigned long long low, long long high) {
	return (static_cast<__int128>(high) << 64) | static_cast<__int128>(low);
}
// op 0: integer -> DECIMAL(dw,ds); 1: DECIMAL(sw,ss) -> DECIMAL(dw,ds); 2: DECIMAL -> integer; 3 / 4: DOUBLE / FLOAT ->
// DECIMAL.  Returns 1 when the cast succeeds.
extern "C" int b2a_test_to_decimal(int op, unsigned long long low, long long high, unsigned sw, unsigned ss, unsigned dw,
                                   unsigned ds, unsigned long long *out_low, long long *out_high) {
	__int128 r = 0;
	bool ok;
	if (op == 0) {
		ok = b2a::castx::int_to_decimal(wide(low, high), dw, ds, &r);
	} else if (op == 1) {
		ok = b2a::castx::decimal_to_decimal(wide(low, high), sw, ss, dw, ds, &r);
	} else if (op == 2) { // DECIMAL(., ss) -> a signed (ds = 1) or unsigned (ds = 0) integer of dw bytes
		const __int128 one = 1;
		const __int128 lo = ds ? -(one << (8 * dw - 1)) : 0;
		const __int128 hi = ds ? (one << (8 * dw - 1)) - 1 : (dw == 16 ? ~static_cast<__int128>(0) : (one << (8 * dw)) - 1);
		ok = b2a::castx::decimal_to_int(wide(low, high), ss, lo, hi, &r);
	} else { // op 3 / 4: the bits of a DOUBLE / FLOAT (in `low`) -> DECIMAL(dw,ds)
		double v;
		if (op == 3) {
			memcpy(&v, &low, sizeof(v));
		} else {
			float f;
			const uint32_t b = static_cast<uint32_t>(low);
			memcpy(&f, &b, sizeof(f));
			v = f;
		}
		ok = b2a::castx::real_to_decimal(v, op == 4, dw, ds, &r);
	}
	*out_low = static_cast<unsigned long long>(r);
	*out_high = static_cast<long long>(r >> 64);
	return ok ? 1 : 0;
}
// DECIMAL of `src_bytes` physical bytes -> the bits of the DOUBLE (as_float = 0) or FLOAT (1)
extern "C" unsigned long long b2a_test_decimal_to_real(unsigned long long low, long long high, int src_bytes, unsigned scale,
                                                       int as_float) {
	if (as_float) {
		const float f = b2a::castx::decimal_to_float(wide(low, high), src_bytes, scale);
		uint32_t b;
		memcpy(&b, &f, sizeof(b));
		return b;
	}
	const double d = b2a::castx::decimal_to_double(wide(low, high), src_bytes, scale);
	uint64_t b;
	memcpy(&b, &d, sizeof(b));
	return b;
}

// kind 0: VARCHAR -> DATE (days), 1: VARCHAR -> TIMESTAMP (microseconds); rc as castt::Result
extern "C" unsigned b2a_test_parse_temporal(int kind, const char *s, size_t n, long long *out) {
	if (kind == 0) {
		int32_t d = 0;
		const uint32_t rc = b2a::castt::varchar_to_date(s, n, &d);
		*out = d;
		return rc;
	}
	int64_t t = 0;
	const uint32_t rc = b2a::castt::varchar_to_timestamp(s, n, &t);
	*out = t;
	return rc;
}

// VARCHAR -> DECIMAL(width, scale): 1 when the cast succeeds
extern "C" int b2a_test_parse_decimal(const char *s, size_t n, unsigned width, unsigned scale, unsigned long long *out_low,
                                      long long *out_high) {
	__int128 r = 0;
	const bool ok = b2a::castq::varchar_to_decimal(s, n, width, scale, &r);
	*out_low = static_cast<unsigned long long>(r);
	*out_high = static_cast<long long>(r >> 64);
	return ok ? 1 : 0;
}
