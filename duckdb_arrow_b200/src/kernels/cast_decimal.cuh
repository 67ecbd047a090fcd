// Casts into and out of DECIMAL as DuckDB does them, on 128-bit integers whatever the physical width (the results are the
// same: every intermediate of the reference's narrower arithmetic fits its type):
//   integer -> DECIMAL(w,s)        StandardNumericToDecimalCast / NumericToHugeDecimalCast
//                                  (duckdb/src/common/operator/cast_operators.cpp:2005-2046): |v| < 10^(w-s), v x 10^s
//   DECIMAL -> DECIMAL             TemplatedDecimalScaleUp / ScaleDown (duckdb/src/function/cast/decimal_cast.cpp:53-190):
//                                  up = x 10^d with a limit check only when the source may not fit; down = round half away
//                                  from zero as (v / (10^d / 2) +- 1) / 2, with the reference's range test (whose rounding
//                                  looks at the WHOLE fraction — 10^source_scale — not at the digits dropped)
//   DECIMAL -> integer             TryCastDecimalToNumeric / TryCastHugeDecimalToNumeric (cast_operators.cpp:2399-2427): round
//                                  half away from zero, then the integer range check
//   FLOAT / DOUBLE -> DECIMAL      DoubleToDecimalCast (cast_operators.cpp:2333-2346): round(v x 10^s) in double, |.| < 10^w,
//                                  then back through the SOURCE type — a FLOAT's product is rounded to float again before
//                                  it becomes an integer, so DECIMAL(9,.) can receive 1000000000 — restated as it is
//   DECIMAL -> DOUBLE / FLOAT      TryCastDecimalToFloatingPoint (cast_operators.cpp:2672-2755): one division when the
//                                  integer is exact in the target (or scale 0), else integer part + fraction / 10^s;
//                                  128-bit integers become floating point as low + high x 2^64 (hugeint.cpp:649-661), not
//                                  correctly rounded — restated as it is
// `__host__ __device__`: pinned on the CPU against the compiled reference (tests/test_cast_parse.py).
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define CASTX_HD __host__ __device__
#else
#define CASTX_HD
#endif

namespace b2a {
namespace castx {

typedef __int128 i128;

CASTX_HD inline i128 pow10(uint32_t k) { // k <= 38
	i128 p = 1;
	for (uint32_t i = 0; i < k; i++) {
		p *= 10;
	}
	return p;
}

// NumericHelper::DOUBLE_POWERS_OF_TEN: the literals 1e0 .. 1e38
#define B2A_DOUBLE_POW10                                                                                               \
	{1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, \
	 1e20, 1e21, 1e22, 1e23, 1e24, 1e25, 1e26, 1e27, 1e28, 1e29, 1e30, 1e31, 1e32, 1e33, 1e34, 1e35, 1e36, 1e37, 1e38}
static const double h_double_pow10[39] = B2A_DOUBLE_POW10;
#if defined(__CUDACC__)
static __device__ const double d_double_pow10[39] = B2A_DOUBLE_POW10;
#endif
CASTX_HD inline double pow10_double(uint32_t k) {
#if defined(__CUDA_ARCH__)
	return d_double_pow10[k];
#else
	return h_double_pow10[k];
#endif
}

// false = "Could not cast value ... to DECIMAL(w,s)"
CASTX_HD inline bool int_to_decimal(i128 input, uint32_t width, uint32_t scale, i128 *out) {
	const i128 max_width = pow10(width - scale);
	if (input >= max_width || input <= -max_width) {
		return false;
	}
	*out = input * pow10(scale);
	return true;
}

// false = "Casting value ... to type DECIMAL(w,s) failed: value is out of range!"
CASTX_HD inline bool decimal_to_decimal(i128 input, uint32_t sw, uint32_t ss, uint32_t dw, uint32_t ds, i128 *out) {
	if (ds >= ss) {
		const uint32_t diff = ds - ss;
		const uint32_t target_width = dw - diff;
		if (sw >= target_width) {
			const i128 limit = pow10(target_width);
			if (input >= limit || input <= -limit) {
				return false;
			}
		}
		*out = input * pow10(diff);
		return true;
	}
	const uint32_t diff = ss - ds;
	const uint32_t target_width = dw + diff;
	if (sw >= target_width) {
		const i128 limit = pow10(target_width), divisor = pow10(ss);
		i128 value = input % divisor, rounded = input;
		if (rounded < 0) {
			rounded = -rounded;
			value = -value;
		}
		if (value >= divisor / 2) {
			rounded += divisor;
		}
		if (!(rounded < limit && rounded > -limit)) {
			return false;
		}
	}
	const i128 scaling = pow10(diff) / 2;
	input /= scaling;
	input += input < 0 ? -1 : 1;
	*out = input / 2;
	return true;
}

// Cast::Operation<SRC, double>: a static_cast for the 16 / 32 / 64-bit physical types, CastBigintToFloating for 128 bits
CASTX_HD inline double to_double(i128 v, int src_bytes) {
	if (src_bytes < 16) {
		return static_cast<double>(static_cast<int64_t>(v));
	}
	const uint64_t lower = static_cast<uint64_t>(v);
	const int64_t upper = static_cast<int64_t>(v >> 64);
	if (upper == -1) {
		return -static_cast<double>(UINT64_MAX - lower) - 1;
	}
	return static_cast<double>(lower) + static_cast<double>(upper) * (static_cast<double>(UINT64_MAX) + 1);
}
CASTX_HD inline float to_float(i128 v, int src_bytes) {
	if (src_bytes < 16) {
		return static_cast<float>(static_cast<int64_t>(v));
	}
	return static_cast<float>(to_double(v, 16));
}

CASTX_HD inline double decimal_to_double(i128 input, int src_bytes, uint32_t scale) {
	const i128 exact = static_cast<i128>(1) << 53;
	if (src_bytes <= 4 || (input <= exact && input >= -exact) || scale == 0) {
		return to_double(input, src_bytes) / pow10_double(scale);
	}
	const i128 p = pow10(scale);
	return to_double(input / p, src_bytes) + to_double(input % p, src_bytes) / pow10_double(scale);
}

CASTX_HD inline float decimal_to_float(i128 input, int src_bytes, uint32_t scale) {
	const i128 exact = static_cast<i128>(1) << 24;
	if (src_bytes <= 2 || (input <= exact && input >= -exact) || scale == 0) {
		return to_float(input, src_bytes) / static_cast<float>(pow10_double(scale));
	}
	const i128 p = pow10(scale);
	return to_float(input / p, src_bytes) + to_float(input % p, src_bytes) / static_cast<float>(pow10_double(scale));
}

// false = "Failed to cast decimal value ... to type ..."; lo / hi = the range of the destination integer type
CASTX_HD inline bool decimal_to_int(i128 input, uint32_t scale, i128 lo, i128 hi, i128 *out) {
	const i128 power = pow10(scale);
	const i128 rounding = (input < 0 ? -power : power) / 2;
	const i128 scaled = (input + rounding) / power;
	if (scaled < lo || scaled > hi) {
		return false;
	}
	*out = scaled;
	return true;
}

// an integer-valued finite double, |v| < 2^127, as a 128-bit integer (what ConvertFloatingToBigint's fmod / division by 2^64
// and a static_cast to int64 both give)
CASTX_HD inline i128 integral_double_to_i128(double v) {
	uint64_t bits;
	memcpy(&bits, &v, sizeof(bits));
	const int exponent = static_cast<int>((bits >> 52) & 0x7FF);
	if (exponent < 1023) {
		return 0; // |v| < 1
	}
	const i128 mantissa = static_cast<i128>((bits & ((uint64_t(1) << 52) - 1)) | (uint64_t(1) << 52));
	const int shift = exponent - 1075;
	const i128 magnitude = shift >= 0 ? mantissa << shift : mantissa >> -shift;
	return (bits >> 63) ? -magnitude : magnitude;
}

CASTX_HD inline bool finite(double v) {
	uint64_t bits;
	memcpy(&bits, &v, sizeof(bits));
	return ((bits >> 52) & 0x7FF) != 0x7FF;
}

// round half away from zero of a double (std::round), without <cmath> differences between host and device
CASTX_HD inline double round_half_away(double v) {
#if defined(__CUDA_ARCH__)
	return ::round(v);
#else
	return __builtin_round(v);
#endif
}
CASTX_HD inline double nearby_even(double v) {
#if defined(__CUDA_ARCH__)
	return ::rint(v);
#else
	return __builtin_rint(v);
#endif
}

// false = "Could not cast value ... to DECIMAL(w,s)".  `from_float`: the source is a FLOAT (passed widened to double)
CASTX_HD inline bool real_to_decimal(double input, bool from_float, uint32_t width, uint32_t scale, i128 *out) {
	const double value = input * pow10_double(scale);
	const double rounded = round_half_away(value);
	if (rounded <= -pow10_double(width) || rounded >= pow10_double(width) || !finite(rounded)) {
		return false;
	}
	// Cast::Operation<SRC, DST>(static_cast<SRC>(rounded)): nearbyint of the value in the SOURCE type
	const double back = from_float ? static_cast<double>(static_cast<float>(rounded)) : rounded;
	*out = integral_double_to_i128(nearby_even(back));
	return true;
}

} // namespace castx
} // namespace b2a
