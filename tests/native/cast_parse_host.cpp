// CPU build of the device VARCHAR -> BIGINT / DOUBLE casts (src/kernels/cast_parse.cuh and cast_parse_double.cuh are
// __host__ __device__) so that they can be pinned against the compiled reference's own casts (SELECT TRY_CAST(s AS BIGINT /
// DOUBLE)) without a GPU.  Test infrastructure only.
#include "kernels/cast_parse.cuh"
#include "kernels/cast_parse_double.cuh"

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
