// CPU build of the device VARCHAR -> BIGINT cast (src/kernels/cast_parse.cuh is __host__ __device__) so that it can be pinned
// against the compiled reference's own cast (SELECT TRY_CAST(s AS BIGINT)) without a GPU.  Test infrastructure only.
#include "kernels/cast_parse.cuh"

extern "C" unsigned b2a_test_varchar_to_bigint(const char *s, size_t n, long long *out) {
	int64_t v = 0;
	const uint32_t rc = b2a::castp::varchar_to_bigint(s, n, &v);
	*out = v;
	return rc;
}
