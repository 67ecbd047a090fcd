// CPU build of the device zstd decoder (src/kernels/zstd_decode.cuh is __host__ __device__) so that the decoder logic can
// be pinned against frames produced by the real libzstd (through pyarrow) without a GPU.  Test infrastructure only.
#include "kernels/zstd_decode.cuh"

#include <new>

extern "C" int b2a_test_zstd_decode(const uint8_t *src, size_t len, uint8_t *dst, size_t dst_size) {
	b2a::zstd::Workspace *w = new (std::nothrow) b2a::zstd::Workspace();
	if (!w) {
		return -1;
	}
	b2a::zstd::Tables *t = new (std::nothrow) b2a::zstd::Tables();
	if (!t) {
		delete w;
		return -1;
	}
	const b2a::zstd::Ctx cx = {0, 1, t};
	int rc = b2a::zstd::decode_frame(cx, *w, src, len, dst, dst_size);
	delete t;
	delete w;
	return rc;
}
