// CPU build of the device zstd decoder (src/kernels/zstd_decode.cuh is __host__ __device__) so that the decoder logic can
// be pinned against frames produced by the real libzstd (through pyarrow) without a GPU.  Test infrastructure only.
#include "kernels/zstd_decode.cuh"

#include <new>

extern "C" int b2a_test_zstd_decode(const uint8_t *src, size_t len, uint8_t *dst, size_t dst_size) {
	b2a::zstd::Workspace *w = new (std::nothrow) b2a::zstd::Workspace();
	if (!w) {
		return -1;
	}
	int rc = b2a::zstd::decode_frame(*w, src, len, dst, dst_size);
	delete w;
	return rc;
}
