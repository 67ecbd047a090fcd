// CPU build of the device LZ4 frame decoder (src/kernels/lz4_decode.cuh is __host__ __device__) so that the decoder logic
// can be pinned against frames produced by the real liblz4 (through pyarrow) without a GPU.  Test infrastructure only.
#include "kernels/lz4_decode.cuh"

extern "C" int b2a_test_lz4_decode(const uint8_t *src, size_t len, uint8_t *dst, size_t dst_size) {
	const b2a::lz4::Ctx cx = {0, 1};
	return b2a::lz4::decode_frame(cx, src, len, dst, dst_size);
}
