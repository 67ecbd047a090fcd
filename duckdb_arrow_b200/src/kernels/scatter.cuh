// b2a_scatter_kernel: places the packed upload ranges of one slice (scanner/host_stage.hpp) at their offsets in the
// device image.  Device-to-device, aligned 16-byte vectors (a piece keeps its phase inside a vector while staged).
#pragma once
#include <stdint.h>

namespace b2a {

struct ScatterPiece {
	uint64_t stage; // offset inside the slice's device staging buffer
	uint64_t dst;   // offset inside the device image
	uint64_t bytes;
};

// returns cudaError_t as int
int LaunchScatter(const ScatterPiece *d_pieces, uint32_t n_pieces, const uint8_t *d_stage, uint8_t *d_image, int sm_count,
                  void *stream);

} // namespace b2a
