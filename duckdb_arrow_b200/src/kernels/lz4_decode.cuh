// LZ4 frame decoder (LZ4 Frame Format 1.6.x + LZ4 Block Format) written for the device: one decoder instance per
// compressed IPC buffer, all state in registers, no workspace.  Arrow's BodyCompression knows two codecs, LZ4_FRAME (0)
// and ZSTD (1) (format/Message.fbs); the reference installs a decompressor for ZSTD only
// (src/ipc/stream_reader/base_stream_reader.cpp:38-50), so an LZ4_FRAME body ends in nanoarrow's "no decompressor"
// error there.  SURVEY §8f N-4 lists the second codec as the next step of the scan path; the arrow-testing fixtures
// 2.0.0-compression/generated_lz4 + generated_uncompressible_lz4 (golden contents in their *.json.gz) pin it.
//
// The code is `__host__ __device__` like zstd_decode.cuh: the CPU test-suite runs the very same decoder against frames
// produced by pyarrow's (real) liblz4; the product only ever calls it from b2a_decompress_kernel.
//
// Execution model on the device: all 32 lanes of a warp run the decoder with identical control flow (every branch
// depends only on bytes all lanes read alike: token, length bytes, offset).  The serial part — the sequence parse —
// costs what one lane would cost; the copies (literals from the frame, matches from the output) are strided over the
// warp.  An overlapping match is periodic in its offset, so every byte of it is fetched from output that existed before
// the match began.  On the host the same code runs with lane 0 of 1.
//
// Supported: version-01 frames with any block maximum size, linked or independent blocks, uncompressed blocks, block /
// content checksums (skipped, not verified), content size field, several concatenated frames.  Not supported (error):
// dictionary ids, skippable and legacy frames.
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define LZ4_HD __host__ __device__
#else
#define LZ4_HD
#endif

namespace b2a {
namespace lz4 {

enum Error : int {
	OK = 0,
	ERR_MAGIC = 1,       // "ERROR_frameType_unknown"
	ERR_TRUNCATED = 2,   // "ERROR_srcSize_wrong"
	ERR_CORRUPT = 3,     // "ERROR_decompressionFailed"
	ERR_DST_SIZE = 4,    // "ERROR_dstMaxSize_tooSmall" / size mismatch
	ERR_UNSUPPORTED = 5  // "ERROR_parameter_unsupported" (dictionary id, reserved bits, version)
};

struct Ctx {
	int lane;   // 0 .. nlanes-1
	int nlanes; // 32 on the device, 1 on the host
};

LZ4_HD inline void lanes_sync(const Ctx &) {
#if defined(__CUDA_ARCH__)
	__syncwarp();
#endif
}

LZ4_HD inline uint32_t load_le32(const uint8_t *p) {
	return static_cast<uint32_t>(p[0]) | static_cast<uint32_t>(p[1]) << 8 | static_cast<uint32_t>(p[2]) << 16 |
	       static_cast<uint32_t>(p[3]) << 24;
}

// d[0..n) = s[0..n), strided over the lanes, four loads in flight per lane; the ranges do not overlap
LZ4_HD inline void copy_strided(const Ctx &cx, const uint8_t *s, uint8_t *d, size_t n) {
	const size_t step = static_cast<size_t>(cx.nlanes);
	size_t i = static_cast<size_t>(cx.lane);
	for (; i + 3 * step < n; i += 4 * step) {
		const uint8_t v0 = s[i], v1 = s[i + step], v2 = s[i + 2 * step], v3 = s[i + 3 * step];
		d[i] = v0;
		d[i + step] = v1;
		d[i + 2 * step] = v2;
		d[i + 3 * step] = v3;
	}
	for (; i < n; i += step) {
		d[i] = s[i];
	}
}

// wide variant for long non-overlapping runs: bytes up to the first 16-byte boundary of d, then 16-byte stores assembled
// from two aligned 16-byte loads of s (funnel shift), then the tail.  Device only; the host takes copy_strided.
LZ4_HD inline void copy_run(const Ctx &cx, const uint8_t *s, uint8_t *d, size_t n) {
#if defined(__CUDA_ARCH__)
	if (n >= 256) {
		const size_t head = (16 - (reinterpret_cast<uintptr_t>(d) & 15)) & 15;
		if (static_cast<size_t>(cx.lane) < head) {
			d[cx.lane] = s[cx.lane];
		}
		const uint8_t *sp = s + head;
		uint8_t *dp = d + head;
		const size_t nv = (n - head) >> 4;
		const uint32_t sh = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(sp) & 7) * 8;
		const uint64_t *s8 = reinterpret_cast<const uint64_t *>(reinterpret_cast<uintptr_t>(sp) & ~uintptr_t(7));
		// the last vector may need one 8-byte word that ends past s + n: those bytes are inside the word that holds
		// s[n-1] whenever sh == 0, and otherwise the word is read only if it starts before s + n (same 8-byte word rule
		// as every aligned load: a word that contains one valid byte is readable)
		for (size_t v = static_cast<size_t>(cx.lane); v < nv; v += 32) {
			const uint64_t a = s8[2 * v], b = s8[2 * v + 1];
			uint64_t lo, hi;
			if (sh == 0) {
				lo = a;
				hi = b;
			} else {
				const uint64_t c = s8[2 * v + 2];
				lo = (a >> sh) | (b << (64 - sh));
				hi = (b >> sh) | (c << (64 - sh));
			}
			*reinterpret_cast<uint4 *>(dp + 16 * v) =
			    make_uint4(static_cast<uint32_t>(lo), static_cast<uint32_t>(lo >> 32), static_cast<uint32_t>(hi),
			               static_cast<uint32_t>(hi >> 32));
		}
		const size_t done = head + (nv << 4);
		if (done + static_cast<size_t>(cx.lane) < n) {
			d[done + cx.lane] = s[done + cx.lane];
		}
		return;
	}
#endif
	copy_strided(cx, s, d, n);
}

// one match of `ml` bytes at distance `off` behind d, strided over the lanes
LZ4_HD inline void copy_match(const Ctx &cx, uint8_t *d, size_t ml, size_t off) {
	const uint8_t *m = d - off;
	const size_t lane = static_cast<size_t>(cx.lane), step = static_cast<size_t>(cx.nlanes);
	if (off >= ml) {
		copy_run(cx, m, d, ml);
	} else if (off == 1) {
		const uint8_t v = m[0];
		for (size_t k = lane; k < ml; k += step) {
			d[k] = v;
		}
	} else {
		// periodic: byte k of the match is byte (k mod off) of the period, all of which existed before the match.
		// k advances by `step` per iteration, so the remainder advances by step mod off without a division per byte
		size_t r = lane % off;
		const size_t dr = step % off;
		for (size_t k = lane; k < ml; k += step) {
			d[k] = m[r];
			r += dr;
			r = r >= off ? r - off : r;
		}
	}
}

// one LZ4 block: src[0..len) -> dst[*pos ..), never past dst_cap; matches may reach back to dst[0] (linked blocks)
LZ4_HD inline int decode_block(const Ctx &cx, const uint8_t *src, size_t len, uint8_t *dst, size_t dst_cap, size_t *pos) {
	size_t ip = 0, op = *pos;
	while (true) {
		if (ip >= len) {
			return ERR_CORRUPT; // a block ends with a literals-only sequence, never between sequences
		}
		const uint32_t token = src[ip++];
		size_t ll = token >> 4;
		if (ll == 15) {
			uint32_t b;
			do {
				if (ip >= len) {
					return ERR_CORRUPT;
				}
				b = src[ip++];
				ll += b;
			} while (b == 255);
		}
		if (ll > len - ip || ll > dst_cap - op) {
			return ll > len - ip ? ERR_CORRUPT : ERR_DST_SIZE;
		}
		copy_run(cx, src + ip, dst + op, ll);
		ip += ll;
		op += ll;
		if (ip == len) {
			break; // last sequence: literals only
		}
		if (len - ip < 2) {
			return ERR_CORRUPT;
		}
		const size_t off = static_cast<size_t>(src[ip]) | static_cast<size_t>(src[ip + 1]) << 8;
		ip += 2;
		size_t ml = token & 15;
		if (ml == 15) {
			uint32_t b;
			do {
				if (ip >= len) {
					return ERR_CORRUPT;
				}
				b = src[ip++];
				ml += b;
			} while (b == 255);
		}
		ml += 4;
		if (off == 0 || off > op) {
			return ERR_CORRUPT;
		}
		if (ml > dst_cap - op) {
			return ERR_DST_SIZE;
		}
		lanes_sync(cx); // the literals just stored and every earlier match are visible to all lanes
		copy_match(cx, dst + op, ml, off);
		op += ml;
	}
	lanes_sync(cx);
	*pos = op;
	return OK;
}

// src[0..len) = one or more LZ4 frames whose contents concatenate to exactly dst_size bytes
LZ4_HD inline int decode_frame(const Ctx &cx, const uint8_t *src, size_t len, uint8_t *dst, size_t dst_size) {
	size_t ip = 0, op = 0;
	bool any = false;
	while (ip < len) {
		if (len - ip < 4) {
			return ERR_TRUNCATED;
		}
		if (load_le32(src + ip) != 0x184D2204u) {
			return ERR_MAGIC; // skippable (0x184D2A5x) and legacy (0x184C2102) frames included
		}
		if (len - ip < 7) {
			return ERR_TRUNCATED;
		}
		const uint32_t flg = src[ip + 4], bd = src[ip + 5];
		ip += 6;
		if ((flg >> 6) != 1 || (flg & 0x02) || (bd & 0x8F)) {
			return ERR_UNSUPPORTED;
		}
		if (flg & 0x01) {
			return ERR_UNSUPPORTED; // dictionary id
		}
		const uint32_t bs_id = (bd >> 4) & 7;
		if (bs_id < 4) {
			return ERR_UNSUPPORTED;
		}
		const size_t block_max = size_t(1) << (8 + 2 * bs_id); // 64 KiB, 256 KiB, 1 MiB, 4 MiB
		const bool block_checksum = (flg & 0x10) != 0, content_size = (flg & 0x08) != 0, content_checksum = (flg & 0x04) != 0;
		uint64_t declared = 0;
		if (content_size) {
			if (len - ip < 8) {
				return ERR_TRUNCATED;
			}
			declared = static_cast<uint64_t>(load_le32(src + ip)) | static_cast<uint64_t>(load_le32(src + ip + 4)) << 32;
			ip += 8;
		}
		if (len - ip < 1) {
			return ERR_TRUNCATED;
		}
		ip += 1; // header checksum byte (not verified)
		const size_t frame_start = op;
		while (true) {
			if (len - ip < 4) {
				return ERR_TRUNCATED;
			}
			const uint32_t word = load_le32(src + ip);
			ip += 4;
			if (word == 0) {
				break; // EndMark
			}
			const size_t bsize = word & 0x7FFFFFFFu;
			if (bsize > block_max) {
				return ERR_CORRUPT;
			}
			if (bsize > len - ip) {
				return ERR_TRUNCATED;
			}
			if (word & 0x80000000u) {
				if (bsize > dst_size - op) {
					return ERR_DST_SIZE;
				}
				copy_run(cx, src + ip, dst + op, bsize);
				op += bsize;
				lanes_sync(cx);
			} else {
				const size_t before = op;
				const int rc = decode_block(cx, src + ip, bsize, dst, dst_size, &op);
				if (rc != OK) {
					return rc;
				}
				if (op - before > block_max) {
					return ERR_CORRUPT;
				}
			}
			ip += bsize;
			if (block_checksum) {
				if (len - ip < 4) {
					return ERR_TRUNCATED;
				}
				ip += 4;
			}
		}
		if (content_checksum) {
			if (len - ip < 4) {
				return ERR_TRUNCATED;
			}
			ip += 4;
		}
		if (content_size && declared != op - frame_start) {
			return ERR_DST_SIZE;
		}
		any = true;
	}
	if (!any && dst_size != 0) {
		return ERR_TRUNCATED;
	}
	return op == dst_size ? OK : ERR_DST_SIZE;
}

} // namespace lz4
} // namespace b2a
