// Single-frame Zstandard decoder (RFC 8878) written for the device: one decoder instance per IPC buffer, driven by one
// lane, no recursion, no dynamic allocation, all state in a caller-provided workspace.  The reference decompresses each
// compressed IPC buffer with ZSTD_decompress (src/ipc/stream_reader/base_stream_reader.cpp:11-33, wired into nanoarrow's
// decoder at :38-50); a compressed IPC buffer is `int64 uncompressed_length` followed by exactly one zstd frame.
//
// The code is `__host__ __device__` so that the CPU test-suite can run the very same decoder against frames produced by
// pyarrow's (real) libzstd; the product only ever calls it from b2a_decompress_kernel.
//
// Supported: everything a frame produced by ZSTD_compress can contain — raw / RLE / compressed blocks, raw / RLE /
// Huffman (1 or 4 streams) / treeless literals, predefined / RLE / FSE / repeat sequence tables, repeat offsets, frame
// content size in all encodings, optional content checksum (skipped, not verified).  Not supported (error): dictionaries,
// skippable frames, multiple concatenated frames.
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define ZSTD_HD __host__ __device__
#else
#define ZSTD_HD
#endif

namespace b2a {
namespace zstd {

enum Error : int {
	OK = 0,
	ERR_MAGIC = 1,
	ERR_TRUNCATED = 2,
	ERR_CORRUPT = 3,
	ERR_DST_SIZE = 4,
	ERR_UNSUPPORTED = 5
};

static constexpr int kMaxBlock = 128 * 1024;
static constexpr int kHufMaxBits = 11;

struct FseEntry {
	uint16_t base;
	uint8_t symbol;
	uint8_t nbits;
};

struct FseTable {
	FseEntry e[512];
	int accuracy;
};

struct Workspace {
	uint8_t huf_symbol[1 << kHufMaxBits];
	uint8_t huf_nbits[1 << kHufMaxBits];
	int huf_max_bits; // 0: no table yet
	FseTable ll, of, ml;
	bool have_ll, have_of, have_ml;
	FseTable tmp; // Huffman weights
	uint64_t rep[3];
	uint8_t literals[kMaxBlock + 32];
};

// ------------------------------------------------------------------------------------------------ bit readers
ZSTD_HD inline int highbit(uint64_t v) { // index of the highest set bit, -1 for 0
	int r = -1;
	while (v) {
		v >>= 1;
		r++;
	}
	return r;
}

// up to 57 bits starting at absolute bit offset `off` of src (little-endian bit order), bytes past `len` read as zero
ZSTD_HD inline uint64_t read_bits_le(const uint8_t *src, size_t len, int nbits, int64_t off) {
	if (nbits == 0) {
		return 0;
	}
	const size_t byte = static_cast<size_t>(off >> 3);
	const int shift = static_cast<int>(off & 7);
	uint64_t v = 0;
	const int nbytes = (nbits + shift + 7) >> 3;
	for (int i = 0; i < nbytes; i++) {
		const size_t p = byte + static_cast<size_t>(i);
		v |= static_cast<uint64_t>(p < len ? src[p] : 0) << (8 * i);
	}
	return (v >> shift) & ((nbits >= 64) ? ~0ull : ((1ull << nbits) - 1));
}

// backward bitstream: `off` is the bit position just above the next bits to read; bits below position 0 read as zero
struct BackBits {
	const uint8_t *src;
	size_t len;
	int64_t off;
};

ZSTD_HD inline int backbits_init(BackBits &b, const uint8_t *src, size_t len) {
	if (len == 0 || src[len - 1] == 0) {
		return ERR_CORRUPT;
	}
	b.src = src;
	b.len = len;
	b.off = static_cast<int64_t>(len) * 8 - (8 - highbit(src[len - 1]));
	return OK;
}

ZSTD_HD inline uint64_t backbits_read(BackBits &b, int nbits) {
	b.off -= nbits;
	if (b.off >= 0) {
		return read_bits_le(b.src, b.len, nbits, b.off);
	}
	const int64_t missing = -b.off;
	if (missing >= nbits) {
		return 0;
	}
	const int have = nbits - static_cast<int>(missing);
	return read_bits_le(b.src, b.len, have, 0) << missing;
}

// ------------------------------------------------------------------------------------------------ FSE
// reads an FSE table description (forward bitstream); returns bytes consumed through *used
ZSTD_HD inline int fse_read_distribution(const uint8_t *src, size_t len, int max_accuracy, int max_symbols, int16_t *freq,
                                         int *n_symbols, int *accuracy, size_t *used) {
	if (len == 0) {
		return ERR_TRUNCATED;
	}
	int64_t off = 0;
	const int acc = static_cast<int>(read_bits_le(src, len, 4, off)) + 5;
	off += 4;
	if (acc > max_accuracy) {
		return ERR_CORRUPT;
	}
	int remaining = 1 << acc;
	int symb = 0;
	while (remaining > 0 && symb < max_symbols) {
		const int bits = highbit(static_cast<uint64_t>(remaining + 1)) + 1;
		int val = static_cast<int>(read_bits_le(src, len, bits, off));
		off += bits;
		const int lower_mask = (1 << (bits - 1)) - 1;
		const int threshold = (1 << bits) - 1 - (remaining + 1);
		if ((val & lower_mask) < threshold) {
			off -= 1;
			val &= lower_mask;
		} else if (val > lower_mask) {
			val -= threshold;
		}
		const int proba = val - 1;
		remaining -= proba < 0 ? -proba : proba;
		freq[symb++] = static_cast<int16_t>(proba);
		if (proba == 0) {
			int repeat = static_cast<int>(read_bits_le(src, len, 2, off));
			off += 2;
			while (true) {
				for (int i = 0; i < repeat && symb < max_symbols; i++) {
					freq[symb++] = 0;
				}
				if (repeat == 3) {
					repeat = static_cast<int>(read_bits_le(src, len, 2, off));
					off += 2;
				} else {
					break;
				}
			}
		}
	}
	if (remaining != 0 || symb > max_symbols) {
		return ERR_CORRUPT;
	}
	const size_t bytes = static_cast<size_t>((off + 7) >> 3);
	if (bytes > len) {
		return ERR_TRUNCATED;
	}
	*n_symbols = symb;
	*accuracy = acc;
	*used = bytes;
	return OK;
}

ZSTD_HD inline int fse_build(FseTable &t, const int16_t *freq, int n_symbols, int accuracy) {
	const int size = 1 << accuracy;
	t.accuracy = accuracy;
	uint16_t next[64]; // per-symbol state counter (max 53 symbols)
	if (n_symbols > 64) {
		return ERR_CORRUPT;
	}
	int high = size;
	for (int s = 0; s < n_symbols; s++) {
		if (freq[s] == -1) {
			t.e[--high].symbol = static_cast<uint8_t>(s);
			next[s] = 1;
		}
	}
	const int step = (size >> 1) + (size >> 3) + 3, mask = size - 1;
	int pos = 0;
	for (int s = 0; s < n_symbols; s++) {
		if (freq[s] <= 0) {
			continue;
		}
		next[s] = static_cast<uint16_t>(freq[s]);
		for (int i = 0; i < freq[s]; i++) {
			t.e[pos].symbol = static_cast<uint8_t>(s);
			do {
				pos = (pos + step) & mask;
			} while (pos >= high);
		}
	}
	if (pos != 0) {
		return ERR_CORRUPT;
	}
	for (int i = 0; i < size; i++) {
		const int s = t.e[i].symbol;
		const int n = next[s]++;
		const int nb = accuracy - highbit(static_cast<uint64_t>(n));
		t.e[i].nbits = static_cast<uint8_t>(nb);
		t.e[i].base = static_cast<uint16_t>((n << nb) - size);
	}
	return OK;
}

ZSTD_HD inline void fse_build_rle(FseTable &t, uint8_t symbol) {
	t.accuracy = 0;
	t.e[0].symbol = symbol;
	t.e[0].nbits = 0;
	t.e[0].base = 0;
}

// ------------------------------------------------------------------------------------------------ Huffman
ZSTD_HD inline int huf_build(Workspace &w, const uint8_t *weights, int n_weights) {
	// the last weight is implied: the sum of 2^(w-1) must be a power of two
	uint32_t sum = 0;
	for (int i = 0; i < n_weights; i++) {
		if (weights[i] > kHufMaxBits) {
			return ERR_CORRUPT;
		}
		sum += weights[i] ? (1u << (weights[i] - 1)) : 0;
	}
	if (sum == 0) {
		return ERR_CORRUPT;
	}
	const int max_bits = highbit(sum) + 1;
	if (max_bits > kHufMaxBits) {
		return ERR_CORRUPT;
	}
	const uint32_t left = (1u << max_bits) - sum;
	if (left == 0 || (left & (left - 1)) != 0) {
		return ERR_CORRUPT;
	}
	uint8_t bits[256];
	const int last_weight = highbit(left) + 1;
	const int n = n_weights + 1;
	if (n > 256) {
		return ERR_CORRUPT;
	}
	for (int i = 0; i < n_weights; i++) {
		bits[i] = weights[i] ? static_cast<uint8_t>(max_bits + 1 - weights[i]) : 0;
	}
	bits[n_weights] = static_cast<uint8_t>(max_bits + 1 - last_weight);
	uint32_t rank_count[kHufMaxBits + 2];
	uint32_t rank_idx[kHufMaxBits + 2];
	for (int i = 0; i <= kHufMaxBits + 1; i++) {
		rank_count[i] = 0;
	}
	for (int i = 0; i < n; i++) {
		rank_count[bits[i]]++;
	}
	rank_idx[max_bits] = 0;
	for (int i = max_bits; i >= 1; i--) {
		rank_idx[i - 1] = rank_idx[i] + rank_count[i] * (1u << (max_bits - i));
		for (uint32_t k = rank_idx[i]; k < rank_idx[i - 1]; k++) {
			w.huf_nbits[k] = static_cast<uint8_t>(i);
		}
	}
	if (rank_idx[0] != (1u << max_bits)) {
		return ERR_CORRUPT;
	}
	for (int i = 0; i < n; i++) {
		if (bits[i] == 0) {
			continue;
		}
		const uint32_t code = rank_idx[bits[i]];
		const uint32_t len = 1u << (max_bits - bits[i]);
		for (uint32_t k = 0; k < len; k++) {
			w.huf_symbol[code + k] = static_cast<uint8_t>(i);
		}
		rank_idx[bits[i]] += len;
	}
	w.huf_max_bits = max_bits;
	return OK;
}

ZSTD_HD inline int huf_read_tree(Workspace &w, const uint8_t *src, size_t len, size_t *used) {
	if (len == 0) {
		return ERR_TRUNCATED;
	}
	const int header = src[0];
	uint8_t weights[256];
	int n = 0;
	if (header >= 128) {
		n = header - 127;
		const size_t bytes = static_cast<size_t>((n + 1) / 2);
		if (1 + bytes > len) {
			return ERR_TRUNCATED;
		}
		for (int i = 0; i < n; i++) {
			const uint8_t b = src[1 + i / 2];
			weights[i] = (i & 1) ? (b & 0xF) : (b >> 4);
		}
		*used = 1 + bytes;
	} else {
		const size_t csize = static_cast<size_t>(header);
		if (1 + csize > len || csize == 0) {
			return ERR_TRUNCATED;
		}
		const uint8_t *p = src + 1;
		int16_t freq[16];
		int nsym = 0, acc = 0;
		size_t hdr = 0;
		int rc = fse_read_distribution(p, csize, 6, 13, freq, &nsym, &acc, &hdr);
		if (rc) {
			return rc;
		}
		rc = fse_build(w.tmp, freq, nsym, acc);
		if (rc) {
			return rc;
		}
		BackBits bb;
		rc = backbits_init(bb, p + hdr, csize - hdr);
		if (rc) {
			return rc;
		}
		uint32_t s1 = static_cast<uint32_t>(backbits_read(bb, acc));
		uint32_t s2 = static_cast<uint32_t>(backbits_read(bb, acc));
		// two interleaved states; stops when the stream runs dry
		while (true) {
			if (n >= 254) {
				return ERR_CORRUPT;
			}
			weights[n++] = w.tmp.e[s1].symbol;
			{
				const FseEntry e = w.tmp.e[s1];
				s1 = e.base + static_cast<uint32_t>(backbits_read(bb, e.nbits));
			}
			if (bb.off < 0) {
				weights[n++] = w.tmp.e[s2].symbol;
				break;
			}
			weights[n++] = w.tmp.e[s2].symbol;
			{
				const FseEntry e = w.tmp.e[s2];
				s2 = e.base + static_cast<uint32_t>(backbits_read(bb, e.nbits));
			}
			if (bb.off < 0) {
				weights[n++] = w.tmp.e[s1].symbol;
				break;
			}
		}
		*used = 1 + csize;
	}
	return huf_build(w, weights, n);
}

ZSTD_HD inline int huf_decode_stream(const Workspace &w, const uint8_t *src, size_t len, uint8_t *dst, size_t n_out) {
	BackBits bb;
	int rc = backbits_init(bb, src, len);
	if (rc) {
		return rc;
	}
	const int mb = w.huf_max_bits;
	const uint32_t mask = (1u << mb) - 1;
	uint32_t state = static_cast<uint32_t>(backbits_read(bb, mb));
	for (size_t i = 0; i < n_out; i++) {
		if (bb.off < -mb) {
			return ERR_CORRUPT;
		}
		dst[i] = w.huf_symbol[state];
		const int nb = w.huf_nbits[state];
		state = ((state << nb) | static_cast<uint32_t>(backbits_read(bb, nb))) & mask;
	}
	if (bb.off != -mb) {
		return ERR_CORRUPT;
	}
	return OK;
}

// ------------------------------------------------------------------------------------------------ sequences tables
ZSTD_HD inline void ll_code(int code, uint32_t *base, int *bits) {
	const uint32_t b[36] = {0,  1,  2,  3,  4,  5,  6,  7,  8,   9,   10,  11,   12,   13,   14,   15,    16,    18,
	                        20, 22, 24, 28, 32, 40, 48, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 65536};
	const uint8_t x[36] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 3, 3, 4, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};
	*base = b[code];
	*bits = x[code];
}
ZSTD_HD inline void ml_code(int code, uint32_t *base, int *bits) {
	const uint32_t b[53] = {3,  4,  5,  6,  7,  8,  9,  10, 11, 12, 13, 14, 15, 16, 17,  18,  19,  20,
	                        21, 22, 23, 24, 25, 26, 27, 28, 29, 30, 31, 32, 33, 34, 35,  37,  39,  41,
	                        43, 47, 51, 59, 67, 83, 99, 131, 259, 515, 1027, 2051, 4099, 8195, 16387, 32771, 65539};
	const uint8_t x[53] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0,
	                       0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 3, 3, 4, 4, 5, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};
	*base = b[code];
	*bits = x[code];
}

ZSTD_HD inline int seq_table(FseTable &t, bool &have, int mode, int which, const uint8_t *src, size_t len, size_t *used) {
	// which: 0 LL, 1 OF, 2 ML
	const int max_acc[3] = {9, 8, 9};
	const int max_sym[3] = {36, 32, 53};
	*used = 0;
	switch (mode) {
	case 0: { // predefined
		const int16_t ll[36] = {4, 3, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 3, 2, 1, 1, 1, 1, 1, -1, -1, -1, -1};
		const int16_t of[29] = {1, 1, 1, 1, 1, 1, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1};
		const int16_t ml[53] = {1, 4, 3, 2, 2, 2, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1,
		                        1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1};
		int rc = which == 0 ? fse_build(t, ll, 36, 6) : which == 1 ? fse_build(t, of, 29, 5) : fse_build(t, ml, 53, 6);
		have = rc == OK;
		return rc;
	}
	case 1: // RLE
		if (len < 1) {
			return ERR_TRUNCATED;
		}
		if (src[0] >= max_sym[which]) {
			return ERR_CORRUPT;
		}
		fse_build_rle(t, src[0]);
		have = true;
		*used = 1;
		return OK;
	case 2: { // FSE compressed
		int16_t freq[64];
		int nsym = 0, acc = 0;
		int rc = fse_read_distribution(src, len, max_acc[which], max_sym[which], freq, &nsym, &acc, used);
		if (rc) {
			return rc;
		}
		rc = fse_build(t, freq, nsym, acc);
		have = rc == OK;
		return rc;
	}
	default: // repeat
		return have ? OK : ERR_CORRUPT;
	}
}

// ------------------------------------------------------------------------------------------------ block
struct Literals {
	const uint8_t *ptr; // raw / decoded literals (ignored when rle)
	size_t size;
	bool rle;
	uint8_t rle_byte;
};

ZSTD_HD inline int decode_literals(Workspace &w, const uint8_t *src, size_t len, Literals &lit, size_t *used) {
	if (len < 1) {
		return ERR_TRUNCATED;
	}
	const int type = src[0] & 3, fmt = (src[0] >> 2) & 3;
	lit.rle = false;
	if (type < 2) {
		size_t hdr, size;
		if ((fmt & 1) == 0) {
			hdr = 1;
			size = src[0] >> 3;
		} else if (fmt == 1) {
			if (len < 2) {
				return ERR_TRUNCATED;
			}
			hdr = 2;
			size = (static_cast<size_t>(src[0]) | (static_cast<size_t>(src[1]) << 8)) >> 4;
		} else {
			if (len < 3) {
				return ERR_TRUNCATED;
			}
			hdr = 3;
			size = (static_cast<size_t>(src[0]) | (static_cast<size_t>(src[1]) << 8) | (static_cast<size_t>(src[2]) << 16)) >> 4;
		}
		if (size > static_cast<size_t>(kMaxBlock)) {
			return ERR_CORRUPT;
		}
		lit.size = size;
		if (type == 0) {
			if (hdr + size > len) {
				return ERR_TRUNCATED;
			}
			lit.ptr = src + hdr;
			*used = hdr + size;
		} else {
			if (hdr + 1 > len) {
				return ERR_TRUNCATED;
			}
			lit.rle = true;
			lit.rle_byte = src[hdr];
			lit.ptr = nullptr;
			*used = hdr + 1;
		}
		return OK;
	}
	// Huffman-compressed (2) or treeless (3)
	size_t hdr, regen, comp;
	int streams;
	if (len < 3) {
		return ERR_TRUNCATED;
	}
	if (fmt <= 1) {
		const uint32_t h = static_cast<uint32_t>(src[0]) | (static_cast<uint32_t>(src[1]) << 8) | (static_cast<uint32_t>(src[2]) << 16);
		hdr = 3;
		regen = (h >> 4) & 0x3FF;
		comp = (h >> 14) & 0x3FF;
		streams = fmt == 0 ? 1 : 4;
	} else if (fmt == 2) {
		if (len < 4) {
			return ERR_TRUNCATED;
		}
		const uint32_t h = static_cast<uint32_t>(src[0]) | (static_cast<uint32_t>(src[1]) << 8) | (static_cast<uint32_t>(src[2]) << 16) |
		                   (static_cast<uint32_t>(src[3]) << 24);
		hdr = 4;
		regen = (h >> 4) & 0x3FFF;
		comp = (h >> 18) & 0x3FFF;
		streams = 4;
	} else {
		if (len < 5) {
			return ERR_TRUNCATED;
		}
		const uint64_t h = static_cast<uint64_t>(src[0]) | (static_cast<uint64_t>(src[1]) << 8) | (static_cast<uint64_t>(src[2]) << 16) |
		                   (static_cast<uint64_t>(src[3]) << 24) | (static_cast<uint64_t>(src[4]) << 32);
		hdr = 5;
		regen = static_cast<size_t>((h >> 4) & 0x3FFFF);
		comp = static_cast<size_t>((h >> 22) & 0x3FFFF);
		streams = 4;
	}
	if (regen > static_cast<size_t>(kMaxBlock) || hdr + comp > len) {
		return ERR_CORRUPT;
	}
	const uint8_t *p = src + hdr;
	size_t left = comp;
	if (type == 2) {
		size_t tree = 0;
		int rc = huf_read_tree(w, p, left, &tree);
		if (rc) {
			return rc;
		}
		p += tree;
		left -= tree;
	} else if (w.huf_max_bits == 0) {
		return ERR_CORRUPT;
	}
	if (streams == 1) {
		int rc = huf_decode_stream(w, p, left, w.literals, regen);
		if (rc) {
			return rc;
		}
	} else {
		if (left < 6) {
			return ERR_CORRUPT;
		}
		const size_t s1 = p[0] | (static_cast<size_t>(p[1]) << 8), s2 = p[2] | (static_cast<size_t>(p[3]) << 8),
		             s3 = p[4] | (static_cast<size_t>(p[5]) << 8);
		if (6 + s1 + s2 + s3 > left) {
			return ERR_CORRUPT;
		}
		const size_t s4 = left - 6 - s1 - s2 - s3;
		const size_t seg = (regen + 3) / 4;
		if (seg * 3 > regen) {
			return ERR_CORRUPT;
		}
		const uint8_t *q = p + 6;
		int rc = huf_decode_stream(w, q, s1, w.literals, seg);
		rc = rc ? rc : huf_decode_stream(w, q + s1, s2, w.literals + seg, seg);
		rc = rc ? rc : huf_decode_stream(w, q + s1 + s2, s3, w.literals + 2 * seg, seg);
		rc = rc ? rc : huf_decode_stream(w, q + s1 + s2 + s3, s4, w.literals + 3 * seg, regen - 3 * seg);
		if (rc) {
			return rc;
		}
	}
	lit.ptr = w.literals;
	lit.size = regen;
	*used = hdr + comp;
	return OK;
}

ZSTD_HD inline void copy_literals(const Literals &lit, size_t from, uint8_t *dst, size_t n) {
	if (lit.rle) {
		for (size_t i = 0; i < n; i++) {
			dst[i] = lit.rle_byte;
		}
	} else {
		const uint8_t *s = lit.ptr + from;
		for (size_t i = 0; i < n; i++) {
			dst[i] = s[i];
		}
	}
}

// decodes one compressed block into dst[pos..); returns the new position through *pos
ZSTD_HD inline int decode_compressed_block(Workspace &w, const uint8_t *src, size_t len, uint8_t *dst, size_t cap,
                                           size_t *pos) {
	Literals lit;
	size_t used = 0;
	int rc = decode_literals(w, src, len, lit, &used);
	if (rc) {
		return rc;
	}
	const uint8_t *p = src + used;
	size_t left = len - used;
	size_t out = *pos;
	// sequences header
	if (left < 1) {
		return ERR_TRUNCATED;
	}
	size_t nseq = p[0];
	size_t hdr = 1;
	if (nseq >= 128) {
		if (nseq == 255) {
			if (left < 3) {
				return ERR_TRUNCATED;
			}
			nseq = static_cast<size_t>(p[1]) + (static_cast<size_t>(p[2]) << 8) + 0x7F00;
			hdr = 3;
		} else {
			if (left < 2) {
				return ERR_TRUNCATED;
			}
			nseq = ((nseq - 128) << 8) + p[1];
			hdr = 2;
		}
	}
	p += hdr;
	left -= hdr;
	size_t lit_pos = 0;
	if (nseq > 0) {
		if (left < 1) {
			return ERR_TRUNCATED;
		}
		const int modes = p[0];
		if (modes & 3) {
			return ERR_CORRUPT;
		}
		p++;
		left--;
		size_t u = 0;
		rc = seq_table(w.ll, w.have_ll, (modes >> 6) & 3, 0, p, left, &u);
		if (rc) {
			return rc;
		}
		p += u;
		left -= u;
		rc = seq_table(w.of, w.have_of, (modes >> 4) & 3, 1, p, left, &u);
		if (rc) {
			return rc;
		}
		p += u;
		left -= u;
		rc = seq_table(w.ml, w.have_ml, (modes >> 2) & 3, 2, p, left, &u);
		if (rc) {
			return rc;
		}
		p += u;
		left -= u;
		BackBits bb;
		rc = backbits_init(bb, p, left);
		if (rc) {
			return rc;
		}
		uint32_t ll_state = static_cast<uint32_t>(backbits_read(bb, w.ll.accuracy));
		uint32_t of_state = static_cast<uint32_t>(backbits_read(bb, w.of.accuracy));
		uint32_t ml_state = static_cast<uint32_t>(backbits_read(bb, w.ml.accuracy));
		for (size_t i = 0; i < nseq; i++) {
			const int of_code = w.of.e[of_state].symbol, llc = w.ll.e[ll_state].symbol, mlc = w.ml.e[ml_state].symbol;
			if (of_code > 31 || llc > 35 || mlc > 52) {
				return ERR_CORRUPT;
			}
			const uint64_t offset_value = (1ull << of_code) + backbits_read(bb, of_code);
			uint32_t mbase, lbase;
			int mbits, lbits;
			ml_code(mlc, &mbase, &mbits);
			ll_code(llc, &lbase, &lbits);
			const size_t match_len = mbase + static_cast<size_t>(backbits_read(bb, mbits));
			const size_t lit_len = lbase + static_cast<size_t>(backbits_read(bb, lbits));
			if (i + 1 < nseq) {
				const FseEntry a = w.ll.e[ll_state];
				ll_state = a.base + static_cast<uint32_t>(backbits_read(bb, a.nbits));
				const FseEntry b = w.ml.e[ml_state];
				ml_state = b.base + static_cast<uint32_t>(backbits_read(bb, b.nbits));
				const FseEntry c = w.of.e[of_state];
				of_state = c.base + static_cast<uint32_t>(backbits_read(bb, c.nbits));
			}
			if (bb.off < 0) {
				return ERR_CORRUPT;
			}
			// repeat offsets
			uint64_t offset;
			if (offset_value > 3) {
				offset = offset_value - 3;
				w.rep[2] = w.rep[1];
				w.rep[1] = w.rep[0];
				w.rep[0] = offset;
			} else {
				uint32_t idx = static_cast<uint32_t>(offset_value) - 1 + (lit_len == 0 ? 1 : 0);
				if (idx == 0) {
					offset = w.rep[0];
				} else {
					offset = idx < 3 ? w.rep[idx] : w.rep[0] - 1;
					if (idx > 1) {
						w.rep[2] = w.rep[1];
					}
					w.rep[1] = w.rep[0];
					w.rep[0] = offset;
				}
			}
			if (lit_pos + lit_len > lit.size || out + lit_len + match_len > cap) {
				return ERR_CORRUPT;
			}
			copy_literals(lit, lit_pos, dst + out, lit_len);
			lit_pos += lit_len;
			out += lit_len;
			if (offset == 0 || offset > out) {
				return ERR_CORRUPT;
			}
			const uint8_t *m = dst + out - offset;
			for (size_t k = 0; k < match_len; k++) {
				dst[out + k] = m[k];
			}
			out += match_len;
		}
		if (bb.off != 0) {
			return ERR_CORRUPT;
		}
	}
	const size_t rest = lit.size - lit_pos;
	if (out + rest > cap) {
		return ERR_CORRUPT;
	}
	copy_literals(lit, lit_pos, dst + out, rest);
	out += rest;
	*pos = out;
	return OK;
}

// ------------------------------------------------------------------------------------------------ frame
ZSTD_HD inline int decode_frame(Workspace &w, const uint8_t *src, size_t len, uint8_t *dst, size_t dst_size) {
	if (len < 6) {
		return ERR_TRUNCATED;
	}
	const uint32_t magic = static_cast<uint32_t>(src[0]) | (static_cast<uint32_t>(src[1]) << 8) |
	                       (static_cast<uint32_t>(src[2]) << 16) | (static_cast<uint32_t>(src[3]) << 24);
	if (magic != 0xFD2FB528u) {
		return (magic & 0xFFFFFFF0u) == 0x184D2A50u ? ERR_UNSUPPORTED : ERR_MAGIC;
	}
	const int fhd = src[4];
	const int fcs_flag = fhd >> 6, single = (fhd >> 5) & 1, checksum = (fhd >> 2) & 1, dict_flag = fhd & 3;
	if (fhd & 0x08) {
		return ERR_CORRUPT;
	}
	size_t p = 5;
	if (!single) {
		p += 1; // window descriptor: the whole output is addressable here, the window size is not needed
	}
	if (dict_flag) {
		return ERR_UNSUPPORTED;
	}
	const int fcs_bytes = fcs_flag == 0 ? (single ? 1 : 0) : (fcs_flag == 1 ? 2 : (fcs_flag == 2 ? 4 : 8));
	if (p + static_cast<size_t>(fcs_bytes) > len) {
		return ERR_TRUNCATED;
	}
	if (fcs_bytes) {
		uint64_t fcs = 0;
		for (int i = 0; i < fcs_bytes; i++) {
			fcs |= static_cast<uint64_t>(src[p + static_cast<size_t>(i)]) << (8 * i);
		}
		if (fcs_bytes == 2) {
			fcs += 256;
		}
		if (fcs != dst_size) {
			return ERR_DST_SIZE;
		}
		p += static_cast<size_t>(fcs_bytes);
	}
	w.huf_max_bits = 0;
	w.have_ll = w.have_of = w.have_ml = false;
	w.rep[0] = 1;
	w.rep[1] = 4;
	w.rep[2] = 8;
	size_t out = 0;
	while (true) {
		if (p + 3 > len) {
			return ERR_TRUNCATED;
		}
		const uint32_t bh = static_cast<uint32_t>(src[p]) | (static_cast<uint32_t>(src[p + 1]) << 8) |
		                    (static_cast<uint32_t>(src[p + 2]) << 16);
		p += 3;
		const bool last = bh & 1;
		const int type = (bh >> 1) & 3;
		const size_t bsize = bh >> 3;
		if (type == 0) {
			if (p + bsize > len) {
				return ERR_TRUNCATED;
			}
			if (out + bsize > dst_size) {
				return ERR_DST_SIZE;
			}
			for (size_t i = 0; i < bsize; i++) {
				dst[out + i] = src[p + i];
			}
			out += bsize;
			p += bsize;
		} else if (type == 1) {
			if (p + 1 > len) {
				return ERR_TRUNCATED;
			}
			if (out + bsize > dst_size) {
				return ERR_DST_SIZE;
			}
			const uint8_t b = src[p];
			for (size_t i = 0; i < bsize; i++) {
				dst[out + i] = b;
			}
			out += bsize;
			p += 1;
		} else if (type == 2) {
			if (p + bsize > len || bsize > static_cast<size_t>(kMaxBlock)) {
				return ERR_TRUNCATED;
			}
			int rc = decode_compressed_block(w, src + p, bsize, dst, dst_size, &out);
			if (rc) {
				return rc;
			}
			p += bsize;
		} else {
			return ERR_CORRUPT;
		}
		if (last) {
			break;
		}
	}
	if (checksum) {
		if (p + 4 > len) {
			return ERR_TRUNCATED;
		}
		p += 4;
	}
	if (out != dst_size) {
		return ERR_DST_SIZE;
	}
	return p == len ? OK : ERR_UNSUPPORTED; // trailing bytes = another frame
}

} // namespace zstd
} // namespace b2a
