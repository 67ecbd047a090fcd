// Single-frame Zstandard decoder (RFC 8878) written for the device: one decoder instance per IPC buffer, driven by one
// lane, no recursion, no dynamic allocation, all state in a caller-provided workspace.  The reference decompresses each
// compressed IPC buffer with ZSTD_decompress (src/ipc/stream_reader/base_stream_reader.cpp:11-33, wired into nanoarrow's
// decoder at :38-50); a compressed IPC buffer is `int64 uncompressed_length` followed by exactly one zstd frame.
//
// The code is `__host__ __device__` so that the CPU test-suite can run the very same decoder against frames produced by
// pyarrow's (real) libzstd; the product only ever calls it from b2a_decompress_kernel.
//
// Execution model on the device: ALL 32 lanes of a warp run the decoder with identical control flow (every branch
// depends only on bytes all lanes read alike), so the serial parts — header parsing, table construction, the FSE state
// machine of the sequences — cost what one lane would cost, while the data-parallel parts split by lane: the four
// Huffman streams of a literals section go to four lanes, and literal / match / raw / RLE copies are strided over the
// warp (an overlapping match is periodic in its offset, so every byte can be fetched from data that already existed).
// On the host the same code runs with lane 0 of 1.
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

struct Tables;
struct Ctx {
	int lane;   // 0 .. nlanes-1
	int nlanes; // 32 on the device, 1 on the host
	Tables *t;  // entropy tables: shared memory on the device (one set per warp), plain memory on the host
};

ZSTD_HD inline void lanes_sync(const Ctx &) {
#if defined(__CUDA_ARCH__)
	__syncwarp();
#endif
}
// the maximum of a per-lane status over the warp (keeps control flow uniform after a lane-parallel section)
ZSTD_HD inline int lanes_max(const Ctx &, int v) {
#if defined(__CUDA_ARCH__)
	return static_cast<int>(__reduce_max_sync(0xFFFFFFFFu, static_cast<unsigned>(v)));
#else
	return v;
#endif
}

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

static constexpr int kChunk = 32; // sequences decoded (serially) before their copies are issued (in parallel)

struct Tables {
	uint16_t huf[1 << kHufMaxBits]; // symbol | nbits << 8
	FseTable ll, of, ml;
	// the current chunk of decoded sequences: lengths, literal / output start within the chunk, match offset
	uint32_t a_ll[kChunk], a_ml[kChunk], a_lp[kChunk], a_o[kChunk];
	uint64_t a_off[kChunk];
};

struct Workspace {
	int huf_max_bits; // 0: no table yet
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

// up to 57 bits starting at absolute bit offset `off` of src (little-endian bit order).
// Host: bytes past `len` read as zero.  Device: two aligned 8-byte loads + funnel shift; may touch up to 15 bytes around
// [src, src+len) — always inside the device image (the frame follows an 8-byte prefix, the image ends with 64 spare bytes).
ZSTD_HD inline uint64_t read_bits_le(const uint8_t *src, size_t len, int nbits, int64_t off) {
	if (nbits == 0) {
		return 0;
	}
#if defined(__CUDA_ARCH__)
	const uintptr_t p = reinterpret_cast<uintptr_t>(src) + static_cast<uintptr_t>(off >> 3);
	const uint64_t *a = reinterpret_cast<const uint64_t *>(p & ~uintptr_t(7));
	const int sh = static_cast<int>(p & 7) * 8 + static_cast<int>(off & 7);
	const uint64_t w0 = a[0], w1 = a[1];
	const uint64_t v = sh ? ((w0 >> sh) | (w1 << (64 - sh))) : w0;
	return v & ((1ull << nbits) - 1);
#else
	const size_t byte = static_cast<size_t>(off >> 3);
	const int shift = static_cast<int>(off & 7);
	uint64_t v = 0;
	const int nbytes = (nbits + shift + 7) >> 3;
	for (int i = 0; i < nbytes; i++) {
		const size_t q = byte + static_cast<size_t>(i);
		v |= static_cast<uint64_t>(q < len ? src[q] : 0) << (8 * i);
	}
	return (v >> shift) & ((1ull << nbits) - 1);
#endif
}

// 64 bits starting at bit position `bit` (>= 0) of src.  Host: bytes past `len` read as zero.  Device: two aligned
// 8-byte loads + funnel shift (same over-read bound as read_bits_le).
ZSTD_HD inline uint64_t load64_at(const uint8_t *src, size_t len, int64_t bit) {
#if defined(__CUDA_ARCH__)
	const uintptr_t p = reinterpret_cast<uintptr_t>(src) + static_cast<uintptr_t>(bit >> 3);
	const uint64_t *a = reinterpret_cast<const uint64_t *>(p & ~uintptr_t(7));
	const int sh = static_cast<int>(p & 7) * 8 + static_cast<int>(bit & 7);
	const uint64_t w0 = a[0], w1 = a[1];
	return sh ? ((w0 >> sh) | (w1 << (64 - sh))) : w0;
#else
	const size_t byte = static_cast<size_t>(bit >> 3);
	const int sh = static_cast<int>(bit & 7);
	uint64_t lo = 0;
	for (int i = 0; i < 8; i++) {
		const size_t q = byte + static_cast<size_t>(i);
		lo |= static_cast<uint64_t>(q < len ? src[q] : 0) << (8 * i);
	}
	const uint64_t ext = byte + 8 < len ? src[byte + 8] : 0;
	return sh ? ((lo >> sh) | (ext << (64 - sh))) : lo;
#endif
}

// Backward bitstream with a 64-bit register window.  `off` is the bit position just above the window's first bit at the
// last refill, `used` the bits consumed since; the window holds stream bits [off-64, off) with bit off-1 on top, bits
// below position 0 read as zero.  The current position is off - used.
struct BackBits {
	const uint8_t *src;
	size_t len;
	int64_t off;
	uint64_t win;
	int used;
};

ZSTD_HD inline void backbits_refill(BackBits &b) {
	b.off -= b.used;
	b.used = 0;
	if (b.off >= 64) {
		b.win = load64_at(b.src, b.len, b.off - 64);
	} else if (b.off > 0) {
		b.win = load64_at(b.src, b.len, 0) << (64 - b.off);
	} else {
		b.win = 0;
	}
}

ZSTD_HD inline int64_t backbits_pos(const BackBits &b) {
	return b.off - b.used;
}

ZSTD_HD inline int backbits_init(BackBits &b, const uint8_t *src, size_t len) {
	if (len == 0 || src[len - 1] == 0) {
		return ERR_CORRUPT;
	}
	b.src = src;
	b.len = len;
	b.off = static_cast<int64_t>(len) * 8 - (8 - highbit(src[len - 1]));
	b.used = 0;
	backbits_refill(b);
	return OK;
}

// unchecked: the caller guarantees used + nbits <= 64 (nbits <= 32)
ZSTD_HD inline uint32_t backbits_take(BackBits &b, int nbits) {
	const uint64_t v = (b.win << b.used) >> 1 >> (63 - nbits); // nbits == 0 -> 0 without a shift by 64
	b.used += nbits;
	return static_cast<uint32_t>(v);
}

ZSTD_HD inline uint64_t backbits_read(BackBits &b, int nbits) {
	if (b.used + nbits > 64) {
		backbits_refill(b);
	}
	return backbits_take(b, nbits);
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
ZSTD_HD inline int huf_build(const Ctx &cx, Workspace &w, const uint8_t *weights, int n_weights) {
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
			cx.t->huf[k] = static_cast<uint16_t>(i << 8);
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
			cx.t->huf[code + k] = static_cast<uint16_t>(cx.t->huf[code + k] | i);
		}
		rank_idx[bits[i]] += len;
	}
	w.huf_max_bits = max_bits;
	return OK;
}

ZSTD_HD inline int huf_read_tree(const Ctx &cx, Workspace &w, const uint8_t *src, size_t len, size_t *used) {
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
			if (backbits_pos(bb) < 0) {
				weights[n++] = w.tmp.e[s2].symbol;
				break;
			}
			weights[n++] = w.tmp.e[s2].symbol;
			{
				const FseEntry e = w.tmp.e[s2];
				s2 = e.base + static_cast<uint32_t>(backbits_read(bb, e.nbits));
			}
			if (backbits_pos(bb) < 0) {
				weights[n++] = w.tmp.e[s1].symbol;
				break;
			}
		}
		*used = 1 + csize;
	}
	return huf_build(cx, w, weights, n);
}

ZSTD_HD inline int huf_decode_stream(const Ctx &cx, const Workspace &w, const uint8_t *src, size_t len, uint8_t *dst,
                                     size_t n_out) {
	BackBits bb;
	int rc = backbits_init(bb, src, len);
	if (rc) {
		return rc;
	}
	const int mb = w.huf_max_bits;
	const uint32_t mask = (1u << mb) - 1;
	const uint16_t *tab = cx.t->huf;
	uint32_t state = backbits_take(bb, mb);
	for (size_t i = 0; i < n_out; i++) {
		if (bb.used > 64 - kHufMaxBits) {
			backbits_refill(bb);
			if (bb.off < -mb) {
				return ERR_CORRUPT;
			}
		}
		const uint32_t e = tab[state];
		dst[i] = static_cast<uint8_t>(e);
		const int nb = static_cast<int>(e >> 8);
		state = ((state << nb) | backbits_take(bb, nb)) & mask;
	}
	if (backbits_pos(bb) != -mb) {
		return ERR_CORRUPT;
	}
	return OK;
}

// ------------------------------------------------------------------------------------------------ sequences tables
// literal-length / match-length codes: baseline | extra bits << 24
#define ZSTD_LL_TAB                                                                                                    \
	{0,          1,          2,          3,          4,          5,          6,          7,          8,                \
	 9,          10,         11,         12,         13,         14,         15,         16 | 1 << 24, 18 | 1 << 24,    \
	 20 | 1 << 24, 22 | 1 << 24, 24 | 2 << 24, 28 | 2 << 24, 32 | 3 << 24, 40 | 3 << 24, 48 | 4 << 24, 64 | 6 << 24,    \
	 128 | 7 << 24, 256 | 8 << 24, 512 | 9 << 24, 1024 | 10 << 24, 2048 | 11 << 24, 4096 | 12 << 24, 8192 | 13 << 24,   \
	 16384 | 14 << 24, 32768 | 15 << 24, 65536 | 16 << 24}
#define ZSTD_ML_TAB                                                                                                    \
	{3,  4,  5,  6,  7,  8,  9,  10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 27, 28, 29, 30, 31,  \
	 32, 33, 34, 35 | 1 << 24, 37 | 1 << 24, 39 | 1 << 24, 41 | 1 << 24, 43 | 2 << 24, 47 | 2 << 24, 51 | 3 << 24,        \
	 59 | 3 << 24, 67 | 4 << 24, 83 | 4 << 24, 99 | 5 << 24, 131 | 7 << 24, 259 | 8 << 24, 515 | 9 << 24,                 \
	 1027 | 10 << 24, 2051 | 11 << 24, 4099 | 12 << 24, 8195 | 13 << 24, 16387 | 14 << 24, 32771 | 15 << 24,              \
	 65539 | 16 << 24}
static const uint32_t h_ll_tab[36] = ZSTD_LL_TAB;
static const uint32_t h_ml_tab[53] = ZSTD_ML_TAB;
#if defined(__CUDACC__)
static __device__ const uint32_t d_ll_tab[36] = ZSTD_LL_TAB;
static __device__ const uint32_t d_ml_tab[53] = ZSTD_ML_TAB;
#endif

ZSTD_HD inline uint32_t ll_code(int code) {
#if defined(__CUDA_ARCH__)
	return d_ll_tab[code];
#else
	return h_ll_tab[code];
#endif
}
ZSTD_HD inline uint32_t ml_code(int code) {
#if defined(__CUDA_ARCH__)
	return d_ml_tab[code];
#else
	return h_ml_tab[code];
#endif
}

// one 32-bit load of an FSE entry: base | symbol << 16 | nbits << 24 (the struct's little-endian layout)
ZSTD_HD inline uint32_t fse_entry(const FseTable &t, uint32_t state) {
	return *reinterpret_cast<const uint32_t *>(&t.e[state]);
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

ZSTD_HD inline int decode_literals(const Ctx &cx, Workspace &w, const uint8_t *src, size_t len, Literals &lit,
                                   size_t *used) {
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
		int rc = huf_read_tree(cx, w, p, left, &tree);
		if (rc) {
			return rc;
		}
		p += tree;
		left -= tree;
	} else if (w.huf_max_bits == 0) {
		return ERR_CORRUPT;
	}
	if (streams == 1) {
		int rc = huf_decode_stream(cx, w, p, left, w.literals, regen);
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
		int rc = OK;
		if (cx.nlanes >= 4) {
			// one stream per lane (lanes 0..3); the others wait at the sync below
			lanes_sync(cx); // table stores of every lane are visible to the decoding lanes
			if (cx.lane < 4) {
				const size_t so[4] = {0, s1, s1 + s2, s1 + s2 + s3}, sl[4] = {s1, s2, s3, s4};
				rc = huf_decode_stream(cx, w, q + so[cx.lane], sl[cx.lane], w.literals + static_cast<size_t>(cx.lane) * seg,
				                       cx.lane < 3 ? seg : regen - 3 * seg);
			}
			rc = lanes_max(cx, rc);
		} else {
			rc = huf_decode_stream(cx, w, q, s1, w.literals, seg);
			rc = rc ? rc : huf_decode_stream(cx, w, q + s1, s2, w.literals + seg, seg);
			rc = rc ? rc : huf_decode_stream(cx, w, q + s1 + s2, s3, w.literals + 2 * seg, seg);
			rc = rc ? rc : huf_decode_stream(cx, w, q + s1 + s2 + s3, s4, w.literals + 3 * seg, regen - 3 * seg);
		}
		if (rc) {
			return rc;
		}
	}
	lanes_sync(cx); // the literal buffer is complete before any lane copies from it
	lit.ptr = w.literals;
	lit.size = regen;
	*used = hdr + comp;
	return OK;
}

// dst[0 .. n) = n literals starting at `from`
ZSTD_HD inline void copy_literals(const Ctx &cx, const Literals &lit, size_t from, uint8_t *dst, uint32_t n) {
	if (lit.rle) {
		for (uint32_t i = static_cast<uint32_t>(cx.lane); i < n; i += static_cast<uint32_t>(cx.nlanes)) {
			dst[i] = lit.rle_byte;
		}
	} else {
		const uint8_t *s = lit.ptr + from;
		const uint32_t step = static_cast<uint32_t>(cx.nlanes);
		uint32_t i = static_cast<uint32_t>(cx.lane);
		for (; i + 3 * step < n; i += 4 * step) {
			const uint8_t v0 = s[i], v1 = s[i + step], v2 = s[i + 2 * step], v3 = s[i + 3 * step];
			dst[i] = v0;
			dst[i + step] = v1;
			dst[i + 2 * step] = v2;
			dst[i + 3 * step] = v3;
		}
		for (; i < n; i += step) {
			dst[i] = s[i];
		}
	}
}

// d[0..n) = s[0..n), 8 loads in flight before the first store; the ranges must not overlap within 8 bytes
ZSTD_HD inline void copy_batched(const uint8_t *s, uint8_t *d, uint32_t n) {
	uint32_t t = 0;
	for (; t + 8 <= n; t += 8) {
		uint8_t v[8];
		for (int u = 0; u < 8; u++) {
			v[u] = s[t + u];
		}
		for (int u = 0; u < 8; u++) {
			d[t + u] = v[u];
		}
	}
	uint8_t v[8];
	for (int u = 0; u < 8; u++) {
		v[u] = t + u < n ? s[t + u] : 0;
	}
	for (int u = 0; u < 8; u++) {
		if (t + u < n) {
			d[t + u] = v[u];
		}
	}
}

// d[k] = s[k mod period] (an overlapping match with offset = period < 8): the period's bytes are loaded once
ZSTD_HD inline void copy_periodic(const uint8_t *s, uint8_t *d, uint32_t n, uint32_t period) {
	uint8_t v[8];
	for (uint32_t u = 0; u < 8; u++) {
		v[u] = u < period ? s[u] : 0;
	}
	uint32_t r = 0;
	for (uint32_t k = 0; k < n; k++) {
		uint8_t x = v[0];
		for (uint32_t u = 1; u < 8; u++) {
			x = r == u ? v[u] : x;
		}
		d[k] = x;
		r = r + 1 == period ? 0 : r + 1;
	}
}

// one match strided over the warp (host: one lane); handles self-overlap through the periodic form
ZSTD_HD inline void copy_match_strided(const Ctx &cx, const uint8_t *m, uint8_t *d, uint32_t ml, size_t off) {
	const uint32_t lane = static_cast<uint32_t>(cx.lane), step = static_cast<uint32_t>(cx.nlanes);
	if (off >= ml) {
		uint32_t k = lane;
		for (; k + 3 * step < ml; k += 4 * step) {
			const uint8_t v0 = m[k], v1 = m[k + step], v2 = m[k + 2 * step], v3 = m[k + 3 * step];
			d[k] = v0;
			d[k + step] = v1;
			d[k + 2 * step] = v2;
			d[k + 3 * step] = v3;
		}
		for (; k < ml; k += step) {
			d[k] = m[k];
		}
	} else if (off == 1) {
		const uint8_t v = m[0];
		for (uint32_t k = lane; k < ml; k += step) {
			d[k] = v;
		}
	} else {
		const uint32_t off32 = static_cast<uint32_t>(off); // off < ml <= 128 KiB + 3
		for (uint32_t k = lane; k < ml; k += step) {
			d[k] = m[k % off32];
		}
	}
}

// decodes one compressed block into dst[pos..); returns the new position through *pos
ZSTD_HD inline int decode_compressed_block(const Ctx &cx, Workspace &w, const uint8_t *src, size_t len, uint8_t *dst,
                                           size_t cap, size_t *pos) {
	Literals lit;
	size_t used = 0;
	int rc = decode_literals(cx, w, src, len, lit, &used);
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
		rc = seq_table(cx.t->ll, w.have_ll, (modes >> 6) & 3, 0, p, left, &u);
		if (rc) {
			return rc;
		}
		p += u;
		left -= u;
		rc = seq_table(cx.t->of, w.have_of, (modes >> 4) & 3, 1, p, left, &u);
		if (rc) {
			return rc;
		}
		p += u;
		left -= u;
		rc = seq_table(cx.t->ml, w.have_ml, (modes >> 2) & 3, 2, p, left, &u);
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
		const FseTable &tll = cx.t->ll, &tof = cx.t->of, &tml = cx.t->ml;
		uint32_t ll_state = static_cast<uint32_t>(backbits_read(bb, tll.accuracy));
		uint32_t of_state = static_cast<uint32_t>(backbits_read(bb, tof.accuracy));
		uint32_t ml_state = static_cast<uint32_t>(backbits_read(bb, tml.accuracy));
		uint64_t rep0 = w.rep[0], rep1 = w.rep[1], rep2 = w.rep[2]; // repeat offsets live in registers inside a block
		constexpr uint32_t kLongRun = 64;
		// per-chunk sequence arrays live with the tables (shared memory on the device: dynamically indexed local arrays
		// would sit in local memory, and every lane stores the same values anyway)
		uint32_t *a_ll = cx.t->a_ll, *a_ml = cx.t->a_ml, *a_lp = cx.t->a_lp, *a_o = cx.t->a_o;
		uint64_t *a_off = cx.t->a_off;
		for (size_t i0 = 0; i0 < nseq; i0 += kChunk) {
			const int cnt = static_cast<int>(nseq - i0 < static_cast<size_t>(kChunk) ? nseq - i0 : kChunk);
			// ---- phase 1: the FSE state machine (serial; every lane runs it alike) for up to 32 sequences.
			// Two window refills per sequence: offset + match + literal extra bits <= 31+16+16, state bits <= 9+9+8.
			size_t o = out, lp = lit_pos;
			for (int j = 0; j < cnt; j++) {
				const uint32_t eo = fse_entry(tof, of_state), el = fse_entry(tll, ll_state), em = fse_entry(tml, ml_state);
				const int of_code = static_cast<int>((eo >> 16) & 0xFF), llc = static_cast<int>((el >> 16) & 0xFF),
				          mlc = static_cast<int>((em >> 16) & 0xFF);
				if (of_code > 31 || llc > 35 || mlc > 52) {
					return ERR_CORRUPT;
				}
				const uint32_t lt = ll_code(llc), mt = ml_code(mlc);
				backbits_refill(bb);
				const uint64_t offset_value = (1ull << of_code) + backbits_take(bb, of_code);
				const uint32_t match_len = (mt & 0xFFFFFFu) + backbits_take(bb, static_cast<int>(mt >> 24));
				const uint32_t lit_len = (lt & 0xFFFFFFu) + backbits_take(bb, static_cast<int>(lt >> 24));
				if (i0 + static_cast<size_t>(j) + 1 < nseq) {
					if (bb.used > 64 - 26) {
						backbits_refill(bb);
					}
					ll_state = (el & 0xFFFFu) + backbits_take(bb, static_cast<int>(el >> 24));
					ml_state = (em & 0xFFFFu) + backbits_take(bb, static_cast<int>(em >> 24));
					of_state = (eo & 0xFFFFu) + backbits_take(bb, static_cast<int>(eo >> 24));
				}
				if (backbits_pos(bb) < 0) {
					return ERR_CORRUPT;
				}
				uint64_t offset;
				if (offset_value > 3) {
					offset = offset_value - 3;
					rep2 = rep1;
					rep1 = rep0;
					rep0 = offset;
				} else {
					const uint32_t idx = static_cast<uint32_t>(offset_value) - 1 + (lit_len == 0 ? 1 : 0);
					if (idx == 0) {
						offset = rep0;
					} else {
						offset = idx == 1 ? rep1 : (idx == 2 ? rep2 : rep0 - 1);
						if (idx > 1) {
							rep2 = rep1;
						}
						rep1 = rep0;
						rep0 = offset;
					}
				}
				a_lp[j] = static_cast<uint32_t>(lp - lit_pos);
				a_o[j] = static_cast<uint32_t>(o - out);
				lp += lit_len;
				o += lit_len;
				if (lp > lit.size || o + match_len > cap || offset == 0 || offset > o) {
					return ERR_CORRUPT;
				}
				o += match_len;
				a_ll[j] = lit_len;
				a_ml[j] = match_len;
				a_off[j] = offset;
			}
			// ---- phase 2: the literal runs of the chunk — independent of each other and of the output.  Short runs go
			// one sequence per lane (loads of all lanes in flight together, batches of 8 bytes: a warp issues in order,
			// so a copy loop that stores each byte right after loading it pays one memory latency per byte), long runs
			// are strided over the warp.
			for (int j = cx.lane; j < cnt; j += cx.nlanes) {
				const uint32_t n = a_ll[j];
				if (n == 0 || n > kLongRun) {
					continue;
				}
				uint8_t *d = dst + out + a_o[j];
				if (lit.rle) {
					for (uint32_t t = 0; t < n; t++) {
						d[t] = lit.rle_byte;
					}
				} else {
					copy_batched(lit.ptr + lit_pos + a_lp[j], d, n);
				}
			}
			for (int j = 0; j < cnt; j++) {
				if (a_ll[j] > kLongRun) {
					copy_literals(cx, lit, lit_pos + a_lp[j], dst + out + a_o[j], a_ll[j]);
				}
			}
			lit_pos = lp;
			lanes_sync(cx);
			// ---- phase 3: matches.  dst[m + k] = dst[m + k - offset] in order of k; an overlapping match (offset <
			// length) is periodic in `offset`, so byte k equals byte (k mod offset) of the window before it.
			// (a) matches whose source lies entirely before this chunk depend on nothing written here: lane per
			//     sequence (short) or warp-strided (long), all in flight together;
			// (b) the others may read what earlier sequences of the chunk wrote: in order, one sync each.
			bool any_dep = false;
			for (int j = cx.lane; j < cnt; j += cx.nlanes) {
				const uint32_t ml = a_ml[j];
				const size_t m0 = out + a_o[j] + a_ll[j], off = static_cast<size_t>(a_off[j]);
				const size_t span = ml < off ? ml : off;
				if (m0 - off + span > out || ml > kLongRun) {
					continue;
				}
				if (off >= 8) {
					copy_batched(dst + m0 - off, dst + m0, ml); // loads of one batch never reach its own stores
				} else {
					copy_periodic(dst + m0 - off, dst + m0, ml, static_cast<uint32_t>(off));
				}
			}
			for (int j = 0; j < cnt; j++) {
				const uint32_t ml = a_ml[j];
				const size_t m0 = out + a_o[j] + a_ll[j], off = static_cast<size_t>(a_off[j]);
				const size_t span = ml < off ? ml : off;
				if (m0 - off + span > out) {
					any_dep = true;
				} else if (ml > kLongRun) {
					copy_match_strided(cx, dst + m0 - off, dst + m0, ml, off);
				}
			}
			lanes_sync(cx);
			if (any_dep) {
				for (int j = 0; j < cnt; j++) {
					const uint32_t ml = a_ml[j];
					const size_t m0 = out + a_o[j] + a_ll[j], off = static_cast<size_t>(a_off[j]);
					const size_t span = ml < off ? ml : off;
					if (m0 - off + span > out) {
						copy_match_strided(cx, dst + m0 - off, dst + m0, ml, off);
						lanes_sync(cx);
					}
				}
			}
			out = o;
		}
		w.rep[0] = rep0;
		w.rep[1] = rep1;
		w.rep[2] = rep2;
		if (backbits_pos(bb) != 0) {
			return ERR_CORRUPT;
		}
	}
	const size_t rest = lit.size - lit_pos;
	if (out + rest > cap) {
		return ERR_CORRUPT;
	}
	copy_literals(cx, lit, lit_pos, dst + out, static_cast<uint32_t>(rest));
	out += rest;
	lanes_sync(cx);
	*pos = out;
	return OK;
}

// ------------------------------------------------------------------------------------------------ frame
ZSTD_HD inline int decode_frame(const Ctx &cx, Workspace &w, const uint8_t *src, size_t len, uint8_t *dst,
                                size_t dst_size) {
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
			for (size_t i = static_cast<size_t>(cx.lane); i < bsize; i += static_cast<size_t>(cx.nlanes)) {
				dst[out + i] = src[p + i];
			}
			lanes_sync(cx);
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
			for (size_t i = static_cast<size_t>(cx.lane); i < bsize; i += static_cast<size_t>(cx.nlanes)) {
				dst[out + i] = b;
			}
			lanes_sync(cx);
			out += bsize;
			p += 1;
		} else if (type == 2) {
			if (p + bsize > len || bsize > static_cast<size_t>(kMaxBlock)) {
				return ERR_TRUNCATED;
			}
			int rc = decode_compressed_block(cx, w, src + p, bsize, dst, dst_size, &out);
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
