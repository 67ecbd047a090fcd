#include "scan_plan.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace b2a {

static const uint64_t kNone = UINT64_MAX;
// bytes moved (in + out) per tile, roughly; B2A_TILE_BYTES overrides for tuning experiments
static uint64_t TileBytes() {
	static uint64_t v = 0;
	if (v == 0) {
		const char *e = getenv("B2A_TILE_BYTES");
		v = e ? strtoull(e, nullptr, 10) : 0;
		if (v < 4096) {
			v = 256 * 1024;
		}
	}
	return v;
}

uint64_t ScanPlan::Alloc(uint64_t bytes) {
	uint64_t off = out_bytes;
	out_bytes += (bytes + 255) & ~uint64_t(255);
	return off;
}

// Buffers of a compressed batch are addressed through VIRTUAL offsets while the plan is built:
// bit 62 | candidate index << 36 | offset inside the uncompressed buffer; Resolve() maps them into the scratch region.
static const uint64_t kVirtBit = uint64_t(1) << 62;
static const int kVirtShift = 36;

RecordBatch ScanPlan::VirtualBatch(const Reader &reader, const RecordBatch &rb) {
	if (rb.compression != 0 && rb.compression != 1) {
		throw Error(B2A_ENOTIMPL, "Arrow IPC: unknown body compression codec " + std::to_string(rb.compression));
	}
	// 1 = ZSTD (the codec the reference installs, base_stream_reader.cpp:38-50); 0 = LZ4_FRAME (the
	// reference fails on it for want of a decompressor; here it is decoded by kernels/lz4_decode.cuh)
	const uint32_t frame_codec = rb.compression == 1 ? B2A_CODEC_ZSTD : B2A_CODEC_LZ4;
	if (alias_direct) {
		throw Error(B2A_EINVAL, "a B2A_SCAN_HOST_ALIAS / B2A_SCAN_DEVICE_ALIAS scan cannot alias compressed bodies: create "
		                        "the scan without it");
	}
	RecordBatch v = rb;
	const uint8_t *image = reader.Image();
	for (auto &b : v.buffers) {
		if (b.length == 0) {
			continue;
		}
		if (b.length < 8 || b.offset + b.length > reader.ImageSize()) {
			throw Error(B2A_EIO, "Arrow IPC: compressed buffer is shorter than its 8-byte length prefix");
		}
		int64_t ulen;
		std::memcpy(&ulen, image + b.offset, 8);
		if (big_endian) {
			ulen = static_cast<int64_t>(__builtin_bswap64(static_cast<uint64_t>(ulen)));
		}
		DecompCand c;
		c.src_off = b.offset + 8;
		c.src_len = b.length - 8;
		c.job = -1;
		if (ulen == -1) {
			c.codec = B2A_CODEC_RAW;
			c.dst_len = c.src_len;
		} else if (ulen < 0) {
			throw Error(B2A_EIO, "Arrow IPC: invalid uncompressed length in a compressed buffer");
		} else {
			c.codec = frame_codec;
			c.dst_len = static_cast<uint64_t>(ulen);
		}
		if (c.dst_len >= (uint64_t(1) << kVirtShift) || cands.size() >= (uint64_t(1) << (62 - kVirtShift))) {
			throw Error(B2A_EINVAL, "compressed scan too large for one plan: split the batch range");
		}
		b.offset = kVirtBit | (static_cast<uint64_t>(cands.size()) << kVirtShift);
		b.length = c.dst_len;
		cands.push_back(c);
	}
	compressed = true;
	return v;
}

uint64_t ScanPlan::Resolve(uint64_t off) const {
	if (off == kNone || !(off & kVirtBit)) {
		return off;
	}
	const DecompCand &c = cands[(off & ~kVirtBit) >> kVirtShift];
	const uint64_t intra = off & ((uint64_t(1) << kVirtShift) - 1);
	if (c.job < 0) {
		return scratch_base; // never read (an untouched / empty buffer)
	}
	return jobs[static_cast<size_t>(c.job)].dst + intra;
}

void ScanPlan::Touch(uint64_t off, uint64_t len) {
	if (len == 0) {
		return;
	}
	if (off & kVirtBit) {
		DecompCand &c = cands[(off & ~kVirtBit) >> kVirtShift];
		if (c.job >= 0) {
			return;
		}
		c.job = static_cast<int64_t>(jobs.size());
		b2a_decomp_job j;
		std::memset(&j, 0, sizeof(j));
		j.src = c.src_off;
		j.src_len = c.src_len;
		j.dst = scratch_bytes; // relative to scratch_base until Build() finishes
		j.dst_len = c.dst_len;
		j.codec = c.codec;
		scratch_bytes += (c.dst_len + 63) & ~uint64_t(63);
		jobs.push_back(j);
		off = c.src_off;
		len = c.src_len;
		if (len == 0) {
			return;
		}
	}
	image_lo = std::min(image_lo, off);
	image_hi = std::max(image_hi, off + len);
	algorithmic_in += len;
	if (cur_ranges) {
		cur_ranges->emplace_back(off, off + len);
	}
}

// sort + merge ranges whose gap is small (one bigger copy beats two small ones); ends rounded to 16 bytes
static void MergeRanges(ScanPlan::RangeList &r, uint64_t limit_hi) {
	static const uint64_t kGap = 256 * 1024;
	std::sort(r.begin(), r.end());
	ScanPlan::RangeList out;
	for (auto &x : r) {
		uint64_t lo = x.first & ~uint64_t(15), hi = std::min(limit_hi, (x.second + 15) & ~uint64_t(15));
		if (!out.empty() && lo <= out.back().second + kGap) {
			out.back().second = std::max(out.back().second, hi);
		} else {
			out.emplace_back(lo, hi);
		}
	}
	r.swap(out);
}

// x16: input bytes per row in 1/16 byte units (bitmaps are 2/16 = 1/8 byte per row)
void ScanPlan::AddTask(b2a_task t, uint64_t in_x16, uint64_t out_per_row) {
	uint64_t rows = t.rows + ((t.op == B2A_OP_VALIDITY && (t.flags & B2A_TASK_F_ADD_NULL)) ? 1 : 0);
	if (rows == 0) {
		return;
	}
	uint64_t per_row_x16 = in_x16 + out_per_row * 16;
	if (per_row_x16 == 0) {
		per_row_x16 = 1;
	}
	uint64_t tile_rows = TileBytes() * 16 / per_row_x16;
	tile_rows = std::max<uint64_t>(B2A_VECTOR_SIZE, (tile_rows / B2A_VECTOR_SIZE) * B2A_VECTOR_SIZE);
	tile_rows = std::min<uint64_t>(tile_rows, uint64_t(1) << 22);
	uint64_t n_tiles = (rows + tile_rows - 1) / tile_rows;
	if (uint64_t(next_tile) + n_tiles > 0xFFFFFFF0ull || tasks.size() > 0xFFFFFFF0ull) {
		throw Error(B2A_EINVAL, "scan too large for one launch: split the batch range");
	}
	t.tile0 = next_tile;
	t.tile_rows = static_cast<uint32_t>(tile_rows);
	uint32_t idx = static_cast<uint32_t>(tasks.size());
	tasks.push_back(t);
	tile_task.insert(tile_task.end(), n_tiles, idx);
	next_tile += static_cast<uint32_t>(n_tiles);
}

static int BatchBuffers(const Field &f, bool ignore_dictionary) {
	if (f.dict_id >= 0 && !ignore_dictionary) {
		return 2;
	}
	return f.n_buffers;
}

void ScanPlan::SkipField(const Field &f, const RecordBatch &rb, size_t &node_idx, size_t &buf_idx, size_t &var_idx,
                         bool ignore_dictionary) {
	node_idx++;
	buf_idx += static_cast<size_t>(BatchBuffers(f, ignore_dictionary));
	if (f.dict_id >= 0 && !ignore_dictionary) {
		return; // the value type's children are not part of the record batch
	}
	if (f.variadic) {
		if (var_idx >= rb.variadic_counts.size()) {
			throw Error(B2A_EIO, "Arrow IPC: missing variadicBufferCounts entry");
		}
		buf_idx += static_cast<size_t>(rb.variadic_counts[var_idx++]);
	}
	for (auto &c : f.children) {
		SkipField(*c, rb, node_idx, buf_idx, var_idx, false);
	}
}

static void NeedBytes(const BufferSlice &b, uint64_t need, const Field &f, const char *what) {
	if (b.length < need) {
		throw Error(B2A_EIO, "Arrow IPC: " + std::string(what) + " buffer of field '" + f.name + "' has " +
		                         std::to_string(b.length) + " bytes, expected at least " + std::to_string(need));
	}
}

// Below a list view only types whose conversion does not depend on where the conversion call starts can share one
// decoded child column: the chunks' child ranges [lowest offset, + sum of sizes) may overlap or leave gaps.
static void CheckListViewChild(const Field &f) {
	if (f.dict_id >= 0) {
		return;
	}
	switch (f.op) {
	case B2A_OP_LIST32:
	case B2A_OP_LIST64:
	case B2A_OP_LISTVIEW32:
	case B2A_OP_LISTVIEW64:
	case B2A_OP_BLOB_FIXED:
		throw Error(B2A_ENOTIMPL, "Arrow IPC: '" + f.format + "' below a list view is not supported by the B200 scan path");
	default:
		break;
	}
	for (auto &c : f.children) {
		CheckListViewChild(*c);
	}
}

void ScanPlan::EmitField(const Reader &reader, const Field &f, const RecordBatch &rb, size_t &node_idx, size_t &buf_idx,
                         size_t &var_idx, const Chain &parent_chain, bool list_child, bool selected,
                         b2a_column_out *slots, bool as_dictionary_values, bool ignore_dictionary) {
	if (!selected) {
		SkipField(f, rb, node_idx, buf_idx, var_idx, ignore_dictionary);
		return;
	}
	const bool is_dict = f.dict_id >= 0 && !ignore_dictionary;
	const int nb = BatchBuffers(f, ignore_dictionary);
	if (node_idx >= rb.nodes.size() || buf_idx + static_cast<size_t>(nb) > rb.buffers.size()) {
		throw Error(B2A_EIO, "Arrow IPC: RecordBatch has too few nodes/buffers for field '" + f.name + "'");
	}
	const NodeSlice node = rb.nodes[node_idx++];
	const BufferSlice *bufs = rb.buffers.data() + buf_idx;
	buf_idx += static_cast<size_t>(nb);
	const uint64_t L = static_cast<uint64_t>(node.length);

	b2a_column_out &slot = slots[f.flat_index];
	std::memset(&slot, 0xFF, sizeof(slot));
	slot.rows = L;
	slot.null_count = static_cast<uint64_t>(node.null_count);
	slot.chars_len = 0;

	// ---- run-end encoded (ColumnArrowToDuckDBRunEndEncoded, arrow_conversion.cpp:721-767): no buffers of its own; the
	// `values` child is decoded like an independent column of R rows (its own validity only — the reference passes
	// neither the parent mask nor the parent chain on), the `run_ends` child is read raw by the gather task, which runs in
	// the second launch because it reads the DECODED values (FlattenRunEnds, :579-663).
	if (f.op == B2A_OP_REE && !is_dict) {
		if (as_dictionary_values) {
			throw Error(B2A_ENOTIMPL, "run-end encoded dictionary values are not supported");
		}
		const Field &re = *f.children.at(0), &vals = *f.children.at(1);
		if (node_idx >= rb.nodes.size() || buf_idx + 2 > rb.buffers.size()) {
			throw Error(B2A_EIO, "Arrow IPC: RecordBatch has too few nodes/buffers for field '" + f.name + "'");
		}
		const NodeSlice rnode = rb.nodes[node_idx++];
		const BufferSlice *rbufs = rb.buffers.data() + buf_idx;
		buf_idx += 2;
		const uint64_t R = static_cast<uint64_t>(rnode.length), rw = static_cast<uint64_t>(re.out_width);
		NeedBytes(rbufs[1], R * rw, re, "run ends");
		Touch(rbufs[1].offset, R * rw);
		Chain empty;
		const size_t vnode = node_idx;
		const bool save_alias = alias_direct;
		alias_direct = false; // the gather reads the values from the arena even in a host-alias scan
		EmitField(reader, vals, rb, node_idx, buf_idx, var_idx, empty, true, true, slots, false, false);
		alias_direct = save_alias;
		if (static_cast<uint64_t>(rb.nodes[vnode].length) != R) {
			throw Error(B2A_EIO, "Arrow IPC: run_ends and values of field '" + f.name + "' differ in length");
		}
		const b2a_column_out vs = slots[vals.flat_index];
		const uint64_t w = static_cast<uint64_t>(f.out_width);
		b2a_task g;
		std::memset(&g, 0, sizeof(g));
		g.op = B2A_OP_REE;
		g.flags = (big_endian ? B2A_TASK_F_BSWAP : 0) | B2A_TASK_F_PHASE2;
		g.rows = L;
		g.src = rbufs[1].offset;
		g.src_width = static_cast<uint8_t>(f.src_width);
		g.dst_width = static_cast<uint8_t>(w);
		g.param = static_cast<int64_t>(R);
		g.src2 = vs.data_off;         // ARENA offset of the decoded values
		g.src2_len = vs.validity_off; // ARENA offset of their ValidityMask, or none
		g.dst = slot.data_off = Alloc(L * w);
		g.aux = kNone;
		if (vs.validity_off != kNone) {
			// the result mask is only written when the values can be NULL (value_format.validity.AllValid(), :603)
			g.aux = slot.validity_off = Alloc(((L + 63) / 64) * 8);
			algorithmic_out += ((L + 63) / 64) * 8;
		}
		if (vs.validity_off == kNone && !list_child && parent_chain.n > 0) {
			// a struct parent ANDs its NULLs into the child's mask BEFORE the child is converted (:1143-1151); FlattenRunEnds
			// leaves that mask alone when no value is NULL (and overwrites it otherwise — the branch above)
			b2a_task v;
			std::memset(&v, 0, sizeof(v));
			v.op = B2A_OP_VALIDITY;
			v.rows = L;
			v.flags = big_endian ? B2A_TASK_F_BSWAP : 0;
			v.n_masks = parent_chain.n;
			for (uint32_t i = 0; i < parent_chain.n; i++) {
				v.masks[i] = parent_chain.m[i];
			}
			v.dst = slot.validity_off = Alloc(((L + 63) / 64) * 8);
			v.aux = slot.stats_off = stats_bytes;
			stats_bytes += 8;
			algorithmic_out += ((L + 63) / 64) * 8;
			AddTask(v, 2 * parent_chain.n, 0);
		}
		slot.chars_off = vs.chars_off;
		slot.chars_len = vs.chars_len;
		algorithmic_out += L * w;
		phase2_pending.push_back(g);
		return;
	}

	// ---- sparse union (arrow_conversion.cpp:1173-1221).  One buffer (type ids).  Every member is converted like an
	// independent column (own validity only: unlike the STRUCT case the parent's NULLs are not ANDed into the members, and
	// the final SetValue loop overwrites the union's own mask row by row), then per row the member `tag` names is boxed
	// and stored back: the union row is valid iff that member's row is, the tag child holds the tag, member m is valid
	// where tag == m and its value is.  The gather task computes tags + all masks in the second launch, from the type
	// ids and the members' DECODED masks.
	if (f.op == B2A_OP_UNION && !is_dict) {
		if (as_dictionary_values) {
			throw Error(B2A_ENOTIMPL, "union dictionary values are not supported");
		}
		if (list_child) {
			// under a list the reference reads type_ids[0 .. child_size) while the members start at the list's offset
			throw Error(B2A_ENOTIMPL, "unions below a list are not supported by the B200 scan path");
		}
		NeedBytes(bufs[0], L, f, "type ids");
		Touch(bufs[0].offset, L);
		const size_t n = f.children.size();
		const uint64_t mask_bytes = ((L + 63) / 64) * 8;
		b2a_task g;
		std::memset(&g, 0, sizeof(g));
		g.op = B2A_OP_UNION;
		g.flags = B2A_TASK_F_PHASE2;
		g.rows = L;
		g.src = bufs[0].offset;
		g.src_width = 1;
		g.dst_width = 1;
		g.param = static_cast<int64_t>(n);
		g.dst = slot.data_off = Alloc(L);
		g.aux = slot.validity_off = Alloc(mask_bytes);
		algorithmic_out += L + mask_bytes;
		std::vector<uint64_t> raw(n), fin(n);
		Chain empty;
		const bool save_alias = alias_direct;
		alias_direct = false;
		for (size_t m = 0; m < n; m++) {
			const Field &c = *f.children[m];
			const size_t child_node = node_idx;
			EmitField(reader, c, rb, node_idx, buf_idx, var_idx, empty, true, true, slots, false, false);
			if (static_cast<uint64_t>(rb.nodes[child_node].length) < L) {
				throw Error(B2A_EIO, "Arrow IPC: union member '" + c.name + "' is shorter than its parent");
			}
			b2a_column_out &cs = slots[c.flat_index];
			raw[m] = cs.validity_off;
			fin[m] = cs.validity_off = Alloc(mask_bytes);
			cs.stats_off = kNone;
			algorithmic_out += mask_bytes;
		}
		alias_direct = save_alias;
		g.variadic0 = static_cast<uint32_t>(variadic.size());
		variadic.insert(variadic.end(), raw.begin(), raw.end());
		variadic.insert(variadic.end(), fin.begin(), fin.end());
		phase2_pending.push_back(g);
		return;
	}

	// ---- effective validity chain: own bitmap (only when null_count != 0 and the buffer exists,
	// arrow_conversion.cpp:71) AND the struct / fixed-size-list ancestors
	Chain chain = parent_chain;
	if (list_child) {
		chain.n = 0; // children of lists do not inherit validity (ColumnArrowToDuckDB is called without a parent mask)
	}
	if (nb > 0 && node.null_count != 0 && bufs[0].length > 0) {
		NeedBytes(bufs[0], (L + 7) / 8, f, "validity");
		if (chain.n >= B2A_MAX_MASKS) {
			throw Error(B2A_ENOTIMPL, "more than 4 nested nullable struct levels are not supported");
		}
		chain.m[chain.n++] = b2a_mask_ref {bufs[0].offset, 1};
		Touch(bufs[0].offset, (L + 7) / 8);
	}
	b2a_task base;
	std::memset(&base, 0, sizeof(base));
	base.aux = kNone;
	base.rows = L;
	base.flags = big_endian ? B2A_TASK_F_BSWAP : 0;
	base.n_masks = chain.n;
	for (uint32_t i = 0; i < chain.n; i++) {
		base.masks[i] = chain.m[i];
	}

	if (chain.n > 0 || as_dictionary_values) {
		b2a_task v = base;
		v.op = B2A_OP_VALIDITY;
		if (as_dictionary_values) {
			v.flags |= B2A_TASK_F_ADD_NULL;
		}
		uint64_t total = L + (as_dictionary_values ? 1 : 0);
		v.dst = Alloc(((total + 63) / 64) * 8);
		v.aux = slot.stats_off = stats_bytes;
		stats_bytes += 8;
		slot.validity_off = v.dst;
		algorithmic_out += ((total + 63) / 64) * 8;
		AddTask(v, 2 * chain.n, 0);
	}

	if (is_dict) {
		if (!f.children.empty()) {
			throw Error(B2A_ENOTIMPL, "dictionary-encoded nested types are not supported");
		}
		auto it = reader.Dictionaries().find(f.dict_id);
		if (it == reader.Dictionaries().end()) {
			throw Error(B2A_EIO, "Arrow IPC: no DictionaryBatch for dictionary id " + std::to_string(f.dict_id));
		}
		NeedBytes(bufs[1], L * static_cast<uint64_t>(f.index_width), f, "dictionary index");
		b2a_task t = base;
		t.op = B2A_OP_DICT_INDEX;
		t.src = bufs[1].offset;
		t.src_width = static_cast<uint8_t>(f.index_width);
		t.dst_width = 4;
		t.param = static_cast<int64_t>(it->second.data.rows);
		bool is_signed = f.index_type_id == B2A_T_TINYINT || f.index_type_id == B2A_T_SMALLINT ||
		                 f.index_type_id == B2A_T_INTEGER || f.index_type_id == B2A_T_BIGINT;
		if (is_signed) {
			t.flags |= B2A_TASK_F_SIGNED;
		}
		t.dst = slot.data_off = Alloc(L * 4);
		Touch(t.src, L * static_cast<uint64_t>(f.index_width));
		algorithmic_out += L * 4;
		AddTask(t, 16 * static_cast<uint64_t>(f.index_width), 4);
		dict_cols[f.flat_index]; // mark: dictionary values must be decoded
		return;
	}

	size_t n_variadic = 0;
	if (f.variadic) {
		if (var_idx >= rb.variadic_counts.size()) {
			throw Error(B2A_EIO, "Arrow IPC: missing variadicBufferCounts entry");
		}
		n_variadic = static_cast<size_t>(rb.variadic_counts[var_idx++]);
		if (buf_idx + n_variadic > rb.buffers.size()) {
			throw Error(B2A_EIO, "Arrow IPC: RecordBatch has too few variadic buffers");
		}
	}

	b2a_task t = base;
	t.op = static_cast<uint16_t>(f.op);
	t.src_width = static_cast<uint8_t>(f.src_width);
	t.dst_width = static_cast<uint8_t>(f.out_width);
	t.param = f.param;
	switch (f.op) {
	case B2A_OP_NONE:
		break;
	case B2A_OP_BOOL:
		NeedBytes(bufs[1], (L + 7) / 8, f, "data");
		t.src = bufs[1].offset;
		t.dst = slot.data_off = Alloc(L);
		Touch(t.src, (L + 7) / 8);
		algorithmic_out += L;
		AddTask(t, 2, 1);
		break;
	case B2A_OP_STRING32:
	case B2A_OP_STRING64: {
		const uint64_t ow = static_cast<uint64_t>(f.src_width);
		if (L > 0) {
			NeedBytes(bufs[1], (L + 1) * ow, f, "offsets");
		}
		t.src = bufs[1].offset;
		t.src2 = bufs[2].offset;
		t.src2_len = bufs[2].length;
		if (f.type_id == B2A_T_BLOB || f.type_id == B2A_T_BIT || f.type_id == B2A_T_VARINT) {
			t.flags |= B2A_TASK_F_BLOB;
		}
		t.dst = slot.data_off = Alloc(L * 16);
		slot.chars_off = bufs[2].offset;
		slot.chars_len = bufs[2].length;
		if (L > 0) {
			Touch(t.src, (L + 1) * ow);
			Touch(t.src2, t.src2_len);
			algorithmic_out += L * 16;
			AddTask(t, 16 * ow + 16 * t.src2_len / L, 16);
		}
		break;
	}
	case B2A_OP_STRING_VIEW: {
		NeedBytes(bufs[1], L * 16, f, "views");
		t.src = bufs[1].offset;
		t.variadic0 = static_cast<uint32_t>(variadic.size());
		t.param = static_cast<int64_t>(n_variadic);
		uint64_t data_bytes = 0;
		for (size_t i = 0; i < n_variadic; i++) {
			const BufferSlice &vb = rb.buffers[buf_idx + i];
			variadic.push_back(vb.offset); // (offset, length) pairs: the kernel checks offset + len of every long view
			variadic.push_back(vb.length);
			Touch(vb.offset, vb.length);
			data_bytes += vb.length;
		}
		t.dst = slot.data_off = Alloc(L * 16);
		slot.chars_off = 0; // pointers are relative to the stream image itself
		slot.chars_len = data_bytes;
		Touch(t.src, L * 16);
		algorithmic_out += L * 16;
		AddTask(t, 16 * 16 + (L ? 16 * data_bytes / L : 0), 16);
		break;
	}
	case B2A_OP_BLOB_FIXED:
		NeedBytes(bufs[1], L * static_cast<uint64_t>(f.param), f, "data");
		t.src = bufs[1].offset;
		t.dst = slot.data_off = Alloc(L * 16);
		if (!chain.levels.empty()) { // conversion-call boundaries under lists (see op_blob_fixed)
			t.variadic0 = static_cast<uint32_t>(variadic.size());
			t.src2_len = chain.levels.size();
			variadic.push_back(chain.top_rows);
			variadic.push_back(chain.levels.size());
			for (auto &lv : chain.levels) {
				variadic.push_back(lv.first);
				variadic.push_back(lv.second);
			}
		}
		slot.chars_off = bufs[1].offset;
		slot.chars_len = bufs[1].length;
		Touch(t.src, L * static_cast<uint64_t>(f.param));
		algorithmic_out += L * 16;
		AddTask(t, 16 * static_cast<uint64_t>(f.param), 16);
		break;
	case B2A_OP_LIST32:
	case B2A_OP_LIST64: {
		const uint64_t ow = static_cast<uint64_t>(f.src_width);
		if (L > 0) {
			NeedBytes(bufs[1], (L + 1) * ow, f, "offsets");
		}
		t.src = bufs[1].offset;
		t.dst = slot.data_off = Alloc(L * 16);
		uint64_t n_chunks_col = (L + B2A_VECTOR_SIZE - 1) / B2A_VECTOR_SIZE;
		t.aux = slot.aux_off = Alloc((n_chunks_col + 1) * 8);
		// the child's node follows in pre-order: the last offset must not exceed its length (nanoarrow FULL validation)
		t.param = node_idx < rb.nodes.size() ? static_cast<int64_t>(rb.nodes[node_idx].length) : 0;
		if (!chain.levels.empty()) {
			t.variadic0 = static_cast<uint32_t>(variadic.size());
			t.src2_len = chain.levels.size();
			variadic.push_back(chain.top_rows);
			variadic.push_back(chain.levels.size());
			for (auto &lv : chain.levels) {
				variadic.push_back(lv.first);
				variadic.push_back(lv.second);
			}
		}
		if (L > 0) {
			Touch(t.src, (L + 1) * ow);
			algorithmic_out += L * 16;
			AddTask(t, 16 * ow, 16);
		}
		break;
	}
	case B2A_OP_LISTVIEW32:
	case B2A_OP_LISTVIEW64: {
		// ConvertArrowListViewOffsetsTemplated (:166-203): per conversion call (= per 2 048-row chunk at the top level)
		// entries {offsets[i] - lowest offset, sizes[i]}, child converted for sum(sizes) rows from the lowest offset
		if (!chain.levels.empty()) {
			throw Error(B2A_ENOTIMPL, "Arrow IPC: list views below a list are not supported by the B200 scan path");
		}
		CheckListViewChild(*f.children.at(0));
		const uint64_t ow = static_cast<uint64_t>(f.src_width);
		NeedBytes(bufs[1], L * ow, f, "offsets");
		NeedBytes(bufs[2], L * ow, f, "sizes");
		if (node_idx >= rb.nodes.size()) {
			throw Error(B2A_EIO, "Arrow IPC: RecordBatch has too few nodes for field '" + f.name + "'");
		}
		t.src = bufs[1].offset;
		t.src2 = bufs[2].offset;
		t.param = rb.nodes[node_idx].length; // child rows: offsets[i] + sizes[i] must stay inside (FULL validation)
		t.flags |= B2A_TASK_F_PHASE2;        // runs in the cold-op kernel (needs a block-wide reduction per chunk)
		t.dst = slot.data_off = Alloc(L * 16);
		uint64_t n_chunks_col = (L + B2A_VECTOR_SIZE - 1) / B2A_VECTOR_SIZE;
		t.aux = slot.aux_off = Alloc((n_chunks_col + 1) * 16);
		if (L > 0) {
			Touch(t.src, L * ow);
			Touch(t.src2, L * ow);
			algorithmic_out += L * 16;
			phase2_pending.push_back(t);
		}
		break;
	}
	default: { // COPY and the elementwise transforms: src_width bytes in, out_width bytes out
		const uint64_t sw = static_cast<uint64_t>(f.src_width), dw = static_cast<uint64_t>(f.out_width);
		NeedBytes(bufs[1], L * sw, f, "data");
		if (alias_direct && f.op == B2A_OP_COPY && (!big_endian || sw == 1)) {
			slot.image_off = bufs[1].offset; // DirectConversion: the vector points at the Arrow buffer itself
			if (alias_on_device) {
				Touch(bufs[1].offset, L * sw); // (read in place by device consumers: part of the upload)
			}
			break;
		}
		t.src = bufs[1].offset;
		t.dst = slot.data_off = Alloc(L * dw);
		Touch(t.src, L * sw);
		algorithmic_out += L * dw;
		if (f.op == B2A_OP_BOOL8) { // rarely used: lives in the second-launch kernel, not in the hot one's switch
			t.flags |= B2A_TASK_F_PHASE2;
			phase2_pending.push_back(t);
			break;
		}
		AddTask(t, 16 * sw, dw);
		break;
	}
	}
	buf_idx += n_variadic;

	// ---- children
	switch (f.type_id) {
	case B2A_T_LIST:
	case B2A_T_MAP: {
		Chain down = chain;
		if (f.op == B2A_OP_LISTVIEW32 || f.op == B2A_OP_LISTVIEW64) {
			// the child is decoded once for the whole child array; each chunk's vector is a window of it (aux pairs)
			EmitField(reader, *f.children.at(0), rb, node_idx, buf_idx, var_idx, down, true, true, slots, false, false);
			break;
		}
		if (down.levels.empty()) {
			down.top_rows = L;
		}
		down.levels.emplace_back(f.op == B2A_OP_LIST32 ? 0 : 1, bufs[1].offset);
		EmitField(reader, *f.children.at(0), rb, node_idx, buf_idx, var_idx, down, true, true, slots, false, false);
		break;
	}
	case B2A_T_STRUCT:
		for (auto &c : f.children) {
			size_t child_node = node_idx;
			EmitField(reader, *c, rb, node_idx, buf_idx, var_idx, chain, false, true, slots, false, false);
			if (static_cast<uint64_t>(rb.nodes[child_node].length) < L) {
				throw Error(B2A_EIO, "Arrow IPC: struct child '" + c->name + "' is shorter than its parent");
			}
		}
		break;
	case B2A_T_ARRAY: {
		Chain scaled = chain;
		if (scaled.levels.empty()) {
			scaled.top_rows = L;
		}
		scaled.levels.emplace_back(2, static_cast<uint64_t>(f.param));
		for (uint32_t i = 0; i < scaled.n; i++) {
			scaled.m[i].div *= static_cast<uint64_t>(f.param);
		}
		size_t child_node = node_idx;
		EmitField(reader, *f.children.at(0), rb, node_idx, buf_idx, var_idx, scaled, false, true, slots, false, false);
		if (static_cast<uint64_t>(rb.nodes[child_node].length) < L * static_cast<uint64_t>(f.param)) {
			throw Error(B2A_EIO, "Arrow IPC: fixed-size-list child is shorter than length * list_size");
		}
		break;
	}
	default:
		for (auto &c : f.children) { // types we do not descend into have no children; keep the cursor right
			SkipField(*c, rb, node_idx, buf_idx, var_idx, false);
		}
		break;
	}
}

void ScanPlan::Build(const Reader &reader, uint64_t lo, uint64_t hi) {
	const Schema &schema = reader.BaseSchema();
	const auto &batches = reader.Batches();
	if (lo > hi || hi > batches.size()) {
		throw Error(B2A_EINVAL, "batch range out of bounds");
	}
	batch_lo = lo;
	batch_hi = hi;
	n_flat = static_cast<uint32_t>(schema.flat.size());
	big_endian = !schema.little_endian;
	top_columns = reader.OutputColumns();
	image_lo = UINT64_MAX;
	image_hi = 0;
	out_bytes = 0;
	std::vector<char> selected(schema.fields.size(), 0);
	for (int c : top_columns) {
		selected[static_cast<size_t>(c)] = 1;
	}
	b2a_column_out none;
	std::memset(&none, 0xFF, sizeof(none));
	none.rows = 0;
	none.null_count = 0;
	none.chars_len = 0;
	cols.assign((hi - lo) * n_flat, none);

	for (uint64_t b = lo; b < hi; b++) {
		RecordBatch virt;
		if (batches[b].compression >= 0) {
			virt = VirtualBatch(reader, batches[b]);
		}
		const RecordBatch &rb = batches[b].compression >= 0 ? virt : batches[b];
		batch_job0.push_back(static_cast<uint32_t>(jobs.size()));
		batch_rows.push_back(rb.rows);
		batch_chunk0.push_back(n_chunks);
		batch_tile0.push_back(next_tile);
		batch_out0.push_back(out_bytes);
		batch_ranges.emplace_back();
		cur_ranges = &batch_ranges.back();
		const uint64_t save_lo = image_lo, save_hi = image_hi;
		image_lo = UINT64_MAX;
		image_hi = 0;
		n_chunks += (rb.rows + B2A_VECTOR_SIZE - 1) / B2A_VECTOR_SIZE;
		total_rows += rb.rows;
		size_t node_idx = 0, buf_idx = 0, var_idx = 0;
		Chain empty;
		for (size_t c = 0; c < schema.fields.size(); c++) {
			size_t node0 = node_idx;
			EmitField(reader, *schema.fields[c], rb, node_idx, buf_idx, var_idx, empty, false, selected[c] != 0,
			          cols.data() + (b - lo) * n_flat, false, false);
			// "arrow_scan: array length mismatch" (arrow_conversion.cpp:1454-1456)
			if (selected[c] && static_cast<uint64_t>(rb.nodes[node0].length) != rb.rows) {
				throw Error(B2A_EINVAL, "arrow_scan: array length mismatch");
			}
		}
		batch_image_lo.push_back(image_lo == UINT64_MAX ? 0 : image_lo);
		batch_image_hi.push_back(image_hi);
		image_lo = std::min(image_lo, save_lo);
		image_hi = std::max(image_hi, save_hi);
	}
	batch_tile0.push_back(next_tile);
	batch_out0.push_back(out_bytes);
	batch_job0.push_back(static_cast<uint32_t>(jobs.size()));
	cur_ranges = &tail_ranges;
	// dictionaries: decoded once per scan (ArrowArrayScanState::AddDictionary, arrow_array_scan_state.cpp:29-52)
	std::vector<int> dict_fields;
	for (auto &kv : dict_cols) {
		dict_fields.push_back(kv.first);
	}
	std::vector<b2a_column_out> tmp(n_flat, none);
	for (int flat : dict_fields) {
		const Field &f = *schema.flat[static_cast<size_t>(flat)];
		const DictionaryBatch &d = reader.Dictionaries().at(f.dict_id);
		RecordBatch virt;
		if (d.data.compression >= 0) {
			virt = VirtualBatch(reader, d.data);
		}
		const RecordBatch &drb = d.data.compression >= 0 ? virt : d.data;
		size_t node_idx = 0, buf_idx = 0, var_idx = 0;
		Chain empty;
		EmitField(reader, f, drb, node_idx, buf_idx, var_idx, empty, false, true, tmp.data(), true, true);
		dict_cols[flat] = tmp[static_cast<size_t>(flat)];
	}
	lean = true;
	for (auto &t : tasks) {
		lean = lean && (t.op == B2A_OP_COPY || t.op == B2A_OP_VALIDITY || t.op == B2A_OP_BOOL || t.op == B2A_OP_STRING32 ||
		                t.op == B2A_OP_STRING64);
	}
	// tasks that read decoded vectors (REE gathers) go last: they are launched after everything else has run
	phase2_tile0 = next_tile;
	for (auto &g : phase2_pending) {
		AddTask(g, 16 * static_cast<uint64_t>(g.op == B2A_OP_BOOL8 ? 1 : 0), g.dst_width);
	}
	phase2_pending.clear();
	// the popcount cells live in one region at the end of the arena (zeroed before every decode)
	stats_base = out_bytes;
	out_bytes += (stats_bytes + 255) & ~uint64_t(255);
	for (auto &t : tasks) {
		if (t.op == B2A_OP_VALIDITY) {
			t.aux += stats_base;
		}
	}
	for (auto &c : cols) {
		if (c.stats_off != kNone) {
			c.stats_off += stats_base;
		}
	}
	for (auto &kv : dict_cols) {
		if (kv.second.stats_off != kNone) {
			kv.second.stats_off += stats_base;
		}
	}
	if (getenv("B2A_DUMP_TASKS")) {
		for (auto &t : tasks) {
			fprintf(stderr, "task op=%u src=%llu src2=%llu src2_len=%llu dst=%llu aux=%llu rows=%llu sw=%u dw=%u flags=%u tile0=%u tile_rows=%u param=%lld\n",
			        t.op, (unsigned long long)t.src, (unsigned long long)t.src2, (unsigned long long)t.src2_len,
			        (unsigned long long)t.dst, (unsigned long long)t.aux, (unsigned long long)t.rows, t.src_width,
			        t.dst_width, t.flags, t.tile0, t.tile_rows, (long long)t.param);
		}
	}
	cur_ranges = nullptr;
	if (image_lo == UINT64_MAX) {
		image_lo = image_hi = 0;
	}
	image_lo &= ~uint64_t(255);
	if (compressed) {
		// the scratch region follows the uploaded slice in the image offset space; virtual offsets become real ones
		scratch_base = (image_hi + 255) & ~uint64_t(255);
		scratch_bytes += 64; // the string kernel fetches whole aligned words around the last characters
		for (auto &j : jobs) {
			j.dst += scratch_base;
		}
		for (auto &t : tasks) {
			t.src = Resolve(t.src);
			t.src2 = Resolve(t.src2);
			for (uint32_t i = 0; i < t.n_masks; i++) {
				t.masks[i].src = Resolve(t.masks[i].src);
			}
		}
		for (auto &v : variadic) {
			v = Resolve(v);
		}
		for (auto &c : cols) {
			c.chars_off = Resolve(c.chars_off);
		}
		for (auto &kv : dict_cols) {
			kv.second.chars_off = Resolve(kv.second.chars_off);
		}
	}
	upload_bytes = 0;
	for (auto &r : batch_ranges) {
		MergeRanges(r, reader.ImageSize());
		for (auto &x : r) {
			upload_bytes += x.second - x.first;
		}
	}
	MergeRanges(tail_ranges, reader.ImageSize());
	for (auto &x : tail_ranges) {
		upload_bytes += x.second - x.first;
	}
	if (out_bytes == 0) {
		out_bytes = 256;
	}
}

} // namespace b2a
