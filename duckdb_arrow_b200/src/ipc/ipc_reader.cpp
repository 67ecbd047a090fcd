// Host framing + slicing.  Reference behaviour restated:
//  * 8-byte alignment skip before every prefix/body     ipc_file_stream_reader.cpp:133-141
//  * "ARROW1\0\0" magic at offset 0 is skipped, the footer is never read   :107-119
//  * prefix = {u32 0xFFFFFFFF, i32 metadata_size}; a different token is an IOException   :121-124
//  * a truncated prefix, or metadata_size == 0, is a clean end of stream   :126-129 + nanoarrow ENODATA :63-66
//  * metadata_size < 0 is an IOException                base_stream_reader.cpp:214-227
//  * first message must be a Schema; afterwards only RecordBatch is accepted by the reference
//    (base_stream_reader.cpp:86-96).  DictionaryBatch is accepted here in addition (BASELINE config 4).
#include "ipc_reader.hpp"
#include "flatbuf.hpp"

#include <algorithm>
#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <unordered_set>

#include <mutex>
#include <thread>
#include <vector>

namespace b2a {

static const uint32_t kContinuationToken = 0xFFFFFFFFu;

// ---- mappings of large files are kept for the next open of the same (unchanged) file: unmapping an 8.5 GB mapping whose
// pages were touched and faulting a fresh one in again cost ~ 30 ms per read_arrow call (tools/time_call_parts.py), and the
// address-space lock held by munmap stalls whoever maps next.  A mapping is identified by {device, inode, size, mtime ns}
// and handed to ONE reader at a time; at most two idle mappings are kept, an evicted one is unmapped by a detached thread,
// 128 MiB at a time.
namespace {
struct IdleMapping {
	dev_t dev;
	ino_t ino;
	off_t size;
	int64_t mtime_ns;
	void *base;
};
std::mutex g_map_mu;
std::vector<IdleMapping> g_idle_maps; // oldest first
constexpr size_t kCachedMappingMin = size_t(64) << 20;
constexpr size_t kIdleMappings = 2;

int64_t MtimeNs(const struct stat &st) {
	return static_cast<int64_t>(st.st_mtim.tv_sec) * 1000000000ll + st.st_mtim.tv_nsec;
}

void UnmapInBackground(void *base_p, size_t len) {
	uint8_t *base = static_cast<uint8_t *>(base_p);
	try {
		std::thread([base, len] {
			const size_t step = size_t(128) << 20;
			for (size_t off = 0; off < len; off += step) {
				munmap(base + off, len - off < step ? len - off : step);
			}
		}).detach();
	} catch (...) {
		munmap(base, len);
	}
}

void *TakeIdleMapping(const struct stat &st) {
	std::lock_guard<std::mutex> lock(g_map_mu);
	for (size_t i = 0; i < g_idle_maps.size(); i++) {
		const IdleMapping &m = g_idle_maps[i];
		if (m.dev == st.st_dev && m.ino == st.st_ino && m.size == st.st_size && m.mtime_ns == MtimeNs(st)) {
			void *base = m.base;
			g_idle_maps.erase(g_idle_maps.begin() + static_cast<std::ptrdiff_t>(i));
			return base;
		}
	}
	return nullptr;
}

void ReleaseMapping(int fd, void *base, size_t len) {
	struct stat st;
	if (len < kCachedMappingMin || fd < 0 || fstat(fd, &st) != 0 || static_cast<size_t>(st.st_size) != len) {
		if (len >= kCachedMappingMin) {
			UnmapInBackground(base, len);
		} else {
			munmap(base, len);
		}
		return;
	}
	IdleMapping evicted {};
	bool evict = false;
	{
		std::lock_guard<std::mutex> lock(g_map_mu);
		g_idle_maps.push_back(IdleMapping {st.st_dev, st.st_ino, st.st_size, MtimeNs(st), base});
		if (g_idle_maps.size() > kIdleMappings) {
			evicted = g_idle_maps.front();
			g_idle_maps.erase(g_idle_maps.begin());
			evict = true;
		}
	}
	if (evict) {
		UnmapInBackground(evicted.base, static_cast<size_t>(evicted.size));
	}
}
} // namespace

Reader::~Reader() {
	if (mmap_base) {
		ReleaseMapping(fd, mmap_base, mmap_len); // (before the descriptor goes: the mapping is filed under the file's identity)
	}
	if (fd >= 0) {
		close(fd);
	}
	if (fd_direct >= 0) {
		close(fd_direct);
	}
}

int Reader::DirectFileDescriptor() {
	if (fd_direct == -2) {
		fd_direct = (fd >= 0 && !path.empty()) ? open(path.c_str(), O_RDONLY | O_DIRECT) : -1;
	}
	return fd_direct;
}

void Reader::OpenFile(const std::string &path_p) {
	path = path_p;
	fd = open(path.c_str(), O_RDONLY);
	if (fd < 0) {
		throw Error(B2A_EIO, "Cannot open file \"" + path + "\": " + std::strerror(errno));
	}
	struct stat st;
	if (fstat(fd, &st) != 0) {
		throw Error(B2A_EIO, "Cannot stat file \"" + path + "\"");
	}
	image_size = static_cast<uint64_t>(st.st_size);
	if (image_size > 0) {
		mmap_base = image_size >= kCachedMappingMin ? TakeIdleMapping(st) : nullptr;
		if (!mmap_base) {
			mmap_base = mmap(nullptr, image_size, PROT_READ, MAP_PRIVATE, fd, 0);
		}
		if (mmap_base == MAP_FAILED) {
			mmap_base = nullptr;
			throw Error(B2A_EIO, "Cannot mmap file \"" + path + "\"");
		}
		mmap_len = image_size;
		image = static_cast<const uint8_t *>(mmap_base);
	}
	Index();
}

void Reader::OpenBuffers(const b2a_ipc_buffer *bufs, uint32_t n) {
	if (n == 0) {
		throw Error(B2A_EINVAL, "scan_arrow_ipc: empty buffer list");
	}
	if (n == 1) {
		image = static_cast<const uint8_t *>(bufs[0].ptr);
		image_size = bufs[0].size;
	} else {
		uint64_t total = 0;
		for (uint32_t i = 0; i < n; i++) {
			total = (total + 7) & ~uint64_t(7);
			total += bufs[i].size;
		}
		owned.assign(total, 0);
		uint64_t pos = 0;
		for (uint32_t i = 0; i < n; i++) {
			pos = (pos + 7) & ~uint64_t(7);
			std::memcpy(owned.data() + pos, bufs[i].ptr, bufs[i].size);
			pos += bufs[i].size;
		}
		image = owned.data();
		image_size = total;
	}
	Index();
}

static uint64_t Align8(uint64_t v) {
	return (v + 7) & ~uint64_t(7);
}

bool Reader::ReadNextMessage(uint64_t &pos) {
	pos = Align8(pos); // EnsureInputStreamAligned
	if (pos + 8 > image_size) {
		return false; // SerializationException on the prefix read == clean EOS
	}
	if (pos == 0 && std::memcmp("ARROW1\0\0", image, 8) == 0) {
		pos = 8;
		return ReadNextMessage(pos);
	}
	uint32_t token;
	int32_t metadata_size;
	std::memcpy(&token, image + pos, 4);
	std::memcpy(&metadata_size, image + pos + 4, 4);
	if (token != kContinuationToken) {
		throw Error(B2A_EIO, "Expected continuation token (0xFFFFFFFF) but got " + std::to_string(token));
	}
	if (metadata_size < 0) {
		throw Error(B2A_EIO, "Expected metadata size >= 0 but got " + std::to_string(metadata_size));
	}
	if (metadata_size == 0) {
		return false; // EOS marker (nanoarrow returns ENODATA)
	}
	uint64_t meta_pos = pos + 8;
	if (meta_pos + static_cast<uint64_t>(metadata_size) > image_size) {
		throw Error(B2A_EIO, "Arrow IPC: message metadata extends past the end of the input");
	}
	const uint8_t *meta = image + meta_pos;
	size_t meta_len = static_cast<size_t>(metadata_size);

	try {
		fb::Table msg = fb::root(meta, meta_len);
		int16_t version = msg.scalar<int16_t>(0, 0);
		if (version < 3) { // MetadataVersion::V4 == 3; nanoarrow rejects older streams
			throw Error(B2A_EIO, "Arrow IPC: unsupported metadata version (expected V4 or V5)");
		}
		uint8_t header_type = msg.scalar<uint8_t>(1, 0);
		int64_t body_length = msg.scalar<int64_t>(3, 0);
		if (body_length < 0) {
			throw Error(B2A_EIO, "Arrow IPC: negative bodyLength");
		}
		uint64_t body_offset = Align8(meta_pos + meta_len);
		if (body_length > 0 && body_offset + static_cast<uint64_t>(body_length) > image_size) {
			throw Error(B2A_EIO, "Arrow IPC: message body extends past the end of the input");
		}
		if (!have_schema) {
			if (header_type != 1) {
				throw Error(B2A_EIO, "Expected Schema Arrow IPC message but got " +
				                         std::string(header_type == 3   ? "RecordBatch"
				                                     : header_type == 2 ? "DictionaryBatch"
				                                                        : "Uninitialized"));
			}
			SchemaParser::ParseMessageSchema(meta, meta_len, schema);
			if (version == 3) {
				// MetadataVersion V4 unions carry a validity buffer of their own (two buffers); only the V5 layout
				// (type ids only) is handled
				for (const Field *fl : schema.flat) {
					if (fl->type_id == B2A_T_UNION) {
						throw Error(B2A_ENOTIMPL, "Arrow IPC: unions of metadata version V4 streams are not supported");
					}
				}
			}
			have_schema = true;
			// nodes / fixed buffers one RecordBatch must carry (dictionary-encoded fields contribute only their
			// index array: validity + indices, and none of the value type's children)
			flat_buffer_start.clear();
			struct Count {
				static void Go(const Field &f, int &nodes, int &bufs, bool &variadic) {
					nodes++;
					if (f.dict_id >= 0) {
						bufs += 2;
						return;
					}
					bufs += f.n_buffers;
					variadic = variadic || f.variadic;
					for (auto &c : f.children) {
						Go(*c, nodes, bufs, variadic);
					}
				}
			};
			int nodes = 0, nbuf = 0;
			for (auto &f : schema.fields) {
				Count::Go(*f, nodes, nbuf, has_variadic);
			}
			flat_buffer_start.push_back(nodes);
			flat_buffer_start.push_back(nbuf);
		} else if (header_type == 3) {
			fb::Table rb = msg.table(2);
			if (!rb.valid) {
				throw Error(B2A_EIO, "Arrow IPC: RecordBatch message without header");
			}
			RecordBatch batch;
			batch.meta_offset = pos;
			ParseRecordBatch(meta, meta_len, rb.pos, body_offset, static_cast<uint64_t>(body_length), batch, false);
			total_rows += batch.rows;
			batches.push_back(std::move(batch));
		} else if (header_type == 2) {
			fb::Table db = msg.table(2);
			if (!db.valid) {
				throw Error(B2A_EIO, "Arrow IPC: DictionaryBatch message without header");
			}
			DictionaryBatch d;
			d.id = db.scalar<int64_t>(0, 0);
			d.is_delta = db.scalar<uint8_t>(2, 0) != 0;
			fb::Table rb = db.table(1);
			if (!rb.valid) {
				throw Error(B2A_EIO, "Arrow IPC: DictionaryBatch without data");
			}
			d.data.meta_offset = pos;
			ParseRecordBatch(meta, meta_len, rb.pos, body_offset, static_cast<uint64_t>(body_length), d.data, true);
			if (d.is_delta) {
				throw Error(B2A_ENOTIMPL, "Arrow IPC: delta dictionaries are not supported");
			}
			if (dictionaries.count(d.id)) {
				// the reference rejects DICTIONARY_REPLACEMENT (base_stream_reader.cpp:59-61)
				throw Error(B2A_EIO, "This stream uses unsupported feature DICTIONARY_REPLACEMENT");
			}
			dictionaries[d.id] = std::move(d);
		} else if (header_type == 1) {
			throw Error(B2A_EIO, "Expected RecordBatch Arrow IPC message but got Schema");
		} else {
			throw Error(B2A_EIO, "Expected RecordBatch Arrow IPC message but got " +
			                         std::string(header_type == 4 ? "Tensor" : "SparseTensor"));
		}
		pos = body_offset + static_cast<uint64_t>(body_length);
	} catch (fb::ParseError &e) {
		throw Error(B2A_EIO, std::string("Arrow IPC: invalid message flatbuffer: ") + e.what());
	}
	return true;
}

// Arrow FILE format: "ARROW1\0\0" <stream incl. EOS> <Footer flatbuffer> <int32 footer length> "ARROW1".
// Footer{version, schema, dictionaries:[Block], recordBatches:[Block]}, Block = struct{offset:int64, metaDataLength:int32,
// bodyLength:int64} (24 bytes).  The reference never reads it (ipc_file_stream_reader.cpp:107-119 skips the magic and
// walks the stream); using it gives the same messages without touching one page per message header in file order —
// SURVEY §8f N-3.  Any inconsistency falls back to the walk, so the results cannot differ.
bool Reader::IndexFromFooter() {
	static const bool disabled = getenv("B2A_NO_FOOTER") != nullptr;
	if (disabled || image_size < 8 + 8 + 10 || std::memcmp("ARROW1\0\0", image, 8) != 0 ||
	    std::memcmp("ARROW1", image + image_size - 6, 6) != 0) {
		return false;
	}
	int32_t footer_len;
	std::memcpy(&footer_len, image + image_size - 10, 4);
	if (footer_len <= 0 || static_cast<uint64_t>(footer_len) + 10 + 8 > image_size) {
		return false;
	}
	const uint8_t *footer = image + image_size - 10 - static_cast<uint64_t>(footer_len);
	try {
		fb::Table ft = fb::root(footer, static_cast<size_t>(footer_len));
		struct Block {
			uint64_t offset;
		};
		std::vector<Block> blocks[2];
		for (int which = 0; which < 2; which++) {
			size_t data;
			uint32_t n = ft.vec(2 + static_cast<unsigned>(which), data, 24);
			fb::Buf fbuf {footer, static_cast<size_t>(footer_len)};
			for (uint32_t i = 0; i < n; i++) {
				const int64_t off = fbuf.rd<int64_t>(data + 24 * size_t(i));
				if (off < 8 || (off & 7) != 0 || static_cast<uint64_t>(off) + 8 > image_size) {
					return false;
				}
				blocks[which].push_back(Block {static_cast<uint64_t>(off)});
			}
		}
		uint64_t pos = 8;
		if (!ReadNextMessage(pos) || !have_schema) { // the stream's own Schema message
			throw Error(B2A_EIO, "no schema");
		}
		for (auto &b : blocks[0]) {
			const size_t before = dictionaries.size();
			uint64_t p = b.offset;
			if (!ReadNextMessage(p) || dictionaries.size() != before + 1) {
				throw Error(B2A_EIO, "footer dictionary block does not hold a DictionaryBatch");
			}
		}
		for (auto &b : blocks[1]) {
			const size_t before = batches.size();
			uint64_t p = b.offset;
			if (!ReadNextMessage(p) || batches.size() != before + 1) {
				throw Error(B2A_EIO, "footer record batch block does not hold a RecordBatch");
			}
		}
		// blocks in file order <=> the order the walk would have produced
		for (size_t i = 1; i < batches.size(); i++) {
			if (batches[i].meta_offset <= batches[i - 1].meta_offset) {
				throw Error(B2A_EIO, "footer blocks out of order");
			}
		}
		return true;
	} catch (const Error &) {
	} catch (const fb::ParseError &) {
	}
	// anything odd: forget what was indexed and walk the stream like the reference does
	schema = Schema();
	have_schema = false;
	has_variadic = false;
	batches.clear();
	dictionaries.clear();
	flat_buffer_start.clear();
	total_rows = 0;
	return false;
}

void Reader::Index() {
	used_footer = IndexFromFooter();
	if (used_footer) {
		return;
	}
	uint64_t pos = 0;
	while (ReadNextMessage(pos)) {
	}
	if (!have_schema) {
		throw Error(B2A_EIO, "Expected Schema Arrow IPC message but got end of stream");
	}
}

void Reader::ParseRecordBatch(const uint8_t *tbl_buf, size_t tbl_len, size_t table_pos, uint64_t body_offset,
                              uint64_t body_length, RecordBatch &out, bool is_dictionary) {
	fb::Buf b {tbl_buf, tbl_len};
	fb::Table rb(b, table_pos);
	int64_t length = rb.scalar<int64_t>(0, 0);
	if (length < 0) {
		throw Error(B2A_EIO, "Arrow IPC: negative RecordBatch length");
	}
	out.rows = static_cast<uint64_t>(length);
	out.body_offset = body_offset;
	out.body_length = body_length;

	size_t ndata;
	uint32_t n_nodes = rb.vec(1, ndata, 16);
	out.nodes.resize(n_nodes);
	for (uint32_t i = 0; i < n_nodes; i++) {
		out.nodes[i].length = b.rd<int64_t>(ndata + 16 * size_t(i));
		out.nodes[i].null_count = b.rd<int64_t>(ndata + 16 * size_t(i) + 8);
		if (out.nodes[i].length < 0) {
			throw Error(B2A_EIO, "Arrow IPC: negative FieldNode length");
		}
	}
	size_t bdata;
	uint32_t n_buffers = rb.vec(2, bdata, 16);
	out.buffers.resize(n_buffers);
	for (uint32_t i = 0; i < n_buffers; i++) {
		int64_t off = b.rd<int64_t>(bdata + 16 * size_t(i));
		int64_t len = b.rd<int64_t>(bdata + 16 * size_t(i) + 8);
		if (off < 0 || len < 0 || static_cast<uint64_t>(off) + static_cast<uint64_t>(len) > body_length) {
			throw Error(B2A_EIO, "Arrow IPC: buffer " + std::to_string(i) + " lies outside the message body");
		}
		if ((off & 7) != 0) {
			throw Error(B2A_EIO, "Arrow IPC: buffer " + std::to_string(i) + " is not 8-byte aligned");
		}
		out.buffers[i].offset = body_offset + static_cast<uint64_t>(off);
		out.buffers[i].length = static_cast<uint64_t>(len);
	}
	fb::Table comp = rb.table(3);
	if (comp.valid) {
		out.compression = comp.scalar<int8_t>(0, 0);
	}
	size_t vdata;
	uint32_t n_var = rb.vec(4, vdata, 8);
	for (uint32_t i = 0; i < n_var; i++) {
		const int64_t c = b.rd<int64_t>(vdata + 8 * size_t(i));
		// a count is a number of buffers of THIS message: negative or larger than the buffer list is malformed (and
		// would wrap the sums below / in the planner)
		if (c < 0 || static_cast<uint64_t>(c) > n_buffers) {
			throw Error(B2A_EIO, "Arrow IPC: variadicBufferCounts[" + std::to_string(i) + "] = " + std::to_string(c) +
			                         " is not a valid buffer count");
		}
		out.variadic_counts.push_back(c);
	}
	if (!is_dictionary) {
		if (n_nodes != static_cast<uint32_t>(flat_buffer_start.front())) {
			throw Error(B2A_EIO, "Arrow IPC: RecordBatch has " + std::to_string(n_nodes) + " field nodes, schema has " +
			                         std::to_string(flat_buffer_start.front()));
		}
		uint64_t expect = static_cast<uint64_t>(flat_buffer_start.back());
		for (auto c : out.variadic_counts) {
			expect += static_cast<uint64_t>(c);
		}
		if (n_buffers != expect) {
			throw Error(B2A_EIO, "Arrow IPC: RecordBatch has " + std::to_string(n_buffers) + " buffers, schema needs " +
			                         std::to_string(expect));
		}
	}
}

std::vector<int> Reader::OutputColumns() const {
	if (!projected_top.empty()) {
		return projected_top;
	}
	std::vector<int> all(schema.fields.size());
	for (size_t i = 0; i < all.size(); i++) {
		all[i] = static_cast<int>(i);
	}
	return all;
}

static std::string Lower(const std::string &s) {
	std::string r = s;
	std::transform(r.begin(), r.end(), r.begin(), [](unsigned char c) { return std::tolower(c); });
	return r;
}

// QueryResult::DeduplicateColumns (duckdb/src/main/query_result.cpp:71-93)
void DeduplicateColumns(std::vector<std::string> &names) {
	std::unordered_map<std::string, uint64_t> name_map;
	for (auto &column_name : names) {
		auto low = Lower(column_name);
		if (name_map.find(low) == name_map.end()) {
			name_map[low]++;
		} else {
			std::string new_name = column_name + "_" + std::to_string(name_map[low]);
			auto new_low = Lower(new_name);
			while (name_map.find(new_low) != name_map.end()) {
				name_map[low]++;
				new_name = column_name + "_" + std::to_string(name_map[low]);
				new_low = Lower(new_name);
			}
			column_name = new_name;
			name_map[new_low]++;
		}
	}
}

void Reader::SetProjection(const std::vector<std::string> &column_names) {
	if (column_names.empty()) {
		throw Error(B2A_EINTERNAL, "Can't request zero fields projected from IpcStreamReader");
	}
	std::vector<std::string> names;
	for (auto &f : schema.fields) {
		names.push_back(f->name);
	}
	DeduplicateColumns(names);
	std::unordered_map<std::string, int> name_to_field;
	std::unordered_set<std::string> duplicates;
	for (size_t i = 0; i < names.size(); i++) {
		if (name_to_field.find(names[i]) != name_to_field.end()) {
			duplicates.insert(names[i]);
		}
		name_to_field.insert({names[i], static_cast<int>(i)});
	}
	std::vector<int> proj;
	for (auto &name : column_names) {
		if (duplicates.count(name)) {
			throw Error(B2A_EINTERNAL, "Field '" + name + "' refers to a duplicate column name in IPC file schema");
		}
		auto it = name_to_field.find(name);
		if (it == name_to_field.end()) {
			throw Error(B2A_EINTERNAL, "Field '" + name + "' does not exist in IPC file schema");
		}
		proj.push_back(it->second);
	}
	projected_top = std::move(proj);
}

} // namespace b2a
