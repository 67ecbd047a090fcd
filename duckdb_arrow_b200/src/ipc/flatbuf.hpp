// Minimal bounds-checked FlatBuffers reader + a tiny builder, enough for Arrow's Message.fbs /
// Schema.fbs (tables, strings, vectors, 16-byte structs).  Replaces the flatcc-generated readers
// nanoarrow_ipc carries (nanoarrow is not in the reference tree: CMakeLists.txt:7-13).
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace b2a {
namespace fb {

struct ParseError : std::runtime_error {
	using std::runtime_error::runtime_error;
};

struct Buf {
	const uint8_t *p = nullptr;
	size_t n = 0;
	void need(size_t off, size_t len) const {
		if (off > n || len > n - off) {
			throw ParseError("flatbuffer: offset out of bounds");
		}
	}
	template <class T>
	T rd(size_t off) const {
		need(off, sizeof(T));
		T v;
		std::memcpy(&v, p + off, sizeof(T));
		return v;
	}
};

struct Table {
	Buf b;
	size_t pos = 0;  // table start
	size_t vt = 0;   // vtable start
	uint16_t vtsize = 0;
	bool valid = false;

	Table() {
	}
	Table(const Buf &buf, size_t table_pos) : b(buf), pos(table_pos), valid(true) {
		int32_t soff = b.rd<int32_t>(pos);
		int64_t v = static_cast<int64_t>(pos) - soff;
		if (v < 0) {
			throw ParseError("flatbuffer: bad vtable offset");
		}
		vt = static_cast<size_t>(v);
		vtsize = b.rd<uint16_t>(vt);
		if (vtsize < 4) {
			throw ParseError("flatbuffer: bad vtable size");
		}
		b.need(vt, vtsize);
	}
	// byte offset of field `id` inside the buffer, 0 if absent
	size_t field(unsigned id) const {
		size_t slot = 4 + 2 * static_cast<size_t>(id);
		if (slot + 2 > vtsize) {
			return 0;
		}
		uint16_t o = b.rd<uint16_t>(vt + slot);
		return o ? pos + o : 0;
	}
	template <class T>
	T scalar(unsigned id, T def) const {
		size_t f = field(id);
		return f ? b.rd<T>(f) : def;
	}
	bool has(unsigned id) const {
		return field(id) != 0;
	}
	// follow a uoffset stored in field `id`
	size_t indirect(unsigned id) const {
		size_t f = field(id);
		if (!f) {
			return 0;
		}
		return f + b.rd<uint32_t>(f);
	}
	Table table(unsigned id) const {
		size_t t = indirect(id);
		return t ? Table(b, t) : Table();
	}
	std::string str(unsigned id) const {
		size_t s = indirect(id);
		if (!s) {
			return std::string();
		}
		uint32_t len = b.rd<uint32_t>(s);
		b.need(s + 4, len);
		return std::string(reinterpret_cast<const char *>(b.p + s + 4), len);
	}
	// vector: returns element count, sets `data` to the first element's offset
	uint32_t vec(unsigned id, size_t &data, size_t elem_size) const {
		size_t v = indirect(id);
		if (!v) {
			data = 0;
			return 0;
		}
		uint32_t len = b.rd<uint32_t>(v);
		b.need(v + 4, static_cast<size_t>(len) * elem_size);
		data = v + 4;
		return len;
	}
	Table vec_table(size_t data, uint32_t i) const {
		size_t e = data + 4 * static_cast<size_t>(i);
		return Table(b, e + b.rd<uint32_t>(e));
	}
};

inline Table root(const uint8_t *p, size_t n) {
	Buf b {p, n};
	uint32_t off = b.rd<uint32_t>(0);
	return Table(b, off);
}

// ---------------------------------------------------------------------------------------------
// Builder: writes back-to-front like the canonical implementations, so that every offset is
// "ahead" of its referrer.  Only what Message/Schema/RecordBatch need.
// ---------------------------------------------------------------------------------------------
class Builder {
public:
	// bytes are accumulated reversed (buf_[0] is the LAST byte of the finished buffer)
	size_t size() const {
		return buf_.size();
	}
	void align(size_t a, size_t upcoming = 0) {
		while ((buf_.size() + upcoming) % a) {
			buf_.push_back(0);
		}
		if (a > min_align_) {
			min_align_ = a;
		}
	}
	template <class T>
	void push(T v) {
		uint8_t tmp[sizeof(T)];
		std::memcpy(tmp, &v, sizeof(T));
		for (size_t i = sizeof(T); i-- > 0;) {
			buf_.push_back(tmp[i]);
		}
	}
	void push_bytes(const void *p, size_t n) {
		const uint8_t *s = static_cast<const uint8_t *>(p);
		for (size_t i = n; i-- > 0;) {
			buf_.push_back(s[i]);
		}
	}
	// all "offsets" handed out are distances from the END of the buffer (== size() at that time)
	uint32_t create_string(const std::string &s) {
		align(4, s.size() + 1);
		buf_.push_back(0);
		push_bytes(s.data(), s.size());
		push<uint32_t>(static_cast<uint32_t>(s.size()));
		return static_cast<uint32_t>(buf_.size());
	}
	// vector of raw structs / scalars (elem bytes already little-endian packed)
	uint32_t create_struct_vector(const void *data, size_t n, size_t elem_size, size_t elem_align) {
		align(elem_align > 4 ? elem_align : 4, n * elem_size);
		// after the length prefix (4 bytes) the data must be elem_align aligned: handled because the
		// buffer end is kept aligned to min_align_ at finish and offsets are measured from the end.
		push_bytes(data, n * elem_size);
		push<uint32_t>(static_cast<uint32_t>(n));
		return static_cast<uint32_t>(buf_.size());
	}
	uint32_t create_offset_vector(const std::vector<uint32_t> &offs) {
		align(4, offs.size() * 4);
		for (size_t i = offs.size(); i-- > 0;) {
			align(4);
			push<uint32_t>(static_cast<uint32_t>(buf_.size() + 4 - offs[i]));
		}
		push<uint32_t>(static_cast<uint32_t>(offs.size()));
		return static_cast<uint32_t>(buf_.size());
	}

	struct FieldLoc {
		unsigned id;
		uint32_t off; // distance from buffer end of the stored field
	};
	void start_table() {
		fields_.clear();
		table_start_ = buf_.size();
	}
	template <class T>
	void add_scalar(unsigned id, T v, T def) {
		if (v == def && !force_defaults_) {
			return;
		}
		align(sizeof(T));
		push<T>(v);
		fields_.push_back({id, static_cast<uint32_t>(buf_.size())});
	}
	template <class T>
	void add_scalar_force(unsigned id, T v) {
		align(sizeof(T));
		push<T>(v);
		fields_.push_back({id, static_cast<uint32_t>(buf_.size())});
	}
	void add_offset(unsigned id, uint32_t target) {
		if (!target) {
			return;
		}
		align(4);
		push<uint32_t>(static_cast<uint32_t>(buf_.size() + 4 - target));
		fields_.push_back({id, static_cast<uint32_t>(buf_.size())});
	}
	uint32_t end_table() {
		align(4);
		// soffset placeholder
		push<int32_t>(0);
		uint32_t table_off = static_cast<uint32_t>(buf_.size());
		unsigned max_id = 0;
		for (auto &f : fields_) {
			if (f.id + 1 > max_id) {
				max_id = f.id + 1;
			}
		}
		uint16_t vtsize = static_cast<uint16_t>(4 + 2 * max_id);
		uint16_t tsize = static_cast<uint16_t>(table_off - table_start_);
		std::vector<uint16_t> slots(max_id, 0);
		for (auto &f : fields_) {
			slots[f.id] = static_cast<uint16_t>(table_off - f.off);
		}
		for (size_t i = slots.size(); i-- > 0;) {
			push<uint16_t>(slots[i]);
		}
		push<uint16_t>(tsize);
		push<uint16_t>(vtsize);
		uint32_t vt_off = static_cast<uint32_t>(buf_.size());
		// patch soffset: table_pos - vtable_pos = (end-table_off) - (end-vt_off) = vt_off - table_off
		int32_t soff = static_cast<int32_t>(vt_off - table_off);
		uint8_t tmp[4];
		std::memcpy(tmp, &soff, 4);
		size_t base = table_off; // reversed index of the first byte of the soffset is table_off-1
		for (int i = 0; i < 4; i++) {
			buf_[base - 1 - i] = tmp[i];
		}
		return table_off;
	}
	// finish: returns the forward buffer, root table at `root`
	std::vector<uint8_t> finish(uint32_t root, size_t final_align = 8) {
		if (final_align < min_align_) {
			final_align = min_align_;
		}
		align(final_align, 4);
		push<uint32_t>(static_cast<uint32_t>(buf_.size() + 4 - root));
		std::vector<uint8_t> out(buf_.rbegin(), buf_.rend());
		return out;
	}

private:
	std::vector<uint8_t> buf_;
	std::vector<FieldLoc> fields_;
	size_t table_start_ = 0;
	size_t min_align_ = 4;
	bool force_defaults_ = false;
};

} // namespace fb
} // namespace b2a
