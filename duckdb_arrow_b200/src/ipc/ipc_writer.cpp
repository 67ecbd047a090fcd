#include "ipc_writer.hpp"

#include "flatbuf.hpp"

#include <cstring>

namespace b2a {

namespace {

// Schema.fbs enum values
enum : int16_t { UNIT_SECOND = 0, UNIT_MILLI = 1, UNIT_MICRO = 2, UNIT_NANO = 3 };
enum : int16_t { DATE_DAY = 0, DATE_MILLI = 1 };
enum : int16_t { INTERVAL_YEAR_MONTH = 0, INTERVAL_DAY_TIME = 1, INTERVAL_MONTH_DAY_NANO = 2 };
enum : int16_t { PREC_HALF = 0, PREC_SINGLE = 1, PREC_DOUBLE = 2 };

int DecimalBitWidth(int width) {
	// DECIMAL physical type: <=4 int16, <=9 int32 (both written as decimal32), <=18 int64, else int128
	// (duckdb/src/common/arrow/arrow_converter.cpp:234-256)
	return width <= 9 ? 32 : width <= 18 ? 64 : 128;
}

struct TypeOut {
	ArrowTypeTag tag;
	uint32_t table; // builder offset of the type table
};

TypeOut BuildType(fb::Builder &b, const WriteField &f, uint32_t tz_string) {
	TypeOut o {ArrowTypeTag::NONE, 0};
	auto int_type = [&](int bits, bool sgn) {
		b.start_table();
		b.add_scalar<int32_t>(0, bits, 0);
		b.add_scalar<uint8_t>(1, sgn ? 1 : 0, 0);
		o.tag = ArrowTypeTag::Int;
		o.table = b.end_table();
	};
	auto empty = [&](ArrowTypeTag tag) {
		b.start_table();
		o.tag = tag;
		o.table = b.end_table();
	};
	auto ts = [&](int16_t unit) {
		b.start_table();
		b.add_scalar<int16_t>(0, unit, UNIT_SECOND);
		b.add_offset(1, tz_string);
		o.tag = ArrowTypeTag::Timestamp;
		o.table = b.end_table();
	};
	auto fixed_binary = [&](int32_t width) {
		b.start_table();
		b.add_scalar<int32_t>(0, width, 0);
		o.tag = ArrowTypeTag::FixedSizeBinary;
		o.table = b.end_table();
	};
	if (f.lossless) { // the storage types of the extension types, arrow_type_extension.cpp:366-379
		switch (f.type_id) {
		case B2A_T_BOOLEAN:
			int_type(8, true);
			return o;
		case B2A_T_HUGEINT:
		case B2A_T_UUID:
			fixed_binary(16);
			return o;
		case B2A_T_TIME_TZ:
			fixed_binary(8);
			return o;
		default:
			break;
		}
	}
	switch (f.type_id) {
	case B2A_T_SQLNULL:
		empty(ArrowTypeTag::Null);
		break;
	case B2A_T_BOOLEAN:
		empty(ArrowTypeTag::Bool);
		break;
	case B2A_T_TINYINT:
		int_type(8, true);
		break;
	case B2A_T_SMALLINT:
		int_type(16, true);
		break;
	case B2A_T_INTEGER:
		int_type(32, true);
		break;
	case B2A_T_BIGINT:
		int_type(64, true);
		break;
	case B2A_T_UTINYINT:
		int_type(8, false);
		break;
	case B2A_T_USMALLINT:
		int_type(16, false);
		break;
	case B2A_T_UINTEGER:
		int_type(32, false);
		break;
	case B2A_T_UBIGINT:
		int_type(64, false);
		break;
	case B2A_T_FLOAT:
	case B2A_T_DOUBLE:
		b.start_table();
		b.add_scalar<int16_t>(0, f.type_id == B2A_T_FLOAT ? PREC_SINGLE : PREC_DOUBLE, PREC_HALF);
		o.tag = ArrowTypeTag::FloatingPoint;
		o.table = b.end_table();
		break;
	case B2A_T_HUGEINT:
	case B2A_T_DECIMAL: {
		const int width = f.type_id == B2A_T_HUGEINT ? 38 : f.decimal_width;
		const int scale = f.type_id == B2A_T_HUGEINT ? 0 : f.decimal_scale;
		b.start_table();
		b.add_scalar<int32_t>(0, width, 0);
		b.add_scalar<int32_t>(1, scale, 0);
		b.add_scalar<int32_t>(2, f.type_id == B2A_T_HUGEINT ? 128 : DecimalBitWidth(width), 128);
		o.tag = ArrowTypeTag::Decimal;
		o.table = b.end_table();
		break;
	}
	case B2A_T_UHUGEINT:
		fixed_binary(16);
		break;
	case B2A_T_ENUM: // a dictionary-encoded field carries the VALUE type; the index type sits in Field.dictionary
		empty(ArrowTypeTag::Utf8);
		break;
	case B2A_T_VARCHAR:
	case B2A_T_JSON:
		empty(f.string_view ? ArrowTypeTag::Utf8View : f.large_offsets ? ArrowTypeTag::LargeUtf8 : ArrowTypeTag::Utf8);
		break;
	case B2A_T_BLOB:
	case B2A_T_VARINT:
	case B2A_T_BIT: // arrow_converter.cpp:268-279 without arrow_lossless_conversion: plain binary
		empty(f.large_offsets ? ArrowTypeTag::LargeBinary : ArrowTypeTag::Binary);
		break;
	case B2A_T_UUID: // arrow_converter.cpp:172-186: a 36-character string unless arrow_lossless_conversion
		empty(f.large_offsets ? ArrowTypeTag::LargeUtf8 : ArrowTypeTag::Utf8);
		break;
	case B2A_T_UNION: { // "+us:0,1,..": sparse, type id i = member i (arrow_converter.cpp:331-361)
		std::vector<int32_t> ids(f.children.size());
		for (size_t i = 0; i < ids.size(); i++) {
			ids[i] = static_cast<int32_t>(i);
		}
		uint32_t ids_vec = b.create_struct_vector(ids.data(), ids.size(), 4, 4);
		b.start_table();
		b.add_scalar<int16_t>(0, 0, 0); // UnionMode::Sparse
		b.add_offset(1, ids_vec);
		o.tag = ArrowTypeTag::Union;
		o.table = b.end_table();
		break;
	}
	case B2A_T_DATE:
		b.start_table();
		b.add_scalar<int16_t>(0, DATE_DAY, DATE_MILLI);
		o.tag = ArrowTypeTag::Date;
		o.table = b.end_table();
		break;
	case B2A_T_TIME:
	case B2A_T_TIME_TZ: // arrow_converter.cpp:202-209: the offset is dropped, "ttu"
		b.start_table();
		b.add_scalar<int16_t>(0, UNIT_MICRO, UNIT_MILLI);
		b.add_scalar<int32_t>(1, 64, 32);
		o.tag = ArrowTypeTag::Time;
		o.table = b.end_table();
		break;
	case B2A_T_TIMESTAMP:
	case B2A_T_TIMESTAMP_TZ:
		ts(UNIT_MICRO);
		break;
	case B2A_T_TIMESTAMP_SEC:
		ts(UNIT_SECOND);
		break;
	case B2A_T_TIMESTAMP_MS:
		ts(UNIT_MILLI);
		break;
	case B2A_T_TIMESTAMP_NS:
		ts(UNIT_NANO);
		break;
	case B2A_T_INTERVAL:
		b.start_table();
		b.add_scalar<int16_t>(0, INTERVAL_MONTH_DAY_NANO, INTERVAL_YEAR_MONTH);
		o.tag = ArrowTypeTag::Interval;
		o.table = b.end_table();
		break;
	case B2A_T_LIST:
		empty(f.list_view ? (f.large_offsets ? ArrowTypeTag::LargeListView : ArrowTypeTag::ListView)
		                  : (f.large_offsets ? ArrowTypeTag::LargeList : ArrowTypeTag::List));
		break;
	case B2A_T_STRUCT:
		empty(ArrowTypeTag::Struct_);
		break;
	case B2A_T_MAP:
		b.start_table();
		b.add_scalar<uint8_t>(0, 0, 0); // keysSorted = false
		o.tag = ArrowTypeTag::Map;
		o.table = b.end_table();
		break;
	case B2A_T_ARRAY:
		b.start_table();
		b.add_scalar<int32_t>(0, static_cast<int32_t>(f.param), 0);
		o.tag = ArrowTypeTag::FixedSizeList;
		o.table = b.end_table();
		break;
	default:
		throw Error(B2A_ENOTIMPL, "Unsupported type in DuckDB -> Arrow Conversion: type id " +
		                              std::to_string(f.type_id));
	}
	return o;
}

// The extension a field is written as: {ARROW:extension:name, ARROW:extension:metadata}; empty name = none.
// arrow_converter.cpp:118-128 (alias types), :130-279 (the arrow_lossless_conversion branches), :387-393 (types
// without a case of their own); names and metadata of arrow_type_extension.cpp:366-390, schema_metadata.cpp:58-73.
std::pair<std::string, std::string> ExtensionOf(const WriteField &f) {
	auto opaque = [](const char *type_name) {
		return std::make_pair(std::string("arrow.opaque"),
		                      std::string("{\"type_name\":\"") + type_name + "\",\"vendor_name\":\"DuckDB\"}");
	};
	switch (f.type_id) {
	case B2A_T_UHUGEINT:
		return opaque("uhugeint");
	case B2A_T_VARINT:
		return opaque("varint");
	default:
		break;
	}
	if (!f.lossless) {
		return {};
	}
	switch (f.type_id) {
	case B2A_T_BOOLEAN:
		return {"arrow.bool8", ""};
	case B2A_T_UUID:
		return {"arrow.uuid", ""};
	case B2A_T_JSON:
		return {"arrow.json", ""};
	case B2A_T_HUGEINT:
		return opaque("hugeint");
	case B2A_T_TIME_TZ:
		return opaque("time_tz");
	case B2A_T_BIT:
		return opaque("bit");
	default:
		return {};
	}
}

uint32_t BuildField(fb::Builder &b, const WriteField &f) {
	std::vector<uint32_t> kids;
	for (auto &c : f.children) {
		kids.push_back(BuildField(b, c));
	}
	uint32_t kids_vec = b.create_offset_vector(kids); // Field.children is always present in arrow-written files
	uint32_t tz = 0;
	if (f.type_id == B2A_T_TIMESTAMP_TZ) {
		tz = b.create_string(f.timezone.empty() ? std::string("UTC") : f.timezone);
	}
	TypeOut ty = BuildType(b, f, tz);
	uint32_t custom = 0;
	const auto ext = ExtensionOf(f);
	if (!ext.first.empty()) { // Field.custom_metadata, in the order ArrowSchemaMetadata::SerializeMetadata emits
		std::vector<uint32_t> kvs;
		const std::pair<const char *, const std::string *> entries[2] = {{"ARROW:extension:metadata", &ext.second},
		                                                                {"ARROW:extension:name", &ext.first}};
		for (auto &e : entries) {
			uint32_t k = b.create_string(e.first);
			uint32_t v = b.create_string(*e.second);
			b.start_table();
			b.add_offset(0, k);
			b.add_offset(1, v);
			kvs.push_back(b.end_table());
		}
		custom = b.create_offset_vector(kvs);
	}
	uint32_t dict = 0;
	if (f.dictionary_id >= 0) { // DictionaryEncoding{id, indexType: Int, isOrdered = false}
		b.start_table();
		b.add_scalar<int32_t>(0, f.index_bits, 0);
		b.add_scalar<uint8_t>(1, 0, 0); // unsigned: "C" / "S" / "I" (arrow_converter.cpp:364-376)
		uint32_t index_type = b.end_table();
		b.start_table();
		b.add_scalar_force<int64_t>(0, f.dictionary_id);
		b.add_offset(1, index_type);
		dict = b.end_table();
	}
	uint32_t name = b.create_string(f.name);
	b.start_table();
	b.add_offset(0, name);
	b.add_scalar<uint8_t>(1, f.nullable ? 1 : 0, 0);
	b.add_scalar<uint8_t>(2, static_cast<uint8_t>(ty.tag), 0);
	b.add_offset(3, ty.table);
	if (dict) {
		b.add_offset(4, dict);
	}
	b.add_offset(5, kids_vec);
	if (custom) {
		b.add_offset(6, custom);
	}
	return b.end_table();
}

std::vector<uint8_t> Encapsulate(const std::vector<uint8_t> &fbuf) {
	// prefix (8 bytes) + metadata padded so that the total is a multiple of 8
	size_t padded = (fbuf.size() + 7) & ~size_t(7);
	std::vector<uint8_t> out(8 + padded, 0);
	uint32_t cont = 0xFFFFFFFFu;
	int32_t size = static_cast<int32_t>(padded);
	std::memcpy(out.data(), &cont, 4);
	std::memcpy(out.data() + 4, &size, 4);
	std::memcpy(out.data() + 8, fbuf.data(), fbuf.size());
	return out;
}

std::vector<uint8_t> FinishMessage(fb::Builder &b, uint8_t header_type, uint32_t header, int64_t body_length) {
	b.start_table();
	b.add_scalar<int16_t>(0, 4, 0); // MetadataVersion::V5
	b.add_scalar<uint8_t>(1, header_type, 0);
	b.add_offset(2, header);
	b.add_scalar_force<int64_t>(3, body_length);
	uint32_t msg = b.end_table();
	return Encapsulate(b.finish(msg, 8));
}

} // namespace

void ResolveArrowFormat(WriteField &f) {
	auto kids = [&](size_t n) {
		if (f.children.size() != n) {
			throw Error(B2A_EINVAL, "writer: field '" + f.name + "' has " + std::to_string(f.children.size()) +
			                            " children, expected " + std::to_string(n));
		}
	};
	if (f.lossless) {
		switch (f.type_id) {
		case B2A_T_BOOLEAN:
			f.format = "c";
			return;
		case B2A_T_HUGEINT:
		case B2A_T_UUID:
			f.format = "w:16";
			return;
		case B2A_T_TIME_TZ:
			f.format = "w:8";
			return;
		default:
			break;
		}
	}
	switch (f.type_id) {
	case B2A_T_SQLNULL:
		f.format = "n";
		break;
	case B2A_T_BOOLEAN:
		f.format = "b";
		break;
	case B2A_T_TINYINT:
		f.format = "c";
		break;
	case B2A_T_SMALLINT:
		f.format = "s";
		break;
	case B2A_T_INTEGER:
		f.format = "i";
		break;
	case B2A_T_BIGINT:
		f.format = "l";
		break;
	case B2A_T_UTINYINT:
		f.format = "C";
		break;
	case B2A_T_USMALLINT:
		f.format = "S";
		break;
	case B2A_T_UINTEGER:
		f.format = "I";
		break;
	case B2A_T_UBIGINT:
		f.format = "L";
		break;
	case B2A_T_FLOAT:
		f.format = "f";
		break;
	case B2A_T_DOUBLE:
		f.format = "g";
		break;
	case B2A_T_HUGEINT:
		f.format = "d:38,0";
		break;
	case B2A_T_DECIMAL:
		if (f.decimal_width < 1 || f.decimal_width > 38) {
			throw Error(B2A_EINVAL, "writer: DECIMAL width out of range");
		}
		f.format = "d:" + std::to_string(f.decimal_width) + "," + std::to_string(f.decimal_scale) + "," +
		           std::to_string(DecimalBitWidth(f.decimal_width));
		break;
	case B2A_T_UHUGEINT:
		f.format = "w:16";
		break;
	case B2A_T_ENUM:
		if (f.dictionary_id < 0 || (f.index_bits != 8 && f.index_bits != 16 && f.index_bits != 32)) {
			throw Error(B2A_EINVAL, "writer: ENUM field '" + f.name + "' without a dictionary");
		}
		f.format = f.index_bits == 8 ? "C" : f.index_bits == 16 ? "S" : "I";
		break;
	case B2A_T_VARCHAR:
	case B2A_T_JSON:
	case B2A_T_UUID:
		f.format = f.string_view ? "vu" : f.large_offsets ? "U" : "u";
		break;
	case B2A_T_BLOB:
	case B2A_T_VARINT:
	case B2A_T_BIT:
		f.format = f.large_offsets ? "Z" : "z";
		break;
	case B2A_T_UNION: {
		if (f.children.empty()) {
			throw Error(B2A_EINVAL, "writer: UNION without members");
		}
		f.format = "+us:";
		for (size_t i = 0; i < f.children.size(); i++) {
			f.format += (i ? "," : "") + std::to_string(i);
		}
		break;
	}
	case B2A_T_DATE:
		f.format = "tdD";
		break;
	case B2A_T_TIME:
	case B2A_T_TIME_TZ:
		f.format = "ttu";
		break;
	case B2A_T_TIMESTAMP:
		f.format = "tsu:";
		break;
	case B2A_T_TIMESTAMP_TZ:
		f.format = "tsu:" + (f.timezone.empty() ? std::string("UTC") : f.timezone);
		break;
	case B2A_T_TIMESTAMP_SEC:
		f.format = "tss:";
		break;
	case B2A_T_TIMESTAMP_MS:
		f.format = "tsm:";
		break;
	case B2A_T_TIMESTAMP_NS:
		f.format = "tsn:";
		break;
	case B2A_T_INTERVAL:
		f.format = "tin";
		break;
	case B2A_T_LIST:
		kids(1);
		f.format = f.list_view ? (f.large_offsets ? "+vL" : "+vl") : (f.large_offsets ? "+L" : "+l");
		f.children[0].name = "l"; // arrow_converter.cpp:330
		break;
	case B2A_T_STRUCT:
		if (f.children.empty()) {
			throw Error(B2A_EINVAL, "writer: STRUCT without fields");
		}
		f.format = "+s";
		break;
	case B2A_T_ARRAY:
		kids(1);
		if (f.param <= 0) {
			throw Error(B2A_EINVAL, "writer: ARRAY size must be positive");
		}
		f.format = "+w:" + std::to_string(f.param);
		f.children[0].name = "";
		break;
	case B2A_T_MAP: {
		kids(1);
		WriteField &entries = f.children[0];
		if (entries.type_id != B2A_T_STRUCT || entries.children.size() != 2) {
			throw Error(B2A_EINVAL, "writer: MAP needs one STRUCT(key, value) child");
		}
		f.format = "+m";
		entries.name = "entries"; // SetArrowMapFormat, arrow_converter.cpp:91-106
		entries.nullable = false;
		entries.children[0].nullable = false; // keys
		break;
	}
	default:
		throw Error(B2A_ENOTIMPL,
		            "Unsupported type in DuckDB -> Arrow Conversion: type id " + std::to_string(f.type_id));
	}
	for (auto &c : f.children) {
		ResolveArrowFormat(c);
	}
}

std::vector<uint8_t> EncodeSchemaMessage(const std::vector<WriteField> &fields,
                                         const std::vector<std::pair<std::string, std::string>> &kv_metadata) {
	fb::Builder b;
	std::vector<uint32_t> kv;
	for (auto &p : kv_metadata) {
		uint32_t v = b.create_string(p.second);
		uint32_t k = b.create_string(p.first);
		b.start_table();
		b.add_offset(0, k);
		b.add_offset(1, v);
		kv.push_back(b.end_table());
	}
	uint32_t kv_vec = kv.empty() ? 0 : b.create_offset_vector(kv);
	std::vector<uint32_t> fs;
	for (auto &f : fields) {
		fs.push_back(BuildField(b, f));
	}
	uint32_t fields_vec = b.create_offset_vector(fs);
	b.start_table();
	b.add_scalar<int16_t>(0, 0, 0); // Endianness::Little
	b.add_offset(1, fields_vec);
	b.add_offset(2, kv_vec);
	uint32_t schema = b.end_table();
	return FinishMessage(b, 1 /* MessageHeader::Schema */, schema, 0);
}

std::vector<uint8_t> EncodeRecordBatchMessage(uint64_t rows, const std::vector<NodeSlice> &nodes,
                                              const std::vector<BufferSlice> &buffers, uint64_t body_length,
                                              size_t n_variadic) {
	fb::Builder b;
	// RecordBatch.variadicBufferCounts (field 4): one entry per view-typed field in pre-order, each with its ONE data buffer
	const std::vector<int64_t> counts(n_variadic, 1);
	uint32_t var_vec = n_variadic ? b.create_struct_vector(counts.data(), counts.size(), 8, 8) : 0;
	static_assert(sizeof(NodeSlice) == 16 && sizeof(BufferSlice) == 16, "FieldNode / Buffer structs are 16 bytes");
	uint32_t buf_vec = b.create_struct_vector(buffers.data(), buffers.size(), 16, 8);
	uint32_t node_vec = b.create_struct_vector(nodes.data(), nodes.size(), 16, 8);
	b.start_table();
	b.add_scalar_force<int64_t>(0, static_cast<int64_t>(rows));
	b.add_offset(1, node_vec);
	b.add_offset(2, buf_vec);
	if (var_vec) {
		b.add_offset(4, var_vec);
	}
	uint32_t rb = b.end_table();
	return FinishMessage(b, 3 /* MessageHeader::RecordBatch */, rb, static_cast<int64_t>(body_length));
}

std::vector<uint8_t> EncodeDictionaryBatchMessage(int64_t id, uint64_t rows, const std::vector<NodeSlice> &nodes,
                                                  const std::vector<BufferSlice> &buffers, uint64_t body_length) {
	fb::Builder b;
	uint32_t buf_vec = b.create_struct_vector(buffers.data(), buffers.size(), 16, 8);
	uint32_t node_vec = b.create_struct_vector(nodes.data(), nodes.size(), 16, 8);
	b.start_table();
	b.add_scalar_force<int64_t>(0, static_cast<int64_t>(rows));
	b.add_offset(1, node_vec);
	b.add_offset(2, buf_vec);
	uint32_t rb = b.end_table();
	b.start_table();
	b.add_scalar_force<int64_t>(0, id);
	b.add_offset(1, rb);
	uint32_t db = b.end_table();
	return FinishMessage(b, 2 /* MessageHeader::DictionaryBatch */, db, static_cast<int64_t>(body_length));
}

size_t DictionaryBatchMessageSize(size_t n_nodes, size_t n_buffers) {
	std::vector<NodeSlice> nodes(n_nodes, NodeSlice {1, 1});
	std::vector<BufferSlice> buffers(n_buffers, BufferSlice {8, 8});
	return EncodeDictionaryBatchMessage(1, 1, nodes, buffers, 8).size();
}

size_t RecordBatchMessageSize(size_t n_nodes, size_t n_buffers, size_t n_variadic) {
	// scalar fields are elided when they equal their default, so size the message with non-default values
	std::vector<NodeSlice> nodes(n_nodes, NodeSlice {1, 1});
	std::vector<BufferSlice> buffers(n_buffers, BufferSlice {8, 8});
	return EncodeRecordBatchMessage(1, nodes, buffers, 8, n_variadic).size();
}

std::vector<uint8_t> EncodeEndOfStream() {
	std::vector<uint8_t> out(8, 0);
	uint32_t cont = 0xFFFFFFFFu;
	std::memcpy(out.data(), &cont, 4);
	return out;
}

} // namespace b2a
