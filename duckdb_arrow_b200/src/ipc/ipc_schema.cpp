// Schema message -> Field tree -> DuckDB type + kernel op.
// Type rules restate duckdb/src/function/table/arrow/arrow_duck_schema.cpp:59-297 (GetTypeFromFormat)
// via the Arrow C-data format string, which is what nanoarrow's ArrowIpcDecoderDecodeSchema would
// have handed to DuckDB (reference call site: src/ipc/stream_reader/base_stream_reader.cpp:52-74).
#include "ipc_schema.hpp"
#include "flatbuf.hpp"

#include <cstdio>

namespace b2a {

int TypeWidth(int t) {
	switch (t) {
	case B2A_T_BOOLEAN:
	case B2A_T_TINYINT:
	case B2A_T_UTINYINT:
		return 1;
	case B2A_T_SMALLINT:
	case B2A_T_USMALLINT:
		return 2;
	case B2A_T_INTEGER:
	case B2A_T_UINTEGER:
	case B2A_T_FLOAT:
	case B2A_T_DATE:
		return 4;
	case B2A_T_BIGINT:
	case B2A_T_UBIGINT:
	case B2A_T_DOUBLE:
	case B2A_T_TIME:
	case B2A_T_TIMESTAMP:
	case B2A_T_TIMESTAMP_SEC:
	case B2A_T_TIMESTAMP_MS:
	case B2A_T_TIMESTAMP_NS:
	case B2A_T_TIMESTAMP_TZ:
	case B2A_T_TIME_TZ:
		return 8;
	case B2A_T_HUGEINT:
	case B2A_T_UHUGEINT:
	case B2A_T_UUID:
	case B2A_T_INTERVAL:
	case B2A_T_VARCHAR:
	case B2A_T_BLOB:
	case B2A_T_LIST:
	case B2A_T_MAP:
		return 16;
	default:
		return 0;
	}
}

static void SetCopy(Field &f, int type_id) {
	f.type_id = type_id;
	f.op = B2A_OP_COPY;
	f.out_width = f.src_width = TypeWidth(type_id);
	f.n_buffers = 2;
}

static const char *TimeUnitChar(int unit) {
	static const char *u[] = {"s", "m", "u", "n"};
	if (unit < 0 || unit > 3) {
		throw Error(B2A_EINVAL, "Arrow IPC: invalid time unit");
	}
	return u[unit];
}

static std::string MetaValue(const Field &f, const char *key) {
	for (auto &kv : f.metadata) {
		if (kv.first == key) {
			return kv.second;
		}
	}
	return std::string();
}

// physical decimal type: width<=4 int16, <=9 int32, <=18 int64, <=38 hugeint
// (duckdb/src/common/types.cpp DecimalType / arrow_converter.cpp:234-256)
int DecimalPhysicalWidth(int width) {
	if (width <= 4) {
		return 2;
	}
	if (width <= 9) {
		return 4;
	}
	if (width <= 18) {
		return 8;
	}
	return 16;
}

static void MapStorageToDuckDB(Field &f);

// value of a string member of a flat JSON object ('{"type_name":"hugeint","vendor_name":"DuckDB"}'); "" when absent
static std::string JsonMember(const std::string &json, const std::string &key) {
	const std::string needle = "\"" + key + "\"";
	size_t p = json.find(needle);
	if (p == std::string::npos) {
		return std::string();
	}
	p = json.find(':', p + needle.size());
	if (p == std::string::npos) {
		return std::string();
	}
	p = json.find('"', p);
	if (p == std::string::npos) {
		return std::string();
	}
	std::string out;
	for (p++; p < json.size() && json[p] != '"'; p++) {
		if (json[p] == '\\' && p + 1 < json.size()) {
			p++;
		}
		out += json[p];
	}
	return out;
}

static const char *IndexFormat(int type_id) {
	switch (type_id) {
	case B2A_T_TINYINT: return "c";
	case B2A_T_SMALLINT: return "s";
	case B2A_T_INTEGER: return "i";
	case B2A_T_BIGINT: return "l";
	case B2A_T_UTINYINT: return "C";
	case B2A_T_USMALLINT: return "S";
	case B2A_T_UINTEGER: return "I";
	default: return "L";
	}
}

// ArrowType::GetTypeFromSchema (arrow_duck_schema.cpp:372-386) + the default registrations
// (arrow_type_extension.cpp:366-390): a field with ARROW:extension:name whose {name, vendor, type, format} — or, for the
// 1:n types json / bit / varint, {name, vendor, type} — is registered gets the extension's type; anything else keeps the
// storage type.  For dictionary-encoded fields the reference looks the extension up with the INDEX format.
static void ApplyExtension(Field &f) {
	const std::string name = MetaValue(f, "ARROW:extension:name");
	if (name.empty()) {
		return;
	}
	const std::string meta = MetaValue(f, "ARROW:extension:metadata");
	const std::string vendor = JsonMember(meta, "vendor_name"), type_name = JsonMember(meta, "type_name");
	const std::string fmt = f.dict_id >= 0 ? IndexFormat(f.index_type_id) : f.format;
	const bool duck = name == "arrow.opaque" && vendor == "DuckDB";
	if (name == "arrow.json" && vendor.empty() && type_name.empty()) {
		if (fmt != "u" && fmt != "U" && fmt != "vu") {
			throw Error(B2A_EINVAL, "Arrow extension type \"" + fmt + "\" not supported for arrow.json");
		}
		f.type_id = B2A_T_JSON;
		return;
	}
	if (duck && (type_name == "bit" || type_name == "varint")) {
		if (fmt != "z" && fmt != "Z") {
			throw Error(B2A_EINVAL, "Arrow extension type \"" + fmt + "\" not supported for " +
			                            (type_name == "bit" ? "BIT type" : "Varint"));
		}
		f.type_id = type_name == "bit" ? B2A_T_BIT : B2A_T_VARINT;
		return;
	}
	if (f.dict_id >= 0) {
		return; // the 1:1 extensions are registered under their storage format, which an index format never matches
	}
	if (name == "arrow.uuid" && vendor.empty() && type_name.empty() && fmt == "w:16") {
		f.type_id = B2A_T_UUID; // conversion: UUIDConversion, arrow_conversion.cpp:480-495
		f.op = B2A_OP_UUID;
		f.src_width = 16;
		f.out_width = 16;
	} else if (name == "arrow.bool8" && vendor.empty() && type_name.empty() && fmt == "c") {
		f.type_id = B2A_T_BOOLEAN; // storage converted as TINYINT, then ArrowBool8::ArrowToDuck (:344-350)
		f.op = B2A_OP_BOOL8;
	} else if (duck && (type_name == "hugeint" || type_name == "uhugeint") && fmt == "w:16") {
		SetCopy(f, type_name == "hugeint" ? B2A_T_HUGEINT : B2A_T_UHUGEINT); // DirectConversion, :902-911
	} else if (duck && type_name == "time_tz" && fmt == "w:8") {
		SetCopy(f, B2A_T_TIME_TZ);
	}
}

void MapFieldToDuckDB(Field &f) {
	MapStorageToDuckDB(f);
	ApplyExtension(f);
}

static void MapStorageToDuckDB(Field &f) {
	// `f.format` and the tag-specific attributes (param, decimal_*, src_width) are set by the parser
	const std::string &fmt = f.format;
	f.op = B2A_OP_NONE;
	f.out_width = 0;
	if (fmt == "n") {
		f.type_id = B2A_T_SQLNULL;
		f.n_buffers = 0;
	} else if (fmt == "b") {
		f.type_id = B2A_T_BOOLEAN;
		f.op = B2A_OP_BOOL;
		f.out_width = 1;
		f.src_width = 0;
		f.n_buffers = 2;
	} else if (fmt == "c") {
		SetCopy(f, B2A_T_TINYINT);
	} else if (fmt == "s") {
		SetCopy(f, B2A_T_SMALLINT);
	} else if (fmt == "i") {
		SetCopy(f, B2A_T_INTEGER);
	} else if (fmt == "l") {
		SetCopy(f, B2A_T_BIGINT);
	} else if (fmt == "C") {
		SetCopy(f, B2A_T_UTINYINT);
	} else if (fmt == "S") {
		SetCopy(f, B2A_T_USMALLINT);
	} else if (fmt == "I") {
		SetCopy(f, B2A_T_UINTEGER);
	} else if (fmt == "L") {
		SetCopy(f, B2A_T_UBIGINT);
	} else if (fmt == "f") {
		SetCopy(f, B2A_T_FLOAT);
	} else if (fmt == "g") {
		SetCopy(f, B2A_T_DOUBLE);
	} else if (fmt == "e") {
		throw Error(B2A_ENOTIMPL, "Unsupported Internal Arrow Type e");
	} else if (fmt[0] == 'd') {
		int bitwidth = f.src_width * 8;
		if (f.decimal_width > 38 || bitwidth > 128) {
			throw Error(B2A_ENOTIMPL, "Unsupported Internal Arrow Type for Decimal " + fmt);
		}
		f.type_id = B2A_T_DECIMAL;
		f.out_width = DecimalPhysicalWidth(f.decimal_width);
		// matching width => DirectConversion (arrow_conversion.cpp:786-790,804-808,822-826)
		f.op = (f.out_width == f.src_width) ? B2A_OP_COPY : B2A_OP_DECIMAL;
		f.n_buffers = 2;
	} else if (fmt == "u" || fmt == "z") {
		f.type_id = fmt == "u" ? B2A_T_VARCHAR : B2A_T_BLOB;
		f.op = B2A_OP_STRING32;
		f.out_width = 16;
		f.src_width = 4;
		f.n_buffers = 3;
	} else if (fmt == "U" || fmt == "Z") {
		f.type_id = fmt == "U" ? B2A_T_VARCHAR : B2A_T_BLOB;
		f.op = B2A_OP_STRING64;
		f.out_width = 16;
		f.src_width = 8;
		f.n_buffers = 3;
	} else if (fmt == "vu") {
		f.type_id = B2A_T_VARCHAR;
		f.op = B2A_OP_STRING_VIEW;
		f.out_width = 16;
		f.src_width = 16;
		f.n_buffers = 2;
		f.variadic = true;
	} else if (fmt == "vz") {
		throw Error(B2A_ENOTIMPL, "Unsupported Internal Arrow Type vz");
	} else if (fmt[0] == 'w') {
		f.type_id = B2A_T_BLOB;
		f.op = B2A_OP_BLOB_FIXED;
		f.out_width = 16;
		f.src_width = static_cast<int>(f.param);
		f.n_buffers = 2;
	} else if (fmt == "tsn:") {
		SetCopy(f, B2A_T_TIMESTAMP_NS);
	} else if (fmt == "tsu:") {
		SetCopy(f, B2A_T_TIMESTAMP);
	} else if (fmt == "tsm:") {
		SetCopy(f, B2A_T_TIMESTAMP_MS);
	} else if (fmt == "tss:") {
		SetCopy(f, B2A_T_TIMESTAMP_SEC);
	} else if (fmt == "tdD") {
		SetCopy(f, B2A_T_DATE);
	} else if (fmt == "tdm") {
		f.type_id = B2A_T_DATE;
		f.op = B2A_OP_DATE64;
		f.out_width = 4;
		f.src_width = 8;
		f.n_buffers = 2;
	} else if (fmt == "tts" || fmt == "ttm") {
		f.type_id = B2A_T_TIME;
		f.op = B2A_OP_MUL_I32;
		f.param = fmt == "tts" ? 1000000 : 1000;
		f.out_width = 8;
		f.src_width = 4;
		f.n_buffers = 2;
	} else if (fmt == "ttu") {
		f.type_id = B2A_T_TIME;
		f.op = B2A_OP_MUL_I64;
		f.param = 1;
		f.out_width = f.src_width = 8;
		f.n_buffers = 2;
	} else if (fmt == "ttn") {
		f.type_id = B2A_T_TIME;
		f.op = B2A_OP_DIV_I64;
		f.param = 1000;
		f.out_width = f.src_width = 8;
		f.n_buffers = 2;
	} else if (fmt == "tDs" || fmt == "tDm" || fmt == "tDu") {
		f.type_id = B2A_T_INTERVAL;
		f.op = B2A_OP_INTERVAL_MUL;
		f.param = fmt == "tDs" ? 1000000 : (fmt == "tDm" ? 1000 : 1);
		f.out_width = 16;
		f.src_width = 8;
		f.n_buffers = 2;
	} else if (fmt == "tDn") {
		f.type_id = B2A_T_INTERVAL;
		f.op = B2A_OP_INTERVAL_DIV;
		f.param = 1000;
		f.out_width = 16;
		f.src_width = 8;
		f.n_buffers = 2;
	} else if (fmt == "tiD") {
		// ArrowDateTimeType::DAYS shares the MILLISECONDS branch: int64 * 1000 (:1041-1046)
		f.type_id = B2A_T_INTERVAL;
		f.op = B2A_OP_INTERVAL_MUL;
		f.param = 1000;
		f.out_width = 16;
		f.src_width = 8;
		f.n_buffers = 2;
	} else if (fmt == "tiM") {
		f.type_id = B2A_T_INTERVAL;
		f.op = B2A_OP_INTERVAL_MONTHS;
		f.out_width = 16;
		f.src_width = 4;
		f.n_buffers = 2;
	} else if (fmt == "tin") {
		f.type_id = B2A_T_INTERVAL;
		f.op = B2A_OP_INTERVAL_MDN;
		f.out_width = 16;
		f.src_width = 16;
		f.n_buffers = 2;
	} else if (fmt.size() >= 4 && fmt[0] == 't' && fmt[1] == 's') {
		f.type_id = B2A_T_TIMESTAMP_TZ;
		f.out_width = f.src_width = 8;
		f.n_buffers = 2;
		switch (fmt[2]) {
		case 'n':
			f.op = B2A_OP_DIV_I64;
			f.param = 1000;
			break;
		case 'u':
			f.op = B2A_OP_COPY;
			break;
		case 'm':
			f.op = B2A_OP_MUL_I64;
			f.param = 1000;
			break;
		case 's':
			f.op = B2A_OP_MUL_I64;
			f.param = 1000000;
			break;
		default:
			throw Error(B2A_ENOTIMPL, " Timestamptz precision of not accepted");
		}
	} else if (fmt == "+l" || fmt == "+L") {
		f.type_id = B2A_T_LIST;
		f.op = fmt == "+l" ? B2A_OP_LIST32 : B2A_OP_LIST64;
		f.out_width = 16;
		f.src_width = fmt == "+l" ? 4 : 8;
		f.n_buffers = 2;
	} else if (fmt == "+m") {
		f.type_id = B2A_T_MAP;
		f.op = B2A_OP_LIST32;
		f.out_width = 16;
		f.src_width = 4;
		f.n_buffers = 2;
	} else if (fmt == "+vl" || fmt == "+vL") {
		// arrow_duck_schema.cpp: ArrowListInfo::ListView; buffers = validity, offsets, sizes
		f.type_id = B2A_T_LIST;
		f.op = fmt == "+vl" ? B2A_OP_LISTVIEW32 : B2A_OP_LISTVIEW64;
		f.out_width = 16;
		f.src_width = fmt == "+vl" ? 4 : 8;
		f.n_buffers = 3;
	} else if (fmt.size() > 2 && fmt[0] == '+' && fmt[1] == 'w') {
		f.type_id = B2A_T_ARRAY;
		f.n_buffers = 1;
	} else if (fmt == "+s") {
		if (f.children.empty()) {
			throw Error(B2A_EINVAL, "Attempted to convert a STRUCT with no fields to DuckDB which is not supported");
		}
		f.type_id = B2A_T_STRUCT;
		f.n_buffers = 1;
	} else if (fmt.size() > 2 && fmt[0] == '+' && fmt[1] == 'u') {
		if (fmt[2] != 's') {
			throw Error(B2A_ENOTIMPL, std::string("Unsupported Internal Arrow Type: \"") + fmt[2] + "\" Union");
		}
		// arrow_duck_schema.cpp:235-259: LogicalType::UNION(members); the type ids of the format string are ignored by the
		// reference (":242 TODO: what are these type ids actually for?") and the type_id byte is used as the member index
		if (f.children.empty()) {
			throw Error(B2A_EINVAL, "Attempted to convert a UNION with no fields to DuckDB which is not supported");
		}
		if (f.children.size() > 256) {
			throw Error(B2A_EINVAL, "Arrow IPC: a union has at most 256 members");
		}
		for (auto &c : f.children) {
			// the reference boxes every member value (GetValue / SetValue, :1216-1217); the device path handles the flat
			// members that boxing leaves bit-identical
			const bool flat = c->children.empty() && c->dict_id < 0 && c->op != B2A_OP_REE && c->op != B2A_OP_STRING_VIEW &&
			                  c->type_id != B2A_T_SQLNULL && c->type_id != B2A_T_BLOB;
			if (!flat) {
				throw Error(B2A_ENOTIMPL, "Arrow IPC: union member '" + c->name + "' of type '" + c->format +
				                              "' is not supported by the B200 scan path (flat members only)");
			}
		}
		f.type_id = B2A_T_UNION;
		f.op = B2A_OP_UNION;
		f.out_width = 1; // the tag
		f.src_width = 1;
		f.n_buffers = 1; // MetadataVersion V5: type ids only
	} else if (fmt == "+r") {
		// arrow_duck_schema.cpp:276-297: the DuckDB type is the type of the `values` child; run ends are int16/32/64
		if (f.children.size() != 2) {
			throw Error(B2A_EINVAL, "Arrow IPC: a run-end encoded field needs exactly two children");
		}
		const Field &re = *f.children[0], &vals = *f.children[1];
		if (re.type_id != B2A_T_SMALLINT && re.type_id != B2A_T_INTEGER && re.type_id != B2A_T_BIGINT) {
			throw Error(B2A_ENOTIMPL, "Type '" + re.format + "' not implemented for RunEndEncoding");
		}
		// FlattenRunEndsSwitch (arrow_conversion.cpp:665-719): flat physical value types only
		const bool flat_value = vals.children.empty() && vals.dict_id < 0 && vals.op != B2A_OP_REE &&
		                        (vals.out_width == 1 || vals.out_width == 2 || vals.out_width == 4 ||
		                         vals.out_width == 8 || vals.out_width == 16) &&
		                        vals.type_id != B2A_T_BLOB && vals.type_id != B2A_T_UHUGEINT && vals.type_id != B2A_T_UUID;
		if (!flat_value) {
			throw Error(B2A_ENOTIMPL, "RunEndEncoded value type '" + vals.format + "' not supported yet");
		}
		f.type_id = vals.type_id;
		f.op = B2A_OP_REE;
		f.out_width = vals.out_width;
		// run ends are read as int16 or int32; int64 run ends are read through an int32 pointer, like the reference
		// does (arrow_conversion.cpp:760-762 instantiates FlattenRunEndsSwitch<int32_t> for INT64)
		f.src_width = re.type_id == B2A_T_SMALLINT ? 2 : 4;
		f.decimal_width = vals.decimal_width;
		f.decimal_scale = vals.decimal_scale;
		f.n_buffers = 0;
	} else {
		throw Error(B2A_ENOTIMPL, "Unsupported Internal Arrow Type " + fmt);
	}
}

// ------------------------------------------------------------------------------------------
// flatbuffer Field -> Field
// ------------------------------------------------------------------------------------------
static void ParseKeyValues(const fb::Table &t, unsigned id, std::vector<std::pair<std::string, std::string>> &out) {
	size_t data;
	uint32_t n = t.vec(id, data, 4);
	for (uint32_t i = 0; i < n; i++) {
		auto kv = t.vec_table(data, i);
		out.emplace_back(kv.str(0), kv.str(1));
	}
}

static void IntTypeInfo(const fb::Table &ty, int &type_id, int &width, std::string &fmt) {
	int bits = ty.scalar<int32_t>(0, 0);
	bool is_signed = ty.scalar<uint8_t>(1, 0) != 0;
	switch (bits) {
	case 8:
		type_id = is_signed ? B2A_T_TINYINT : B2A_T_UTINYINT;
		fmt = is_signed ? "c" : "C";
		break;
	case 16:
		type_id = is_signed ? B2A_T_SMALLINT : B2A_T_USMALLINT;
		fmt = is_signed ? "s" : "S";
		break;
	case 32:
		type_id = is_signed ? B2A_T_INTEGER : B2A_T_UINTEGER;
		fmt = is_signed ? "i" : "I";
		break;
	case 64:
		type_id = is_signed ? B2A_T_BIGINT : B2A_T_UBIGINT;
		fmt = is_signed ? "l" : "L";
		break;
	default:
		throw Error(B2A_EINVAL, "Arrow IPC: invalid Int bitWidth " + std::to_string(bits));
	}
	width = bits / 8;
}

static std::unique_ptr<Field> ParseField(const fb::Table &t, int depth) {
	if (depth > 64) {
		throw Error(B2A_EINVAL, "Arrow IPC: schema nesting too deep");
	}
	auto f = std::unique_ptr<Field>(new Field());
	f->name = t.str(0);
	f->nullable = t.scalar<uint8_t>(1, 0) != 0;
	f->tag = static_cast<ArrowTypeTag>(t.scalar<uint8_t>(2, 0));
	fb::Table ty = t.table(3);
	ParseKeyValues(t, 6, f->metadata);

	size_t cdata;
	uint32_t nchild = t.vec(5, cdata, 4);
	for (uint32_t i = 0; i < nchild; i++) {
		f->children.push_back(ParseField(t.vec_table(cdata, i), depth + 1));
	}

	char buf[96];
	switch (f->tag) {
	case ArrowTypeTag::Null:
		f->format = "n";
		break;
	case ArrowTypeTag::Int: {
		int tid, w;
		IntTypeInfo(ty, tid, w, f->format);
		break;
	}
	case ArrowTypeTag::FloatingPoint: {
		int prec = ty.valid ? ty.scalar<int16_t>(0, 0) : 0;
		f->format = prec == 0 ? "e" : (prec == 1 ? "f" : "g");
		break;
	}
	case ArrowTypeTag::Binary:
		f->format = "z";
		break;
	case ArrowTypeTag::Utf8:
		f->format = "u";
		break;
	case ArrowTypeTag::LargeBinary:
		f->format = "Z";
		break;
	case ArrowTypeTag::LargeUtf8:
		f->format = "U";
		break;
	case ArrowTypeTag::BinaryView:
		f->format = "vz";
		break;
	case ArrowTypeTag::Utf8View:
		f->format = "vu";
		break;
	case ArrowTypeTag::Bool:
		f->format = "b";
		break;
	case ArrowTypeTag::Decimal: {
		int precision = ty.scalar<int32_t>(0, 0);
		int scale = ty.scalar<int32_t>(1, 0);
		int bits = ty.scalar<int32_t>(2, 128);
		f->decimal_width = precision;
		f->decimal_scale = scale;
		f->src_width = bits / 8;
		if (bits == 128) {
			snprintf(buf, sizeof(buf), "d:%d,%d", precision, scale);
		} else {
			snprintf(buf, sizeof(buf), "d:%d,%d,%d", precision, scale, bits);
		}
		f->format = buf;
		break;
	}
	case ArrowTypeTag::Date: {
		int unit = ty.valid ? ty.scalar<int16_t>(0, 1) : 1;
		f->format = unit == 0 ? "tdD" : "tdm";
		break;
	}
	case ArrowTypeTag::Time: {
		int unit = ty.valid ? ty.scalar<int16_t>(0, 1) : 1;
		f->format = std::string("tt") + TimeUnitChar(unit);
		break;
	}
	case ArrowTypeTag::Timestamp: {
		int unit = ty.valid ? ty.scalar<int16_t>(0, 0) : 0;
		std::string tz = ty.valid ? ty.str(1) : std::string();
		f->format = std::string("ts") + TimeUnitChar(unit) + ":" + tz;
		break;
	}
	case ArrowTypeTag::Duration: {
		int unit = ty.valid ? ty.scalar<int16_t>(0, 1) : 1;
		f->format = std::string("tD") + TimeUnitChar(unit);
		break;
	}
	case ArrowTypeTag::Interval: {
		int unit = ty.valid ? ty.scalar<int16_t>(0, 0) : 0;
		f->format = unit == 0 ? "tiM" : (unit == 1 ? "tiD" : "tin");
		break;
	}
	case ArrowTypeTag::List:
		f->format = "+l";
		break;
	case ArrowTypeTag::LargeList:
		f->format = "+L";
		break;
	case ArrowTypeTag::ListView:
		f->format = "+vl";
		break;
	case ArrowTypeTag::LargeListView:
		f->format = "+vL";
		break;
	case ArrowTypeTag::Struct_:
		f->format = "+s";
		break;
	case ArrowTypeTag::Map:
		f->format = "+m";
		break;
	case ArrowTypeTag::FixedSizeBinary: {
		int w = ty.scalar<int32_t>(0, 0);
		f->param = w;
		snprintf(buf, sizeof(buf), "w:%d", w);
		f->format = buf;
		break;
	}
	case ArrowTypeTag::FixedSizeList: {
		int n = ty.scalar<int32_t>(0, 0);
		f->param = n;
		snprintf(buf, sizeof(buf), "+w:%d", n);
		f->format = buf;
		break;
	}
	case ArrowTypeTag::Union: {
		int mode = ty.valid ? ty.scalar<int16_t>(0, 0) : 0;
		f->format = mode == 0 ? "+us:" : "+ud:";
		break;
	}
	case ArrowTypeTag::RunEndEncoded:
		f->format = "+r";
		break;
	default:
		throw Error(B2A_EINVAL, "Arrow IPC: unknown Type tag " + std::to_string(static_cast<int>(f->tag)));
	}

	// dictionary: the field's own type is the VALUE type; indices come from DictionaryEncoding.indexType
	fb::Table dict = t.table(4);
	if (dict.valid) {
		f->dict_id = dict.scalar<int64_t>(0, 0);
		fb::Table it = dict.table(1);
		std::string ifmt;
		if (it.valid) {
			IntTypeInfo(it, f->index_type_id, f->index_width, ifmt);
		} else {
			f->index_type_id = B2A_T_INTEGER; // signed int32 is the default index type
			f->index_width = 4;
		}
	}
	MapFieldToDuckDB(*f);
	return f;
}

void Schema::Flatten() {
	flat.clear();
	top_flat.clear();
	top_count.clear();
	struct Rec {
		static void Go(Schema &s, Field &f, int parent) {
			f.flat_index = static_cast<int>(s.flat.size());
			f.parent_flat = parent;
			s.flat.push_back(&f);
			for (auto &c : f.children) {
				Go(s, *c, f.flat_index);
			}
		}
	};
	for (auto &f : fields) {
		int start = static_cast<int>(flat.size());
		Rec::Go(*this, *f, -1);
		top_flat.push_back(start);
		top_count.push_back(static_cast<int>(flat.size()) - start);
	}
}

void SchemaParser::ParseMessageSchema(const uint8_t *meta, size_t meta_len, Schema &out) {
	try {
		fb::Table msg = fb::root(meta, meta_len);
		uint8_t header_type = msg.scalar<uint8_t>(1, 0);
		if (header_type != 1) {
			throw Error(B2A_EIO, "Expected Schema Arrow IPC message");
		}
		fb::Table sch = msg.table(2);
		if (!sch.valid) {
			throw Error(B2A_EIO, "Arrow IPC: Schema message without header");
		}
		out.little_endian = sch.scalar<int16_t>(0, 0) == 0;
		size_t fdata;
		uint32_t nf = sch.vec(1, fdata, 4);
		for (uint32_t i = 0; i < nf; i++) {
			out.fields.push_back(ParseField(sch.vec_table(fdata, i), 0));
		}
		ParseKeyValues(sch, 2, out.metadata);
		out.Flatten();
	} catch (fb::ParseError &e) {
		throw Error(B2A_EIO, std::string("Arrow IPC: invalid Schema flatbuffer: ") + e.what());
	}
}

} // namespace b2a
