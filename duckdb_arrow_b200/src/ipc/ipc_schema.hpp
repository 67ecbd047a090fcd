// Schema model produced by the host slicer: the Arrow field tree with the DuckDB type each node
// maps to (rules of duckdb/src/function/table/arrow/arrow_duck_schema.cpp:59-297) and the kernel op
// that converts its value buffer.
#pragma once
#include "b2arrow.h"

#include <cstdint>
#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace b2a {

struct Status {
	int code = B2A_OK;
	std::string msg;
	bool ok() const {
		return code == B2A_OK;
	}
};

struct Error {
	int code;
	std::string msg;
	Error(int c, std::string m) : code(c), msg(std::move(m)) {
	}
};

// Arrow flatbuffer Type union tags (format/Schema.fbs)
enum class ArrowTypeTag : uint8_t {
	NONE = 0,
	Null = 1,
	Int = 2,
	FloatingPoint = 3,
	Binary = 4,
	Utf8 = 5,
	Bool = 6,
	Decimal = 7,
	Date = 8,
	Time = 9,
	Timestamp = 10,
	Interval = 11,
	List = 12,
	Struct_ = 13,
	Union = 14,
	FixedSizeBinary = 15,
	FixedSizeList = 16,
	Map = 17,
	Duration = 18,
	LargeBinary = 19,
	LargeUtf8 = 20,
	LargeList = 21,
	RunEndEncoded = 22,
	BinaryView = 23,
	Utf8View = 24,
	ListView = 25,
	LargeListView = 26
};

struct Field {
	std::string name;
	std::string format; // C data interface format string
	bool nullable = true;
	ArrowTypeTag tag = ArrowTypeTag::NONE;
	std::vector<std::pair<std::string, std::string>> metadata;
	std::vector<std::unique_ptr<Field>> children;

	// dictionary encoding (Field.dictionary)
	int64_t dict_id = -1;
	int index_type_id = B2A_T_INVALID;
	int index_width = 0;

	// DuckDB mapping
	int type_id = B2A_T_INVALID;
	int op = B2A_OP_NONE;
	int out_width = 0;
	int src_width = 0;
	int n_buffers = 0;
	int decimal_width = 0;
	int decimal_scale = 0;
	int64_t param = 0;
	bool variadic = false; // view types: buffer count comes from variadicBufferCounts

	// flattened pre-order index (assigned by Schema::Flatten)
	int flat_index = -1;
	int parent_flat = -1;
};

struct Schema {
	bool little_endian = true;
	std::vector<std::unique_ptr<Field>> fields;
	std::vector<std::pair<std::string, std::string>> metadata;

	// pre-order view of all nodes
	std::vector<Field *> flat;
	// for each top-level column: its flat index and the number of nodes below it (incl. itself)
	std::vector<int> top_flat;
	std::vector<int> top_count;

	void Flatten();
};

// Parse a Schema flatbuffer table (the header of a Schema message).  Throws b2a::Error.
struct SchemaParser {
	static void ParseMessageSchema(const uint8_t *meta, size_t meta_len, Schema &out);
};

// Computes type_id/op/widths/format for a field whose tag-specific attributes have been read.
void MapFieldToDuckDB(Field &f);

int TypeWidth(int type_id);
int DecimalPhysicalWidth(int decimal_width); // bytes of DuckDB's physical type for DECIMAL(width, .)

} // namespace b2a
