// Host-side Arrow IPC message encoder: Schema message, RecordBatch message headers, end-of-stream marker.
// Replaces what the reference gets from nanoarrow's encoder (ArrowIpcEncoderEncodeSchema /
// EncodeSimpleRecordBatch / FinalizeBuffer, call sites src/writer/column_data_collection_serializer.cpp:57-95)
// and ArrowStreamWriter::{WriteSchema,Finalize} (src/writer/arrow_stream_writer.cpp:55-82).  Bodies are NOT built
// here: the encode kernels write every buffer at the offset this header announces.
#pragma once
#include "ipc_reader.hpp"

#include <string>
#include <utility>
#include <vector>

namespace b2a {

// Arrow type of one written field, derived from the DuckDB LogicalType by the rules of
// duckdb/src/common/arrow/arrow_converter.cpp:115-432 (SetArrowFormat) with default ClientProperties
// (regular 32-bit offsets, no lossless conversion, no string views, no list views).
struct WriteField {
	std::string name;
	bool nullable = true;
	int type_id = B2A_T_INVALID; // DuckDB type
	int decimal_width = 0, decimal_scale = 0;
	int64_t param = 0; // ARRAY size
	std::string timezone;
	bool string_view = false;   // produce_arrow_string_view: "vu"
	bool list_view = false;     // arrow_output_list_view: "+vl" / "+vL"
	bool large_offsets = false; // arrow_large_buffer_size: "U" / "Z" / "+L" (set by the write planner)
	bool lossless = false;      // arrow_lossless_conversion: extension types (set by the write planner)
	int64_t dictionary_id = -1; // ENUM: id of its DictionaryBatch (set by the write planner)
	int index_bits = 0;         // ENUM: 8 / 16 / 32, unsigned ("C" / "S" / "I")
	std::vector<WriteField> children;
	std::string format; // the C-data-interface format string the reference would set (for tests / diagnostics)
};

// Fills `format` (recursively) and validates that the type is writable; throws b2a::Error(B2A_ENOTIMPL) otherwise.
void ResolveArrowFormat(WriteField &f);

// Encapsulated messages: 0xFFFFFFFF, int32 metadata size (padded so the body starts 8-byte aligned), flatbuffer.
std::vector<uint8_t> EncodeSchemaMessage(const std::vector<WriteField> &fields,
                                         const std::vector<std::pair<std::string, std::string>> &kv_metadata);
std::vector<uint8_t> EncodeRecordBatchMessage(uint64_t rows, const std::vector<NodeSlice> &nodes,
                                              const std::vector<BufferSlice> &buffers, uint64_t body_length,
                                              size_t n_variadic = 0);
// size of a RecordBatch message header with the given node / buffer counts (independent of the values)
size_t RecordBatchMessageSize(size_t n_nodes, size_t n_buffers, size_t n_variadic = 0);
// DictionaryBatch{id, data: RecordBatch, isDelta = false}: the values of one dictionary (here always one utf8 column:
// validity, int32 offsets, characters); the reference's appender builds it in ArrowEnumData::Initialize
// (appender/enum_data.hpp:67-75), the message form is Arrow's (Message.header_type = DictionaryBatch)
std::vector<uint8_t> EncodeDictionaryBatchMessage(int64_t id, uint64_t rows, const std::vector<NodeSlice> &nodes,
                                                  const std::vector<BufferSlice> &buffers, uint64_t body_length);
size_t DictionaryBatchMessageSize(size_t n_nodes, size_t n_buffers);
// ArrowStreamWriter::Finalize: FFFFFFFF 00000000
std::vector<uint8_t> EncodeEndOfStream();

} // namespace b2a
