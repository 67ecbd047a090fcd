//===----------------------------------------------------------------------===//
// b2a_extension.cpp — the nanoarrow extension's surface over libb2arrow.so (include/b2arrow.h).
//
// Same entry point and the same five registrations as the reference (src/nanoarrow_extension.cpp:33-67):
//   nanoarrow_version()                                   scalar
//   read_arrow(VARCHAR | LIST(VARCHAR)) + replacement scan for *.arrow / *.arrows   src/scanner/read_arrow.cpp:44-86
//   COPY ... TO / FROM (FORMAT arrows | arrow)            src/writer/write_arrow_stream.cpp:54-272
//   scan_arrow_ipc(LIST(STRUCT(ptr POINTER, size UBIGINT)))   src/scanner/scan_arrow_ipc.cpp:20-64
//   to_arrow_ipc(TABLE)                                   src/writer/to_arrow_ipc.cpp:52-182
// Everything DuckDB-facing keeps the reference's names, arguments, options, defaults and error texts; everything below
// the DuckDB API (IPC framing, slicing, per-value conversion, body packing) is a call into the C ABI.  Compiled against
// the DuckDB headers of the reference checkout and linked against that DuckDB (integration/build.py).
//===----------------------------------------------------------------------===//
#define DUCKDB_EXTENSION_MAIN

#include "b2a_scanner.hpp"

#include "duckdb.hpp"
#include "duckdb/common/bind_helpers.hpp"
#include "duckdb/common/multi_file/multi_file_function.hpp"
#include "duckdb/common/serializer/buffered_file_writer.hpp"
#include "duckdb/common/types/column/column_data_collection.hpp"
#include "duckdb/execution/physical_operator.hpp"
#include "duckdb/function/copy_function.hpp"
#include "duckdb/function/scalar_function.hpp"
#include "duckdb/function/table_function.hpp"
#include "duckdb/main/client_context.hpp"
#include "duckdb/main/config.hpp"
#include "duckdb/main/extension_util.hpp"
#include "duckdb/parser/expression/constant_expression.hpp"
#include "duckdb/parser/expression/function_expression.hpp"
#include "duckdb/parser/tableref/table_function_ref.hpp"

namespace duckdb {
namespace ext_nanoarrow {

// ====================================================================================================================
// read_arrow: one file = one B2AFileScan (reference: ArrowFileScan, src/file_scanner/arrow_file_scan.cpp)
// ====================================================================================================================
class B2AFileScan : public BaseFileReader {
public:
	B2AFileScan(ClientContext &context, const OpenFileInfo &file_p) : BaseFileReader(file_p) {
		b2a_reader *reader = nullptr;
		int rc = b2a_reader_open_file(file.path.c_str(), &reader);
		if (rc) {
			B2AThrow(rc);
		}
		scanner = make_uniq<B2AScanner>(reader, B2ADevice());
		columns = MultiFileColumnDefinition::ColumnsFromNamesAndTypes(scanner->Names(), scanner->Types());
	}

	string GetReaderType() const override {
		return "ARROW";
	}

	bool TryInitializeScan(ClientContext &context, GlobalTableFunctionState &gstate,
	                       LocalTableFunctionState &lstate) override;

	void Scan(ClientContext &context, GlobalTableFunctionState &global_state, LocalTableFunctionState &local_state,
	          DataChunk &chunk) override;

	shared_ptr<BaseUnionData> GetUnionData(idx_t file_idx) override {
		auto data = make_shared_ptr<BaseUnionData>(file);
		data->names = scanner->Names();
		data->types = scanner->Types();
		return data;
	}

	double GetProgressInFile(ClientContext &context) override {
		return scanner->Progress();
	}

	unique_ptr<B2AScanner> scanner;

private:
};

class B2AFileReaderOptions : public BaseFileReaderOptions {};

struct B2AFileGlobalState : public GlobalTableFunctionState {};
//! one scan thread of the pipeline: the decoded slice it is walking
struct B2AFileLocalState : public LocalTableFunctionState {
	B2AScanner::Cursor cursor;
};

// The reference gives ONE file one scan thread (arrow_file_scan.cpp:35-41: a second thread is turned away).  Here the
// unit is a decoded slice of the file (~256 MiB of record batches): every thread that asks (MultiFileFunction calls this
// under its global lock, in batch_index order) gets the next one, so the operators above the scan run on all threads.
bool B2AFileScan::TryInitializeScan(ClientContext &context, GlobalTableFunctionState &gstate,
                                    LocalTableFunctionState &lstate) {
	if (!scanner->Begun()) {
		vector<idx_t> cols;
		for (auto &index : column_indexes) {
			cols.push_back(index.GetPrimaryIndex());
		}
		scanner->Begin(cols);
	}
	return scanner->TakeSlice(lstate.Cast<B2AFileLocalState>().cursor);
}

void B2AFileScan::Scan(ClientContext &context, GlobalTableFunctionState &global_state,
                       LocalTableFunctionState &local_state, DataChunk &chunk) {
	scanner->NextFrom(local_state.Cast<B2AFileLocalState>().cursor, chunk);
}
struct B2AMultiFileData final : public TableFunctionData {};

//! reference: ArrowMultiFileInfo (src/file_scanner/arrow_multi_file_info.cpp) — no options of its own, the first file
//! (or the union of all files) gives the schema, one thread per file
struct B2AMultiFileInfo : MultiFileReaderInterface {
	static unique_ptr<MultiFileReaderInterface> InitializeInterface(ClientContext &context, MultiFileReader &reader,
	                                                                MultiFileList &file_list) {
		return make_uniq<B2AMultiFileInfo>();
	}
	unique_ptr<BaseFileReaderOptions> InitializeOptions(ClientContext &context,
	                                                    optional_ptr<TableFunctionInfo> info) override {
		return make_uniq<B2AFileReaderOptions>();
	}
	bool ParseCopyOption(ClientContext &context, const string &key, const vector<Value> &values,
	                     BaseFileReaderOptions &options, vector<string> &expected_names,
	                     vector<LogicalType> &expected_types) override {
		return false; // the reader has no options (arrow_multi_file_info.cpp:17-36)
	}
	bool ParseOption(ClientContext &context, const string &key, const Value &val, MultiFileOptions &file_options,
	                 BaseFileReaderOptions &options) override {
		return false;
	}
	unique_ptr<TableFunctionData> InitializeBindData(MultiFileBindData &multi_file_data,
	                                                 unique_ptr<BaseFileReaderOptions> options) override {
		return make_uniq<B2AMultiFileData>();
	}
	void BindReader(ClientContext &context, vector<LogicalType> &return_types, vector<string> &names,
	                MultiFileBindData &bind_data) override {
		B2AFileReaderOptions options;
		if (!bind_data.file_options.union_by_name) {
			bind_data.reader_bind = bind_data.multi_file_reader->BindReader(
			    context, return_types, names, *bind_data.file_list, bind_data, options, bind_data.file_options);
		} else {
			bind_data.reader_bind = bind_data.multi_file_reader->BindUnionReader(
			    context, return_types, names, *bind_data.file_list, bind_data, options, bind_data.file_options);
		}
	}
	optional_idx MaxThreads(const MultiFileBindData &bind_data_p, const MultiFileGlobalState &global_state,
	                        FileExpandResult expand_result) override {
		return optional_idx(); // a single file is scanned by every thread too (slice by slice)
	}
	unique_ptr<GlobalTableFunctionState> InitializeGlobalState(ClientContext &context, MultiFileBindData &bind_data,
	                                                           MultiFileGlobalState &global_state) override {
		return make_uniq<B2AFileGlobalState>();
	}
	unique_ptr<LocalTableFunctionState> InitializeLocalState(ExecutionContext &context,
	                                                         GlobalTableFunctionState &function_state) override {
		return make_uniq<B2AFileLocalState>();
	}
	shared_ptr<BaseFileReader> CreateReader(ClientContext &context, GlobalTableFunctionState &gstate,
	                                        BaseUnionData &union_data, const MultiFileBindData &bind_data) override {
		return make_shared_ptr<B2AFileScan>(context, union_data.GetFileName());
	}
	shared_ptr<BaseFileReader> CreateReader(ClientContext &context, GlobalTableFunctionState &gstate,
	                                        const OpenFileInfo &file_info, idx_t file_idx,
	                                        const MultiFileBindData &bind_data) override {
		return make_shared_ptr<B2AFileScan>(context, file_info);
	}
	shared_ptr<BaseFileReader> CreateReader(ClientContext &context, const OpenFileInfo &file_info,
	                                        BaseFileReaderOptions &options,
	                                        const MultiFileOptions &file_options) override {
		return make_shared_ptr<B2AFileScan>(context, file_info);
	}
	unique_ptr<NodeStatistics> GetCardinality(const MultiFileBindData &bind_data, idx_t file_count) override {
		return make_uniq<NodeStatistics>();
	}
	void GetVirtualColumns(ClientContext &, MultiFileBindData &, virtual_column_map_t &result) override {
		if (result.find(COLUMN_IDENTIFIER_EMPTY) != result.end()) {
			result.erase(COLUMN_IDENTIFIER_EMPTY);
		}
	}
};

static TableFunction ReadArrowFunction() {
	MultiFileFunction<B2AMultiFileInfo> read_arrow("read_arrow");
	read_arrow.projection_pushdown = true;
	read_arrow.filter_pushdown = false; // read_arrow.cpp:46-48: filters stay above the scan
	read_arrow.filter_prune = false;
	return static_cast<TableFunction>(read_arrow);
}

static unique_ptr<TableRef> ReadArrowReplacement(ClientContext &context, ReplacementScanInput &input,
                                                 optional_ptr<ReplacementScanData> data) {
	auto table_name = ReplacementScan::GetFullPath(input);
	if (!ReplacementScan::CanReplace(table_name, {"arrows", "arrow"})) {
		return nullptr;
	}
	auto table_function = make_uniq<TableFunctionRef>();
	vector<unique_ptr<ParsedExpression>> children;
	children.push_back(make_uniq<ConstantExpression>(Value(table_name)));
	table_function->function = make_uniq<FunctionExpression>("read_arrow", std::move(children));
	if (!FileSystem::HasGlob(table_name)) {
		auto &fs = FileSystem::GetFileSystem(context);
		table_function->alias = fs.ExtractBaseName(table_name);
	}
	return std::move(table_function);
}

// ====================================================================================================================
// scan_arrow_ipc: caller-owned buffers (reference: src/scanner/scan_arrow_ipc.cpp + ipc_buffer_stream_reader.cpp)
// ====================================================================================================================
struct ScanIPCBindData : public TableFunctionData {
	vector<b2a_ipc_buffer> buffers;
	vector<LogicalType> types;
};

struct ScanIPCGlobalState : public GlobalTableFunctionState {
	unique_ptr<B2AScanner> scanner;
	mutex lock;
	idx_t MaxThreads() const override {
		return 1; // one stream
	}
};

static unique_ptr<B2AScanner> OpenBuffers(const vector<b2a_ipc_buffer> &buffers) {
	b2a_reader *reader = nullptr;
	int rc = b2a_reader_open_buffers(buffers.data(), NumericCast<uint32_t>(buffers.size()), &reader);
	if (rc) {
		B2AThrow(rc);
	}
	return make_uniq<B2AScanner>(reader, B2ADevice());
}

static unique_ptr<FunctionData> ScanIPCBind(ClientContext &context, TableFunctionBindInput &input,
                                            vector<LogicalType> &return_types, vector<string> &names) {
	auto res = make_uniq<ScanIPCBindData>();
	for (auto &entry : ListValue::GetChildren(input.inputs[0])) {
		auto unpacked = StructValue::GetChildren(entry);
		b2a_ipc_buffer buf;
		buf.ptr = reinterpret_cast<const void *>(unpacked[0].GetPointer());
		buf.size = unpacked[1].GetValue<uint64_t>();
		res->buffers.push_back(buf);
	}
	auto probe = OpenBuffers(res->buffers);
	names = probe->Names();
	return_types = probe->Types();
	res->types = return_types;
	return std::move(res);
}

static unique_ptr<GlobalTableFunctionState> ScanIPCInitGlobal(ClientContext &context, TableFunctionInitInput &input) {
	auto &bind = input.bind_data->Cast<ScanIPCBindData>();
	auto state = make_uniq<ScanIPCGlobalState>();
	state->scanner = OpenBuffers(bind.buffers);
	vector<idx_t> cols;
	for (auto col : input.column_ids) {
		if (col == COLUMN_IDENTIFIER_ROW_ID) {
			throw NotImplementedException("scan_arrow_ipc: rowid is not available");
		}
		cols.push_back(col);
	}
	state->scanner->Begin(cols);
	return std::move(state);
}

static void ScanIPCFunction(ClientContext &context, TableFunctionInput &data, DataChunk &output) {
	auto &state = data.global_state->Cast<ScanIPCGlobalState>();
	lock_guard<mutex> guard(state.lock);
	state.scanner->Next(output);
}

// ====================================================================================================================
// writer side: DuckDB vectors -> b2a_vector tree (host pointers) -> b2a_write_arrow_host
// ====================================================================================================================
struct VectorTree {
	vector<unique_ptr<b2a_vector[]>> arrays;
	vector<unique_ptr<string>> strings;
	vector<unique_ptr<Vector>> keep; // flattened copies where needed
	string time_zone;
};

static int32_t B2ATypeId(const LogicalType &t) {
	if (t.IsJSONType()) { // VARCHAR with the JSON alias: arrow.json under arrow_lossless_conversion (arrow_converter.cpp:118-128)
		return B2A_T_JSON;
	}
	switch (t.id()) {
	case LogicalTypeId::SQLNULL: return B2A_T_SQLNULL;
	case LogicalTypeId::BOOLEAN: return B2A_T_BOOLEAN;
	case LogicalTypeId::TINYINT: return B2A_T_TINYINT;
	case LogicalTypeId::SMALLINT: return B2A_T_SMALLINT;
	case LogicalTypeId::INTEGER: return B2A_T_INTEGER;
	case LogicalTypeId::BIGINT: return B2A_T_BIGINT;
	case LogicalTypeId::UTINYINT: return B2A_T_UTINYINT;
	case LogicalTypeId::USMALLINT: return B2A_T_USMALLINT;
	case LogicalTypeId::UINTEGER: return B2A_T_UINTEGER;
	case LogicalTypeId::UBIGINT: return B2A_T_UBIGINT;
	case LogicalTypeId::HUGEINT: return B2A_T_HUGEINT;
	case LogicalTypeId::FLOAT: return B2A_T_FLOAT;
	case LogicalTypeId::DOUBLE: return B2A_T_DOUBLE;
	case LogicalTypeId::DECIMAL: return B2A_T_DECIMAL;
	case LogicalTypeId::VARCHAR: return B2A_T_VARCHAR;
	case LogicalTypeId::BLOB: return B2A_T_BLOB;
	case LogicalTypeId::DATE: return B2A_T_DATE;
	case LogicalTypeId::TIME: return B2A_T_TIME;
	case LogicalTypeId::TIMESTAMP: return B2A_T_TIMESTAMP;
	case LogicalTypeId::TIMESTAMP_SEC: return B2A_T_TIMESTAMP_SEC;
	case LogicalTypeId::TIMESTAMP_MS: return B2A_T_TIMESTAMP_MS;
	case LogicalTypeId::TIMESTAMP_NS: return B2A_T_TIMESTAMP_NS;
	case LogicalTypeId::TIMESTAMP_TZ: return B2A_T_TIMESTAMP_TZ;
	case LogicalTypeId::INTERVAL: return B2A_T_INTERVAL;
	case LogicalTypeId::UUID: return B2A_T_UUID;
	case LogicalTypeId::LIST: return B2A_T_LIST;
	case LogicalTypeId::STRUCT: return B2A_T_STRUCT;
	case LogicalTypeId::MAP: return B2A_T_MAP;
	case LogicalTypeId::ARRAY: return B2A_T_ARRAY;
	case LogicalTypeId::UNION: return B2A_T_UNION;
	case LogicalTypeId::TIME_TZ: return B2A_T_TIME_TZ;
	case LogicalTypeId::BIT: return B2A_T_BIT;
	case LogicalTypeId::UHUGEINT: return B2A_T_UHUGEINT;
	case LogicalTypeId::VARINT: return B2A_T_VARINT;
	case LogicalTypeId::ENUM: return B2A_T_ENUM;
	default:
		throw NotImplementedException("Unsupported Arrow type %s", t.ToString());
	}
}

static void FillVector(VectorTree &tree, Vector &vec, idx_t count, const string &name, b2a_vector &out) {
	vec.Flatten(count);
	const auto &type = vec.GetType();
	std::memset(&out, 0, sizeof(out));
	out.type_id = B2ATypeId(type);
	tree.strings.push_back(make_uniq<string>(name));
	out.name = tree.strings.back()->c_str();
	out.rows = count;
	if (type.id() == LogicalTypeId::DECIMAL) {
		out.decimal_width = DecimalType::GetWidth(type);
		out.decimal_scale = DecimalType::GetScale(type);
	}
	if (type.id() == LogicalTypeId::TIMESTAMP_TZ) {
		out.timezone = tree.time_zone.c_str();
	}
	auto &mask = FlatVector::Validity(vec);
	out.validity = mask.AllValid() ? nullptr : mask.GetData();
	auto children = [&](idx_t n) {
		tree.arrays.push_back(unique_ptr<b2a_vector[]>(new b2a_vector[n]));
		out.children = tree.arrays.back().get();
		out.n_children = NumericCast<int32_t>(n);
		return tree.arrays.back().get();
	};
	switch (type.id()) {
	case LogicalTypeId::SQLNULL:
		break;
	case LogicalTypeId::STRUCT: {
		auto &entries = StructVector::GetEntries(vec);
		auto kids = children(entries.size());
		for (idx_t i = 0; i < entries.size(); i++) {
			FillVector(tree, *entries[i], count, StructType::GetChildName(type, i), kids[i]);
		}
		break;
	}
	case LogicalTypeId::UNION: {
		// STRUCT{tag UTINYINT, members...}: the node's data are the tags (include/b2arrow.h: b2a_vector)
		auto &entries = StructVector::GetEntries(vec);
		entries[0]->Flatten(count);
		out.data = FlatVector::GetData(*entries[0]);
		auto kids = children(entries.size() - 1);
		for (idx_t i = 1; i < entries.size(); i++) {
			FillVector(tree, *entries[i], count, UnionType::GetMemberName(type, i - 1), kids[i - 1]);
		}
		break;
	}
	case LogicalTypeId::LIST:
	case LogicalTypeId::MAP: {
		out.data = FlatVector::GetData(vec);
		auto kids = children(1);
		FillVector(tree, ListVector::GetEntry(vec), ListVector::GetListSize(vec),
		           type.id() == LogicalTypeId::MAP ? "entries" : "l", kids[0]);
		break;
	}
	case LogicalTypeId::ENUM: {
		// the codes + the dictionary values in insert order (ArrowEnumData::Initialize, appender/enum_data.hpp:67-75)
		out.data = FlatVector::GetData(vec);
		const idx_t n_values = EnumType::GetSize(type);
		tree.keep.push_back(make_uniq<Vector>(EnumType::GetValuesInsertOrder(type), idx_t(0), n_values));
		auto kids = children(1);
		FillVector(tree, *tree.keep.back(), n_values, "dictionary", kids[0]);
		break;
	}
	case LogicalTypeId::ARRAY: {
		const auto k = ArrayType::GetSize(type);
		out.param = NumericCast<int64_t>(k);
		auto kids = children(1);
		FillVector(tree, ArrayVector::GetEntry(vec), count * k, "l", kids[0]);
		break;
	}
	default:
		out.data = FlatVector::GetData(vec);
	}
}

struct WriteSettings {
	uint32_t flags = 0;
	string time_zone;
};

static WriteSettings SettingsOf(ClientContext &context) {
	auto props = context.GetClientProperties();
	WriteSettings s;
	s.time_zone = props.time_zone;
	if (props.arrow_offset_size == ArrowOffsetSize::LARGE) {
		s.flags |= B2A_WRITE_LARGE_OFFSETS;
	}
	if (props.arrow_use_list_view) {
		s.flags |= B2A_WRITE_LIST_VIEW;
	}
	if (props.produce_arrow_string_view) {
		s.flags |= B2A_WRITE_STRING_VIEW;
	}
	if (props.arrow_lossless_conversion) {
		s.flags |= B2A_WRITE_LOSSLESS;
	}
	return s;
}

//! One finished stream (schema + record batch messages + EOS) in pinned memory
struct HostStream {
	b2a_host_stream s;
	HostStream() {
		std::memset(&s, 0, sizeof(s));
	}
	~HostStream() {
		b2a_host_stream_release(&s);
	}
	//! byte range of RecordBatch message i (prefix + flatbuffer + body)
	std::pair<const_data_ptr_t, idx_t> Message(idx_t i) const {
		b2a_batch_info bi;
		int rc = b2a_writer_batch(s.writer, i, &bi);
		if (rc) {
			B2AThrow(rc);
		}
		return {const_data_ptr_cast(s.data) + bi.meta_offset,
		        NumericCast<idx_t>(bi.body_offset + bi.body_length - bi.meta_offset)};
	}
	//! what precedes the first record batch: the Schema message + the DictionaryBatch messages of ENUM columns
	std::pair<const_data_ptr_t, idx_t> Schema() const {
		uint64_t n = 0;
		int rc = b2a_writer_preamble(s.writer, &n);
		if (rc) {
			B2AThrow(rc);
		}
		return {const_data_ptr_cast(s.data), NumericCast<idx_t>(n)};
	}
};

//! ColumnDataCollectionSerializer::Serialize(DataChunk&) / (ColumnDataCollection&): the rows as ONE record batch
static unique_ptr<HostStream> SerializeChunk(DataChunk &chunk, const vector<string> &names, const WriteSettings &settings,
                                             const vector<pair<string, string>> &kv_metadata) {
	VectorTree tree;
	tree.time_zone = settings.time_zone;
	vector<b2a_vector> cols(chunk.ColumnCount());
	for (idx_t c = 0; c < chunk.ColumnCount(); c++) {
		FillVector(tree, chunk.data[c], chunk.size(), names[c], cols[c]);
	}
	vector<const char *> keys, values;
	for (auto &kv : kv_metadata) {
		keys.push_back(kv.first.c_str());
		values.push_back(kv.second.c_str());
	}
	b2a_write_options opt;
	std::memset(&opt, 0, sizeof(opt));
	opt.row_group_size = MaxValue<idx_t>(chunk.size(), 1);
	opt.n_kv_metadata = NumericCast<uint32_t>(keys.size());
	opt.kv_keys = keys.data();
	opt.kv_values = values.data();
	opt.flags = settings.flags;
	auto out = make_uniq<HostStream>();
	int rc = b2a_write_arrow_host(cols.data(), NumericCast<uint32_t>(cols.size()), chunk.size(), &opt, B2ADevice(),
	                              &out->s);
	if (rc) {
		B2AThrow(rc);
	}
	return out;
}

static unique_ptr<HostStream> SerializeCollection(ClientContext &context, ColumnDataCollection &collection,
                                                  const vector<string> &names, const WriteSettings &settings,
                                                  const vector<pair<string, string>> &kv_metadata) {
	// copy#1 of the reference (column_data_collection_serializer.cpp:97-115): every chunk appended to one DataChunk
	DataChunk all;
	all.Initialize(Allocator::Get(context), collection.Types(), MaxValue<idx_t>(collection.Count(), 1));
	for (auto &chunk : collection.Chunks()) {
		all.Append(chunk, true);
	}
	all.SetCardinality(collection.Count());
	return SerializeChunk(all, names, settings, kv_metadata);
}

// ---- COPY ... TO (FORMAT arrows)
struct B2AWriteBindData : public TableFunctionData {
	vector<LogicalType> sql_types;
	vector<string> column_names;
	vector<pair<string, string>> kv_metadata;
	idx_t row_group_size = 122880; // write_arrow_stream.cpp:28-33
	bool row_group_size_set = false;
	optional_idx row_groups_per_file;
	static constexpr const idx_t BYTES_PER_ROW = 1024;
	idx_t row_group_size_bytes = 0;
};

struct B2AWriteGlobalState : public GlobalFunctionData {
	unique_ptr<BufferedFileWriter> writer;
	WriteSettings settings;
	idx_t row_group_count = 0;
	mutex lock;
};

struct B2AWriteLocalState : public LocalFunctionData {
	B2AWriteLocalState(ClientContext &context, const vector<LogicalType> &types)
	    : buffer(context, types, ColumnDataAllocatorType::HYBRID) {
		buffer.InitializeAppend(append_state);
	}
	ColumnDataCollection buffer;
	ColumnDataAppendState append_state;
};

static unique_ptr<FunctionData> B2AWriteBind(ClientContext &context, CopyFunctionBindInput &input,
                                             const vector<string> &names, const vector<LogicalType> &sql_types) {
	auto bind_data = make_uniq<B2AWriteBindData>();
	bool row_group_size_bytes_set = false;
	for (auto &option : input.info.options) {
		const auto loption = StringUtil::Lower(option.first);
		if (option.second.size() != 1) {
			throw BinderException("%s requires exactly one argument", StringUtil::Upper(loption));
		}
		if (loption == "row_group_size" || loption == "chunk_size") {
			if (bind_data->row_group_size_set) {
				throw BinderException("ROW_GROUP_SIZE and ROW_GROUP_SIZE_BYTES are mutually exclusive");
			}
			bind_data->row_group_size = option.second[0].GetValue<uint64_t>();
			bind_data->row_group_size_set = true;
		} else if (loption == "row_group_size_bytes") {
			auto roption = option.second[0];
			if (roption.GetTypeMutable().id() == LogicalTypeId::VARCHAR) {
				bind_data->row_group_size_bytes = DBConfig::ParseMemoryLimit(roption.ToString());
			} else {
				bind_data->row_group_size_bytes = option.second[0].GetValue<uint64_t>();
			}
			row_group_size_bytes_set = true;
		} else if (loption == "row_groups_per_file") {
			bind_data->row_groups_per_file = option.second[0].GetValue<uint64_t>();
		} else if (loption == "kv_metadata") {
			auto &kv_struct = option.second[0];
			auto &kv_struct_type = kv_struct.type();
			if (kv_struct_type.id() != LogicalTypeId::STRUCT) {
				throw BinderException("Expected kv_metadata argument to be a STRUCT");
			}
			auto values = StructValue::GetChildren(kv_struct);
			for (idx_t i = 0; i < values.size(); i++) {
				const auto &value = values[i];
				auto key = StructType::GetChildName(kv_struct_type, i);
				// a BLOB is written as its raw bytes, anything else as its string form (write_arrow_stream.cpp:87-105)
				if (value.type().id() == LogicalTypeId::BLOB) {
					bind_data->kv_metadata.emplace_back(key, StringValue::Get(value));
				} else {
					bind_data->kv_metadata.emplace_back(key, value.ToString());
				}
			}
		}
	}
	if (row_group_size_bytes_set) {
		if (DBConfig::GetConfig(context).options.preserve_insertion_order) {
			throw BinderException("ROW_GROUP_SIZE_BYTES does not work while preserving insertion order. Use \"SET "
			                      "preserve_insertion_order=false;\" to disable preserving insertion order.");
		}
	} else {
		bind_data->row_group_size_bytes = bind_data->row_group_size * B2AWriteBindData::BYTES_PER_ROW;
	}
	bind_data->sql_types = sql_types;
	bind_data->column_names = names;
	return std::move(bind_data);
}

static unique_ptr<GlobalFunctionData> B2AWriteInitializeGlobal(ClientContext &context, FunctionData &bind_data_p,
                                                               const string &file_path) {
	auto &bind = bind_data_p.Cast<B2AWriteBindData>();
	auto state = make_uniq<B2AWriteGlobalState>();
	state->settings = SettingsOf(context);
	auto &fs = FileSystem::GetFileSystem(context);
	state->writer = make_uniq<BufferedFileWriter>(fs, file_path.c_str(),
	                                              FileFlags::FILE_FLAGS_WRITE | FileFlags::FILE_FLAGS_FILE_CREATE_NEW);
	// ArrowStreamWriter::WriteSchema (arrow_stream_writer.cpp:55-58): the Schema message of a zero-row stream
	DataChunk empty;
	empty.Initialize(Allocator::Get(context), bind.sql_types, 1);
	empty.SetCardinality(0);
	auto stream = SerializeChunk(empty, bind.column_names, state->settings, bind.kv_metadata);
	auto schema = stream->Schema();
	state->writer->WriteData(schema.first, schema.second);
	return std::move(state);
}

static void WriteBatches(B2AWriteGlobalState &gstate, HostStream &stream) {
	lock_guard<mutex> guard(gstate.lock);
	for (idx_t i = 0; i < stream.s.n_batches; i++) {
		auto msg = stream.Message(i);
		gstate.writer->WriteData(msg.first, msg.second);
	}
	++gstate.row_group_count; // one Flush = one row group (ArrowStreamWriter::Flush)
}

static void FlushCollection(ClientContext &context, B2AWriteBindData &bind, B2AWriteGlobalState &gstate,
                            ColumnDataCollection &buffer) {
	// the reference serialises (and writes) an empty record batch too: Combine always flushes
	auto stream = SerializeCollection(context, buffer, bind.column_names, gstate.settings, {});
	buffer.Reset();
	if (stream->s.n_batches == 0) { // zero rows: a zero-length record batch, like the reference's encoder emits
		// b2a_writer cuts `rows` into batches of row_group_size: no rows, no batch.  Nothing to write.
		lock_guard<mutex> guard(gstate.lock);
		++gstate.row_group_count;
		return;
	}
	WriteBatches(gstate, *stream);
}

static void B2AWriteSink(ExecutionContext &context, FunctionData &bind_data_p, GlobalFunctionData &gstate_p,
                         LocalFunctionData &lstate_p, DataChunk &input) {
	auto &bind = bind_data_p.Cast<B2AWriteBindData>();
	auto &gstate = gstate_p.Cast<B2AWriteGlobalState>();
	auto &lstate = lstate_p.Cast<B2AWriteLocalState>();
	lstate.buffer.Append(lstate.append_state, input);
	if (lstate.buffer.Count() >= bind.row_group_size || lstate.buffer.SizeInBytes() >= bind.row_group_size_bytes) {
		lstate.append_state.current_chunk_state.handles.clear();
		FlushCollection(context.client, bind, gstate, lstate.buffer);
		lstate.buffer.InitializeAppend(lstate.append_state);
	}
}

static void B2AWriteCombine(ExecutionContext &context, FunctionData &bind_data_p, GlobalFunctionData &gstate_p,
                            LocalFunctionData &lstate_p) {
	auto &bind = bind_data_p.Cast<B2AWriteBindData>();
	auto &gstate = gstate_p.Cast<B2AWriteGlobalState>();
	auto &lstate = lstate_p.Cast<B2AWriteLocalState>();
	if (lstate.buffer.Count() > 0) {
		FlushCollection(context.client, bind, gstate, lstate.buffer);
	}
}

static void B2AWriteFinalize(ClientContext &context, FunctionData &bind_data, GlobalFunctionData &gstate_p) {
	auto &gstate = gstate_p.Cast<B2AWriteGlobalState>();
	const uint8_t end_of_stream[] = {0xFF, 0xFF, 0xFF, 0xFF, 0x00, 0x00, 0x00, 0x00}; // arrow_stream_writer.cpp:78-82
	gstate.writer->WriteData(end_of_stream, sizeof(end_of_stream));
	gstate.writer->Close();
}

static unique_ptr<LocalFunctionData> B2AWriteInitializeLocal(ExecutionContext &context, FunctionData &bind_data_p) {
	auto &bind = bind_data_p.Cast<B2AWriteBindData>();
	return make_uniq<B2AWriteLocalState>(context.client, bind.sql_types);
}

static CopyFunctionExecutionMode B2AWriteExecutionMode(bool preserve_insertion_order, bool supports_batch_index) {
	if (!preserve_insertion_order) {
		return CopyFunctionExecutionMode::PARALLEL_COPY_TO_FILE;
	}
	if (supports_batch_index) {
		return CopyFunctionExecutionMode::BATCH_COPY_TO_FILE;
	}
	return CopyFunctionExecutionMode::REGULAR_COPY_TO_FILE;
}

static idx_t B2AWriteDesiredBatchSize(ClientContext &context, FunctionData &bind_data_p) {
	return bind_data_p.Cast<B2AWriteBindData>().row_group_size;
}

static bool B2AWriteRotateFiles(FunctionData &bind_data_p, const optional_idx &file_size_bytes) {
	auto &bind = bind_data_p.Cast<B2AWriteBindData>();
	return file_size_bytes.IsValid() || bind.row_groups_per_file.IsValid();
}

static bool B2AWriteRotateNextFile(GlobalFunctionData &gstate_p, FunctionData &bind_data_p,
                                   const optional_idx &file_size_bytes) {
	auto &gstate = gstate_p.Cast<B2AWriteGlobalState>();
	auto &bind = bind_data_p.Cast<B2AWriteBindData>();
	if (file_size_bytes.IsValid() && gstate.writer->GetTotalWritten() > file_size_bytes.GetIndex()) {
		return true;
	}
	if (bind.row_groups_per_file.IsValid() && gstate.row_group_count >= bind.row_groups_per_file.GetIndex()) {
		return true;
	}
	return false;
}

struct B2AWriteBatchData : public PreparedBatchData {
	unique_ptr<HostStream> stream;
};

//! called concurrently for large writes: converts, does not touch the file
static unique_ptr<PreparedBatchData> B2AWritePrepareBatch(ClientContext &context, FunctionData &bind_data_p,
                                                          GlobalFunctionData &gstate_p,
                                                          unique_ptr<ColumnDataCollection> collection) {
	auto &bind = bind_data_p.Cast<B2AWriteBindData>();
	auto &gstate = gstate_p.Cast<B2AWriteGlobalState>();
	auto batch = make_uniq<B2AWriteBatchData>();
	batch->stream = SerializeCollection(context, *collection, bind.column_names, gstate.settings, {});
	collection->Reset();
	return std::move(batch);
}

static void B2AWriteFlushBatch(ClientContext &context, FunctionData &bind_data, GlobalFunctionData &gstate_p,
                               PreparedBatchData &batch_p) {
	auto &gstate = gstate_p.Cast<B2AWriteGlobalState>();
	auto &batch = batch_p.Cast<B2AWriteBatchData>();
	WriteBatches(gstate, *batch.stream);
}

// ---- to_arrow_ipc(TABLE) -> (ipc BLOB, header BOOLEAN)
//! BLOB rows point straight into the pinned stream: the buffer keeps it alive (the reference's ArrowStringVectorBuffer,
//! src/include/writer/to_arrow_ipc.hpp:16-23)
class B2AStreamVectorBuffer : public VectorBuffer {
public:
	explicit B2AStreamVectorBuffer(shared_ptr<HostStream> stream_p)
	    : VectorBuffer(VectorBufferType::OPAQUE_BUFFER), stream(std::move(stream_p)) {
	}

private:
	shared_ptr<HostStream> stream;
};

struct ToIPCBindData : public TableFunctionData {
	vector<LogicalType> types;
	vector<string> names;
	const idx_t chunk_size = 120 * STANDARD_VECTOR_SIZE; // ToArrowIPCFunction::DEFAULT_CHUNK_SIZE vectors
};

struct ToIPCGlobalState : public GlobalTableFunctionState {
	atomic<bool> sent_schema {false};
	mutex lock;
};

struct ToIPCLocalState : public LocalTableFunctionState {
	unique_ptr<DataChunk> pending;
	idx_t current_count = 0;
	bool checked_schema = false;
	WriteSettings settings;
};

static unique_ptr<FunctionData> ToIPCBind(ClientContext &context, TableFunctionBindInput &input,
                                          vector<LogicalType> &return_types, vector<string> &names) {
	auto result = make_uniq<ToIPCBindData>();
	return_types.emplace_back(LogicalType::BLOB);
	names.emplace_back("ipc");
	return_types.emplace_back(LogicalType::BOOLEAN);
	names.emplace_back("header");
	result->types = input.input_table_types;
	result->names = input.input_table_names;
	for (auto &t : result->types) {
		B2ATypeId(t); // unsupported types fail at bind time, like ToArrowSchema would
	}
	return std::move(result);
}

static unique_ptr<GlobalTableFunctionState> ToIPCInitGlobal(ClientContext &context, TableFunctionInitInput &input) {
	return make_uniq<ToIPCGlobalState>();
}

static unique_ptr<LocalTableFunctionState> ToIPCInitLocal(ExecutionContext &context, TableFunctionInitInput &input,
                                                          GlobalTableFunctionState *global_state) {
	auto state = make_uniq<ToIPCLocalState>();
	state->settings = SettingsOf(context.client);
	return std::move(state);
}

static void EmitBlob(shared_ptr<HostStream> stream, std::pair<const_data_ptr_t, idx_t> bytes, bool header,
                     DataChunk &output) {
	auto &vector = output.data[0];
	StringVector::AddBuffer(vector, make_buffer<B2AStreamVectorBuffer>(std::move(stream)));
	FlatVector::GetData<string_t>(vector)[0] =
	    string_t(const_char_ptr_cast(bytes.first), NumericCast<uint32_t>(bytes.second));
	output.data[1].SetValue(0, Value::BOOLEAN(header));
	output.SetCardinality(1);
}

static shared_ptr<HostStream> FlushPending(const ToIPCBindData &data, ToIPCLocalState &lstate) {
	lstate.pending->SetCardinality(lstate.current_count);
	shared_ptr<HostStream> stream = SerializeChunk(*lstate.pending, data.names, lstate.settings, {});
	lstate.pending.reset();
	lstate.current_count = 0;
	return stream;
}

static OperatorResultType ToIPCFunction(ExecutionContext &context, TableFunctionInput &data_p, DataChunk &input,
                                        DataChunk &output) {
	auto &data = data_p.bind_data->Cast<ToIPCBindData>();
	auto &lstate = data_p.local_state->Cast<ToIPCLocalState>();
	auto &gstate = data_p.global_state->Cast<ToIPCGlobalState>();
	bool sending_schema = false;
	const bool caching_disabled = !PhysicalOperator::OperatorCachingAllowed(context);
	if (!lstate.checked_schema) {
		if (!gstate.sent_schema) {
			lock_guard<mutex> init_lock(gstate.lock);
			if (!gstate.sent_schema) { // this run sends the schema, the other threads only send record batches
				gstate.sent_schema = true;
				sending_schema = true;
			}
		}
		lstate.checked_schema = true;
	}
	if (sending_schema) {
		DataChunk empty;
		empty.Initialize(Allocator::Get(context.client), data.types, 1);
		empty.SetCardinality(0);
		shared_ptr<HostStream> stream = SerializeChunk(empty, data.names, lstate.settings, {});
		auto schema = stream->Schema();
		EmitBlob(stream, schema, true, output);
		return OperatorResultType::HAVE_MORE_OUTPUT; // the same input comes back for the data message
	}
	if (!lstate.pending) {
		lstate.pending = make_uniq<DataChunk>();
		lstate.pending->Initialize(Allocator::Get(context.client), data.types, data.chunk_size + STANDARD_VECTOR_SIZE);
	}
	lstate.pending->SetCardinality(lstate.current_count);
	lstate.pending->Append(input, true);
	lstate.current_count += input.size();
	if (caching_disabled || lstate.current_count >= data.chunk_size) {
		auto stream = FlushPending(data, lstate);
		EmitBlob(stream, stream->Message(0), false, output);
	}
	return OperatorResultType::NEED_MORE_INPUT;
}

static OperatorFinalizeResultType ToIPCFunctionFinal(ExecutionContext &context, TableFunctionInput &data_p,
                                                     DataChunk &output) {
	auto &data = data_p.bind_data->Cast<ToIPCBindData>();
	auto &lstate = data_p.local_state->Cast<ToIPCLocalState>();
	if (lstate.pending && lstate.current_count > 0) {
		auto stream = FlushPending(data, lstate);
		EmitBlob(stream, stream->Message(0), false, output);
	}
	return OperatorFinalizeResultType::FINISHED;
}

// ====================================================================================================================
// registration
// ====================================================================================================================
static void VersionFn(DataChunk &args, ExpressionState &state, Vector &result) {
	// the reference reports the nanoarrow it was built with ("0.7.0-SNAPSHOT", test/sql/nanoarrow.test:18); the decoder
	// here is libb2arrow
	result.SetValue(0, StringVector::AddString(result, b2a_version()));
	result.SetVectorType(VectorType::CONSTANT_VECTOR);
}

void LoadInternal(DatabaseInstance &db) {
	ExtensionUtil::RegisterFunction(db, ScalarFunction("nanoarrow_version", {}, LogicalType::VARCHAR, VersionFn));

	auto read_arrow = ReadArrowFunction();
	ExtensionUtil::RegisterFunction(db, read_arrow);
	read_arrow.arguments = {LogicalType::LIST(LogicalType::VARCHAR)}; // ['a.arrow', 'b.arrow']
	ExtensionUtil::RegisterFunction(db, read_arrow);
	DBConfig::GetConfig(db).replacement_scans.emplace_back(ReadArrowReplacement);

	CopyFunction copy("arrows");
	copy.copy_to_bind = B2AWriteBind;
	copy.copy_to_initialize_global = B2AWriteInitializeGlobal;
	copy.copy_to_initialize_local = B2AWriteInitializeLocal;
	copy.copy_to_sink = B2AWriteSink;
	copy.copy_to_combine = B2AWriteCombine;
	copy.copy_to_finalize = B2AWriteFinalize;
	copy.execution_mode = B2AWriteExecutionMode;
	copy.copy_from_bind = MultiFileFunction<B2AMultiFileInfo>::MultiFileBindCopy;
	copy.copy_from_function = ReadArrowFunction();
	copy.prepare_batch = B2AWritePrepareBatch;
	copy.flush_batch = B2AWriteFlushBatch;
	copy.desired_batch_size = B2AWriteDesiredBatchSize;
	copy.rotate_files = B2AWriteRotateFiles;
	copy.rotate_next_file = B2AWriteRotateNextFile;
	copy.extension = "arrows";
	ExtensionUtil::RegisterFunction(db, copy);
	copy.name = "arrow";
	copy.extension = "arrow";
	ExtensionUtil::RegisterFunction(db, copy);

	child_list_t<LogicalType> buffer_struct {{"ptr", LogicalType::POINTER}, {"size", LogicalType::UBIGINT}};
	TableFunction scan_ipc("scan_arrow_ipc", {LogicalType::LIST(LogicalType::STRUCT(buffer_struct))}, ScanIPCFunction,
	                       ScanIPCBind, ScanIPCInitGlobal);
	scan_ipc.projection_pushdown = true;
	scan_ipc.filter_pushdown = false;
	scan_ipc.filter_prune = false;
	ExtensionUtil::RegisterFunction(db, scan_ipc);

	TableFunction to_ipc("to_arrow_ipc", {LogicalType::TABLE}, nullptr, ToIPCBind, ToIPCInitGlobal, ToIPCInitLocal);
	to_ipc.in_out_function = ToIPCFunction;
	to_ipc.in_out_function_final = ToIPCFunctionFinal;
	ExtensionUtil::RegisterFunction(db, to_ipc);
}

} // namespace ext_nanoarrow
} // namespace duckdb

extern "C" {

DUCKDB_EXTENSION_API void nanoarrow_init(duckdb::DatabaseInstance &db) {
	duckdb::ext_nanoarrow::LoadInternal(db);
}

DUCKDB_EXTENSION_API const char *nanoarrow_version() {
	return duckdb::DuckDB::LibraryVersion();
}
}
