// libref_oracle.so — the reference's OWN converters behind a small C surface, for the tests and the CPU baseline.
//
// TEST INFRASTRUCTURE ONLY (see oracle/build_ref.py).  Links oracle/_ref/libduckdb_ref.so, which is the reference's
// DuckDB compiled from /root/reference/duckdb.  Nothing here restates an algorithm: every conversion is a call into
//   ArrowTableFunction::{PopulateArrowTableType, ArrowScanInitGlobal, ArrowScanInitLocalInternal, ArrowScanFunction}
//       (duckdb/src/function/table/arrow.cpp:20-37,126-201 -> ArrowToDuckDB, arrow_conversion.cpp:1421-1484)
//   ArrowConverter::{ToArrowSchema, ToArrowArray}   (duckdb/src/common/arrow/arrow_converter.cpp:19-25,115-432
//       -> ArrowAppender, arrow_appender.cpp:37-131)
// driven the way the extension drives them (src/file_scanner/arrow_file_scan.cpp:30-72 for the scan,
// src/writer/column_data_collection_serializer.cpp:80-115 for the writer), plus DuckDB's SQL front end for dbgen and
// the Q6 query of benchmark/lineitem.py:41.
//
// The Arrow input comes through the C stream interface from pyarrow's IPC reader (nanoarrow, the reference's slicer,
// is not on disk: CMakeLists.txt:7-13).
#include "duckdb.hpp"
#include "duckdb/common/arrow/arrow_converter.hpp"
#include "duckdb/common/arrow/arrow_wrapper.hpp"
#include "duckdb/common/file_system.hpp"
#include "duckdb/common/serializer/buffered_file_reader.hpp"
#include "duckdb/common/types/vector.hpp"
#include "duckdb/function/table/arrow.hpp"
#include "duckdb/function/table_function.hpp"
#include "duckdb/main/client_context.hpp"
#include "duckdb/main/config.hpp"
#include "duckdb/main/database.hpp"
#include "core_functions_extension.hpp"
#include "tpch_extension.hpp"

#include <chrono>
#include <dlfcn.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

using namespace duckdb;

extern "C" {
// same layout as OSlab in oracle/oracle.h (oracle_py.Slab reads both)
typedef struct RSlab {
	uint64_t rows;
	uint64_t cap;
	int32_t width;
	int32_t type_id;
	uint8_t *data;
	uint8_t *valid;
	int32_t n_children;
	int32_t all_null;
	struct RSlab **children;
	uint32_t *sel;
	struct RSlab *dict;
} RSlab;
}

namespace {

struct Session {
	unique_ptr<DuckDB> db;
	unique_ptr<Connection> con;
	unique_ptr<Connection> meta; // ClientProperties / type lookups while `con` holds an open streaming result
	std::string last_result;
};

Session &session() {
	static Session s;
	if (!s.db) {
		DBConfig config;
		config.options.maximum_threads = 1;
		s.db = make_uniq<DuckDB>(nullptr, &config);
		s.db->LoadStaticExtension<CoreFunctionsExtension>();
		s.db->LoadStaticExtension<TpchExtension>();
		s.con = make_uniq<Connection>(*s.db);
		s.meta = make_uniq<Connection>(*s.db);
	}
	return s;
}

int fail(char *err, int errlen, const std::exception &e) {
	ErrorData ed(e);
	if (err && errlen > 0) {
		snprintf(err, (size_t)errlen, "%s: %s", Exception::ExceptionTypeToString(ed.Type()).c_str(),
		         ed.RawMessage().c_str());
	}
	switch (ed.Type()) {
	case ExceptionType::CONVERSION:
		return 1;
	case ExceptionType::NOT_IMPLEMENTED:
		return 2;
	case ExceptionType::INVALID_INPUT:
		return 3;
	default:
		return 4;
	}
}

// ---- a factory over a borrowed ArrowArrayStream (the role of ArrowIPCStreamFactory, src/ipc/stream_factory.cpp:14-30)
struct BorrowedStreamFactory {
	ArrowArrayStream *src;

	static unique_ptr<ArrowArrayStreamWrapper> Produce(uintptr_t factory_ptr, ArrowStreamParameters &) {
		auto &f = *reinterpret_cast<BorrowedStreamFactory *>(factory_ptr);
		auto w = make_uniq<ArrowArrayStreamWrapper>();
		w->arrow_array_stream = *f.src; // the wrapper takes the stream over and releases it
		f.src->release = nullptr;
		return w;
	}
	static void GetSchema(ArrowArrayStream *factory_ptr, ArrowSchema &schema) {
		auto &f = *reinterpret_cast<BorrowedStreamFactory *>(factory_ptr);
		if (f.src->get_schema(f.src, &schema)) {
			throw InvalidInputException("arrow_scan: get_schema failed(): %s", f.src->get_last_error(f.src));
		}
	}
};

// ---- one reference scan, set up like ArrowFileScan::TryInitializeScan (src/file_scanner/arrow_file_scan.cpp:30-67)
struct RefScan {
	BorrowedStreamFactory factory;
	unique_ptr<ArrowScanFunctionData> bind;
	vector<string> names;
	vector<LogicalType> types;
	vector<idx_t> projection_ids;
	unique_ptr<TableFunctionInitInput> init;
	unique_ptr<GlobalTableFunctionState> gstate;
	unique_ptr<LocalTableFunctionState> lstate;
	unique_ptr<TableFunctionInput> input;
	DataChunk chunk;

	RefScan(ClientContext &context, ArrowArrayStream *stream) {
		factory.src = stream;
		bind = make_uniq<ArrowScanFunctionData>(&BorrowedStreamFactory::Produce, reinterpret_cast<uintptr_t>(&factory));
		BorrowedStreamFactory::GetSchema(reinterpret_cast<ArrowArrayStream *>(&factory), bind->schema_root.arrow_schema);
		ArrowTableFunction::PopulateArrowTableType(DBConfig::GetConfig(context), bind->arrow_table, bind->schema_root,
		                                           names, types);
		QueryResult::DeduplicateColumns(names);
		bind->all_types = types;
		if (types.empty()) {
			throw InvalidInputException("Provided table/dataframe must have at least one column");
		}
		vector<column_t> column_ids;
		for (idx_t i = 0; i < types.size(); i++) {
			column_ids.push_back(i);
		}
		init = make_uniq<TableFunctionInitInput>(bind.get(), column_ids, projection_ids, nullptr);
		gstate = ArrowTableFunction::ArrowScanInitGlobal(context, *init);
		lstate = ArrowTableFunction::ArrowScanInitLocalInternal(context, *init, gstate.get());
		input = make_uniq<TableFunctionInput>(bind.get(), lstate.get(), gstate.get());
		chunk.Initialize(Allocator::DefaultAllocator(), types);
	}
	// MultiFileScan's loop body (multi_file_function.hpp:592-600): Reset, then ArrowFileScan::Scan
	bool Next(ClientContext &context) {
		chunk.Reset();
		ArrowTableFunction::ArrowScanFunction(context, *input, chunk);
		return chunk.size() > 0;
	}
};

// ---- slabs --------------------------------------------------------------------------------------------------------
int type_id_of(const LogicalType &t) {
	// numbering of oracle.h / include/b2arrow.h
	switch (t.id()) {
	case LogicalTypeId::SQLNULL: return 1;
	case LogicalTypeId::BOOLEAN: return 2;
	case LogicalTypeId::TINYINT: return 3;
	case LogicalTypeId::SMALLINT: return 4;
	case LogicalTypeId::INTEGER: return 5;
	case LogicalTypeId::BIGINT: return 6;
	case LogicalTypeId::UTINYINT: return 7;
	case LogicalTypeId::USMALLINT: return 8;
	case LogicalTypeId::UINTEGER: return 9;
	case LogicalTypeId::UBIGINT: return 10;
	case LogicalTypeId::HUGEINT: return 11;
	case LogicalTypeId::UHUGEINT: return 12;
	case LogicalTypeId::FLOAT: return 13;
	case LogicalTypeId::DOUBLE: return 14;
	case LogicalTypeId::DECIMAL: return 15;
	case LogicalTypeId::VARCHAR: return t.IsJSONType() ? 35 : 16;
	case LogicalTypeId::BLOB: return 17;
	case LogicalTypeId::DATE: return 18;
	case LogicalTypeId::TIME: return 19;
	case LogicalTypeId::TIMESTAMP: return 20;
	case LogicalTypeId::TIMESTAMP_SEC: return 21;
	case LogicalTypeId::TIMESTAMP_MS: return 22;
	case LogicalTypeId::TIMESTAMP_NS: return 23;
	case LogicalTypeId::TIMESTAMP_TZ: return 24;
	case LogicalTypeId::INTERVAL: return 25;
	case LogicalTypeId::UUID: return 26;
	case LogicalTypeId::LIST: return 27;
	case LogicalTypeId::STRUCT: return 28;
	case LogicalTypeId::MAP: return 29;
	case LogicalTypeId::ARRAY: return 30;
	case LogicalTypeId::UNION: return 31;
	case LogicalTypeId::TIME_TZ: return 32;
	case LogicalTypeId::BIT: return 33;
	case LogicalTypeId::VARINT: return 34;
	default: return 0;
	}
}

bool is_nested(const LogicalType &t) {
	switch (t.id()) {
	case LogicalTypeId::LIST:
	case LogicalTypeId::STRUCT:
	case LogicalTypeId::MAP:
	case LogicalTypeId::ARRAY:
	case LogicalTypeId::UNION:
		return true;
	default:
		return false;
	}
}

int width_of(const LogicalType &t) {
	switch (t.id()) {
	case LogicalTypeId::SQLNULL:
	case LogicalTypeId::STRUCT:
	case LogicalTypeId::ARRAY:
		return 0;
	case LogicalTypeId::UNION:
		return 1; // the UTINYINT tags
	default:
		return (int)GetTypeIdSize(t.InternalType());
	}
}

RSlab *slab_new(const LogicalType &t, bool dictionary = false) {
	auto s = (RSlab *)calloc(1, sizeof(RSlab));
	s->width = dictionary ? 0 : width_of(t);
	s->type_id = type_id_of(t);
	if (!dictionary && is_nested(t)) {
		vector<LogicalType> kids;
		switch (t.id()) {
		case LogicalTypeId::LIST:
		case LogicalTypeId::MAP:
			kids.push_back(ListType::GetChildType(t));
			break;
		case LogicalTypeId::ARRAY:
			kids.push_back(ArrayType::GetChildType(t));
			break;
		case LogicalTypeId::STRUCT:
			for (auto &c : StructType::GetChildTypes(t)) {
				kids.push_back(c.second);
			}
			break;
		default: // UNION: the members (the tag child is the slab's own data)
			for (idx_t i = 0; i < UnionType::GetMemberCount(t); i++) {
				kids.push_back(UnionType::GetMemberType(t, i));
			}
		}
		s->n_children = (int32_t)kids.size();
		s->children = (RSlab **)calloc(kids.size(), sizeof(RSlab *));
		for (idx_t i = 0; i < kids.size(); i++) {
			s->children[i] = slab_new(kids[i]);
		}
	}
	return s;
}

void slab_free(RSlab *s) {
	if (!s) {
		return;
	}
	for (int i = 0; i < s->n_children; i++) {
		slab_free(s->children[i]);
	}
	free(s->children);
	free(s->data);
	free(s->valid);
	free(s->sel);
	slab_free(s->dict);
	free(s);
}

void slab_reserve(RSlab *s, uint64_t extra, bool with_sel) {
	uint64_t need = s->rows + extra;
	if (need <= s->cap && s->valid && (!with_sel || s->sel)) {
		return;
	}
	uint64_t cap = s->cap ? s->cap : 2048;
	while (cap < need) {
		cap *= 2;
	}
	if (s->width) {
		s->data = (uint8_t *)realloc(s->data, cap * (uint64_t)s->width);
	}
	s->valid = (uint8_t *)realloc(s->valid, cap);
	if (with_sel) {
		s->sel = (uint32_t *)realloc(s->sel, cap * 4);
	}
	s->cap = cap;
}

// Bytes of heap-owned strings, kept for the last 512 scans (the Python side reads them after the scan returns).
std::vector<std::vector<char *>> g_kept(1);
char *keep_bytes(const char *p, uint32_t len) {
	char *c = (char *)malloc(len);
	memcpy(c, p, len);
	g_kept.back().push_back(c);
	return c;
}
void keep_next_scan() {
	if (g_kept.size() >= 512) {
		for (auto c : g_kept.front()) {
			free(c);
		}
		g_kept.erase(g_kept.begin());
	}
	g_kept.emplace_back();
}

// Appends rows [0, count) of the vector the reference produced.  FLAT / CONSTANT(NULL) / DICTIONARY vectors are what
// ArrowToDuckDB leaves behind; anything else is flattened first (never seen on this path).
void slab_append(RSlab *s, Vector &v, idx_t count, idx_t dict_len_hint) {
	const auto &t = v.GetType();
	if (v.GetVectorType() == VectorType::DICTIONARY_VECTOR) {
		slab_reserve(s, count, true);
		auto &sel = DictionaryVector::SelVector(v);
		auto &child = DictionaryVector::Child(v);
		for (idx_t i = 0; i < count; i++) {
			s->sel[s->rows + i] = sel.get_index(i);
			s->valid[s->rows + i] = 1;
		}
		if (!s->dict) {
			s->dict = slab_new(t);
			// dictionary values + the NULL slot SetSelectionVector points NULL indices at (arrow_conversion.cpp:1227-1348)
			idx_t n = dict_len_hint;
			if (n == 0) {
				// a dictionary below a list / struct: DuckDB's dictionary Vector carries no length, take what sel can reach
				for (idx_t i = 0; i < count; i++) {
					n = MaxValue<idx_t>(n, sel.get_index(i) + 1);
				}
			}
			auto &mask = FlatVector::Validity(child);
			bool null_slot = mask.GetData() != nullptr && mask.Capacity() > dict_len_hint;
			slab_append(s->dict, child, n, 0);
			if (null_slot) {
				slab_reserve(s->dict, 1, false);
				if (s->dict->width) {
					memset(s->dict->data + s->dict->rows * (uint64_t)s->dict->width, 0, (size_t)s->dict->width);
				}
				s->dict->valid[s->dict->rows] = (uint8_t)mask.RowIsValid(dict_len_hint);
				s->dict->rows++;
			}
		}
		s->rows += count;
		return;
	}
	if (v.GetVectorType() == VectorType::CONSTANT_VECTOR) {
		slab_reserve(s, count, false);
		bool is_null = ConstantVector::IsNull(v);
		if (is_null) {
			for (idx_t i = 0; i < count; i++) {
				s->valid[s->rows + i] = 0;
			}
			if (s->width) {
				memset(s->data + s->rows * (uint64_t)s->width, 0, count * (uint64_t)s->width);
			}
			s->all_null = 1;
			// children of a constant NULL nested vector: nothing was converted
			if (t.id() == LogicalTypeId::STRUCT || t.id() == LogicalTypeId::UNION) {
				v.Flatten(count);
			} else {
				s->rows += count;
				return;
			}
		} else {
			v.Flatten(count);
		}
	} else if (v.GetVectorType() != VectorType::FLAT_VECTOR) {
		v.Flatten(count);
	}
	slab_reserve(s, count, false);
	auto &mask = FlatVector::Validity(v);
	for (idx_t i = 0; i < count; i++) {
		s->valid[s->rows + i] = (uint8_t)mask.RowIsValid(i);
	}
	switch (t.id()) {
	case LogicalTypeId::STRUCT: {
		auto &kids = StructVector::GetEntries(v);
		for (idx_t c = 0; c < kids.size(); c++) {
			slab_append(s->children[c], *kids[c], count, 0);
		}
		break;
	}
	case LogicalTypeId::UNION: {
		auto &kids = StructVector::GetEntries(v); // {tag, members...}
		Vector &tags = *kids[0];
		tags.Flatten(count);
		memcpy(s->data + s->rows, FlatVector::GetData(tags), count);
		for (idx_t c = 1; c < kids.size(); c++) {
			slab_append(s->children[c - 1], *kids[c], count, 0);
		}
		break;
	}
	case LogicalTypeId::LIST:
	case LogicalTypeId::MAP: {
		memcpy(s->data + s->rows * 16, FlatVector::GetData(v), count * 16);
		idx_t n = ListVector::GetListSize(v);
		if (n) {
			slab_append(s->children[0], ListVector::GetEntry(v), n, 0);
		}
		break;
	}
	case LogicalTypeId::ARRAY:
		slab_append(s->children[0], ArrayVector::GetEntry(v), count * ArrayType::GetSize(t), 0);
		break;
	default:
		if (s->width) {
			auto dst = s->data + s->rows * (uint64_t)s->width;
			memcpy(dst, FlatVector::GetData(v), count * (uint64_t)s->width);
			// strings the reference copied into the vector's own string heap (ArrowToDuckDBBlob -> AddStringOrBlob,
			// arrow_conversion.cpp:342-390) die with the DataChunk: keep their bytes alive for the caller
			if (t.InternalType() == PhysicalType::VARCHAR && v.GetAuxiliary() &&
			    v.GetAuxiliary()->GetBufferType() == VectorBufferType::STRING_BUFFER) {
				for (idx_t i = 0; i < count; i++) {
					if (!mask.RowIsValid(i)) {
						continue;
					}
					uint32_t len;
					memcpy(&len, dst + i * 16, 4);
					if (len > 12) {
						const char *ptr;
						memcpy(&ptr, dst + i * 16 + 8, 8);
						char *keep = keep_bytes(ptr, len);
						memcpy(dst + i * 16 + 8, &keep, 8);
					}
				}
			}
		}
	}
	s->rows += count;
}

} // namespace

extern "C" {

// 0 = ok; 1 conversion, 2 not implemented, 3 invalid input, 4 other (message in err)
int ref_open(char *err, int errlen) {
	try {
		session();
		return 0;
	} catch (std::exception &e) {
		return fail(err, errlen, e);
	}
}

// Reference scan of a stream whose batches are struct arrays with ONE child (the column under test): every 2 048-row
// DataChunk the reference produces is appended to the slab, like oracle_scan_column does for the restatement.
// `dictionary_length` = length of the column's dictionary (0 if none): sizes the dumped dictionary vector.
int ref_scan_column(ArrowArrayStream *stream, uint64_t dictionary_length, RSlab **out, char *err, int errlen) {
	*out = nullptr;
	RSlab *slab = nullptr;
	keep_next_scan();
	try {
		auto &s = session();
		auto &context = *s.con->context;
		s.con->BeginTransaction();
		try {
			RefScan scan(context, stream);
			slab = slab_new(scan.types[0], false);
			slab_reserve(slab, 0, false);
			if (scan.lstate) {
				while (scan.Next(context)) {
					auto &v = scan.chunk.data[0];
					if (v.GetVectorType() == VectorType::DICTIONARY_VECTOR && slab->width) {
						slab->width = 0; // dictionary column: rows live in sel + dict
					}
					slab_append(slab, v, scan.chunk.size(), dictionary_length);
				}
			}
		} catch (...) {
			s.con->Rollback();
			throw;
		}
		s.con->Commit();
		*out = slab;
		return 0;
	} catch (std::exception &e) {
		slab_free(slab);
		return fail(err, errlen, e);
	}
}

void ref_slab_free(RSlab *s) {
	slab_free(s);
}

// Timed CPU baseline: the reference scan of a whole stream, all columns, nothing consuming the vectors (what
// ArrowFileScan::Scan delivers to the executor).  Returns rows; *checksum folds one word of every vector so that the
// conversions cannot be elided.
int64_t ref_scan_stream(ArrowArrayStream *stream, uint64_t *checksum, char *err, int errlen) {
	try {
		auto &s = session();
		Connection con(*s.db); // own connection: callable from several threads at once (one stream each)
		auto &context = *con.context;
		con.BeginTransaction();
		int64_t rows = 0;
		uint64_t sum = 0;
		try {
			RefScan scan(context, stream);
			if (scan.lstate) {
				while (scan.Next(context)) {
					rows += (int64_t)scan.chunk.size();
					for (auto &v : scan.chunk.data) {
						if (v.GetVectorType() == VectorType::FLAT_VECTOR && GetTypeIdSize(v.GetType().InternalType()) >= 4) {
							auto p = FlatVector::GetData(v);
							if (p) {
								uint32_t w;
								memcpy(&w, p + (scan.chunk.size() - 1) * GetTypeIdSize(v.GetType().InternalType()), 4);
								sum += w;
							}
						}
					}
				}
			}
		} catch (...) {
			con.Rollback();
			throw;
		}
		con.Commit();
		if (checksum) {
			*checksum = sum;
		}
		return rows;
	} catch (std::exception &e) {
		fail(err, errlen, e);
		return -1;
	}
}

// a1 of the reference's read path: IPCFileStreamReader reads every message through DuckDB's BufferedFileReader
// (src/ipc/stream_reader/ipc_file_stream_reader.cpp:47-138 -> duckdb/src/common/serializer/buffered_file_reader.cpp:25-49,
// a 4 KiB bounce buffer).  ref_file_read does exactly that ReadData call for one message at `location`.
void *ref_file_open(const char *path, char *err, int errlen) {
	try {
		auto &s = session();
		auto &fs = FileSystem::GetFileSystem(*s.db->instance);
		return new BufferedFileReader(fs, path);
	} catch (std::exception &e) {
		fail(err, errlen, e);
		return nullptr;
	}
}

int ref_file_read(void *handle, uint64_t location, uint64_t size, void *dst, char *err, int errlen) {
	try {
		auto r = (BufferedFileReader *)handle;
		if (r->CurrentOffset() != location) {
			r->Seek(location);
		}
		r->ReadData((data_ptr_t)dst, size);
		return 0;
	} catch (std::exception &e) {
		return fail(err, errlen, e);
	}
}

void ref_file_close(void *handle) {
	delete (BufferedFileReader *)handle;
}

// The writer side: scans the whole stream with the reference's reader into ONE DataChunk (what
// ColumnDataCollectionSerializer::Serialize(ColumnDataCollection&) builds, column_data_collection_serializer.cpp:97-115)
// and hands it to ArrowConverter::ToArrowArray / ToArrowSchema under the given ClientProperties.
//   flags: 1 arrow_large_buffer_size, 2 arrow_output_list_view, 4 produce_arrow_string_view, 8 arrow_lossless_conversion
// The extension types' PopulateSchema callbacks read the settings from the context, not from the ClientProperties
// handed to ToArrowSchema (arrow_type_extension.cpp:306-316): keep the two in step.
static void ApplyFlags(Connection &con, uint32_t flags) {
	const char *names[4] = {"arrow_large_buffer_size", "arrow_output_list_view", "produce_arrow_string_view",
	                        "arrow_lossless_conversion"};
	for (int i = 0; i < 4; i++) {
		auto r = con.Query(string("SET ") + names[i] + " = " + ((flags >> i) & 1 ? "true" : "false"));
		if (r->HasError()) {
			r->GetErrorObject().Throw();
		}
	}
}

int ref_append(ArrowArrayStream *stream, uint32_t flags, ArrowArray *out_array, ArrowSchema *out_schema, char *err,
               int errlen) {
	try {
		auto &s = session();
		auto &context = *s.con->context;
		ApplyFlags(*s.con, flags);
		s.con->BeginTransaction();
		try {
			RefScan scan(context, stream);
			ClientProperties options = context.GetClientProperties();
			options.arrow_offset_size = (flags & 1) ? ArrowOffsetSize::LARGE : ArrowOffsetSize::REGULAR;
			options.arrow_use_list_view = (flags & 2) != 0;
			options.produce_arrow_string_view = (flags & 4) != 0;
			options.arrow_lossless_conversion = (flags & 8) != 0;
			vector<unique_ptr<DataChunk>> chunks;
			idx_t total = 0;
			if (scan.lstate) {
				while (scan.Next(context)) {
					auto c = make_uniq<DataChunk>();
					c->Initialize(Allocator::DefaultAllocator(), scan.types, scan.chunk.size());
					scan.chunk.Copy(*c);
					total += c->size();
					chunks.push_back(std::move(c));
				}
			}
			DataChunk all;
			all.Initialize(Allocator::DefaultAllocator(), scan.types, MaxValue<idx_t>(total, 1));
			for (auto &c : chunks) {
				all.Append(*c, true);
			}
			all.SetCardinality(total);
			ArrowConverter::ToArrowSchema(out_schema, scan.types, scan.names, options);
			auto ext = ArrowTypeExtensionData::GetExtensionTypes(context, scan.types);
			ArrowConverter::ToArrowArray(all, out_array, options, ext);
		} catch (...) {
			s.con->Rollback();
			throw;
		}
		s.con->Commit();
		return 0;
	} catch (std::exception &e) {
		return fail(err, errlen, e);
	}
}

// Registers the stream as view `name` (arrow_scan_dumb: no projection pushdown into the producer, which cannot
// re-project a borrowed stream; callers project the stream themselves, which is what the reference's projection
// pushdown amounts to).  The stream is consumed by the FIRST query that scans the view.
int ref_register_stream(const char *name, ArrowArrayStream *stream, char *err, int errlen) {
	try {
		auto &s = session();
		auto factory = new BorrowedStreamFactory(); // lives as long as the session (a few bytes per call)
		factory->src = stream;
		vector<Value> params;
		params.push_back(Value::POINTER((uintptr_t)factory));
		params.push_back(Value::POINTER((uintptr_t)&BorrowedStreamFactory::Produce));
		params.push_back(Value::POINTER((uintptr_t)&BorrowedStreamFactory::GetSchema));
		s.con->TableFunction("arrow_scan_dumb", params)->CreateView(name, true, true);
		return 0;
	} catch (std::exception &e) {
		return fail(err, errlen, e);
	}
}

// Runs SQL; the result (text, one row per line, columns separated by '|') stays valid until the next call.
int ref_query(const char *sql, const char **result, char *err, int errlen) {
	try {
		auto &s = session();
		auto r = s.con->Query(sql);
		if (r->HasError()) {
			r->ThrowError();
		}
		s.last_result.clear();
		for (idx_t row = 0; row < r->RowCount(); row++) {
			for (idx_t col = 0; col < r->ColumnCount(); col++) {
				if (col) {
					s.last_result += '|';
				}
				s.last_result += r->GetValue(col, row).ToString();
			}
			s.last_result += '\n';
		}
		if (result) {
			*result = s.last_result.c_str();
		}
		return 0;
	} catch (std::exception &e) {
		return fail(err, errlen, e);
	}
}

int ref_set_threads(int n, char *err, int errlen) {
	std::string q = "SET threads = " + std::to_string(n);
	return ref_query(q.c_str(), nullptr, err, errlen);
}

// SELECT ... -> Arrow record batches of `batch_rows` rows through the reference's appender, the way COPY ... (FORMAT
// ARROWS) cuts them (row_group_size, src/writer/write_arrow_stream.cpp:33,141-159).  ref_export_begin runs the query,
// ref_export_next fills one batch (returns 1) or signals the end (0).
struct Export {
	unique_ptr<QueryResult> result;
	vector<LogicalType> types;
	vector<string> names;
	unique_ptr<DataChunk> pending; // rows fetched but not yet emitted
	idx_t pending_offset = 0;
	bool done = false;
	ClientProperties options;
	unordered_map<idx_t, const shared_ptr<ArrowTypeExtensionData>> ext;
	Export() : options("UTC", ArrowOffsetSize::REGULAR, false, false, false, nullptr) {
	}
};

void *ref_export_begin(const char *sql, uint32_t flags, ArrowSchema *out_schema, char *err, int errlen) {
	try {
		auto &s = session();
		auto e = new Export();
		ApplyFlags(*s.meta, flags);
		e->options = s.meta->context->GetClientProperties();
		e->result = s.con->SendQuery(sql);
		if (e->result->HasError()) {
			auto ex = e->result->GetErrorObject();
			delete e;
			ex.Throw();
		}
		e->types = e->result->types;
		e->names = e->result->names;
		e->options.arrow_offset_size = (flags & 1) ? ArrowOffsetSize::LARGE : ArrowOffsetSize::REGULAR;
		e->options.arrow_use_list_view = (flags & 2) != 0;
		e->options.produce_arrow_string_view = (flags & 4) != 0;
		e->options.arrow_lossless_conversion = (flags & 8) != 0;
		e->ext = ArrowTypeExtensionData::GetExtensionTypes(*s.meta->context, e->types);
		ArrowConverter::ToArrowSchema(out_schema, e->types, e->names, e->options);
		return e;
	} catch (std::exception &ex) {
		fail(err, errlen, ex);
		return nullptr;
	}
}

int ref_export_next(void *handle, uint64_t batch_rows, ArrowArray *out_array, char *err, int errlen) {
	try {
		auto e = (Export *)handle;
		DataChunk all;
		all.Initialize(Allocator::DefaultAllocator(), e->types, batch_rows);
		idx_t have = 0;
		while (have < batch_rows && !e->done) {
			if (!e->pending || e->pending_offset >= e->pending->size()) {
				e->pending = e->result->Fetch();
				e->pending_offset = 0;
				if (!e->pending || e->pending->size() == 0) {
					e->pending.reset();
					e->done = true; // a closed stream result must not be fetched again
					break;
				}
			}
			idx_t take = MinValue<idx_t>(batch_rows - have, e->pending->size() - e->pending_offset);
			if (e->pending_offset == 0 && take == e->pending->size()) {
				all.Append(*e->pending, false);
			} else {
				DataChunk part;
				part.Initialize(Allocator::DefaultAllocator(), e->types);
				e->pending->Copy(part, e->pending_offset);
				part.SetCardinality(take);
				all.Append(part, false);
			}
			e->pending_offset += take;
			have += take;
		}
		if (have == 0) {
			return 0;
		}
		ArrowConverter::ToArrowArray(all, out_array, e->options, e->ext);
		return 1;
	} catch (std::exception &ex) {
		fail(err, errlen, ex);
		return -1;
	}
}

void ref_export_end(void *handle) {
	delete (Export *)handle;
}

// Loads a DuckDB extension library into the session's database the way DuckDB's loader does once it has verified the
// file (duckdb/src/main/extension/extension_load.cpp:529-548): dlopen, dlsym("<name>_init"), call it with the
// DatabaseInstance.  Used by the tests to run the reference's SQL statements against integration/duckdb's
// libnanoarrow_b2a.so — the drop-in surface over libb2arrow.so — inside the reference's own DuckDB.
int ref_load_extension(const char *path, const char *name, char *err, int errlen) {
	try {
		auto &s = session();
		void *h = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
		if (!h) {
			throw IOException("dlopen(%s): %s", path, dlerror());
		}
		std::string sym = std::string(name) + "_init";
		typedef void (*init_t)(DatabaseInstance &);
		auto init = (init_t)dlsym(h, sym.c_str());
		if (!init) {
			throw IOException("%s has no symbol %s", path, sym);
		}
		init(*s.db->instance);
		return 0;
	} catch (std::exception &e) {
		return fail(err, errlen, e);
	}
}

// scan_arrow_ipc takes LIST(STRUCT(ptr POINTER, size UBIGINT)): POINTER values cannot be written in SQL, clients build
// them through the API (duckdb/tools/pythonpkg/src/python_replacement_scan.cpp:24-55).  Creates view `name` over
// scan_arrow_ipc(buffers) on the session's connection.
int ref_scan_arrow_ipc_view(const char *name, const uint64_t *ptrs, const uint64_t *sizes, uint32_t n, char *err,
                            int errlen) {
	try {
		auto &s = session();
		vector<Value> entries;
		for (uint32_t i = 0; i < n; i++) {
			child_list_t<Value> fields;
			fields.emplace_back("ptr", Value::POINTER(ptrs[i]));
			fields.emplace_back("size", Value::UBIGINT(sizes[i]));
			entries.push_back(Value::STRUCT(std::move(fields)));
		}
		child_list_t<LogicalType> st {{"ptr", LogicalType::POINTER}, {"size", LogicalType::UBIGINT}};
		vector<Value> params;
		params.push_back(Value::LIST(LogicalType::STRUCT(st), std::move(entries)));
		s.con->TableFunction("scan_arrow_ipc", params)->CreateView(name, true, true);
		return 0;
	} catch (std::exception &e) {
		return fail(err, errlen, e);
	}
}

const char *ref_source_id() {
	return DuckDB::SourceID();
}

} // extern "C"
