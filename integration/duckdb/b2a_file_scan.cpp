#include "b2a_file_scan.hpp"

#include "duckdb/common/exception.hpp"
#include "duckdb/common/exception/conversion_exception.hpp"
#include "duckdb/common/multi_file/multi_file_reader.hpp"
#include "duckdb/common/types/vector.hpp"
#include "duckdb/main/query_result.hpp"

#include <cstdlib>

namespace duckdb {
namespace ext_nanoarrow {

// ---------------------------------------------------------------------------------------------- slab
B2ASlab::B2ASlab(idx_t bytes) : size(bytes) {
	// production: cudaHostAlloc (pinned) so that b2a_read_arrow_host moves data at full PCIe rate
	data = reinterpret_cast<data_ptr_t>(std::aligned_alloc(256, (bytes + 255) & ~idx_t(255)));
	if (!data) {
		throw OutOfMemoryException("B2ASlab: cannot allocate %llu bytes", bytes);
	}
}
B2ASlab::~B2ASlab() {
	std::free(data);
}

// ---------------------------------------------------------------------------------------------- errors
void B2AFileScan::Throw(int status) const {
	const string msg = b2a_last_error();
	switch (status) { // b2a_status names the DuckDB exception class the reference would have thrown
	case B2A_EINVAL:
		throw InvalidInputException(msg);
	case B2A_EIO:
		throw IOException(msg);
	case B2A_ENOTIMPL:
		throw NotImplementedException(msg);
	case B2A_ECONVERSION:
		throw ConversionException(msg);
	case B2A_EOOM:
		throw OutOfMemoryException(msg);
	default:
		throw InternalException(msg);
	}
}

// ---------------------------------------------------------------------------------------------- schema
LogicalType B2AFileScan::TypeOf(uint32_t flat) const {
	const b2a_field_info &f = fields[flat];
	switch (f.type_id) {
	case B2A_T_SQLNULL:
		return LogicalType::SQLNULL;
	case B2A_T_BOOLEAN:
		return LogicalType::BOOLEAN;
	case B2A_T_TINYINT:
		return LogicalType::TINYINT;
	case B2A_T_SMALLINT:
		return LogicalType::SMALLINT;
	case B2A_T_INTEGER:
		return LogicalType::INTEGER;
	case B2A_T_BIGINT:
		return LogicalType::BIGINT;
	case B2A_T_UTINYINT:
		return LogicalType::UTINYINT;
	case B2A_T_USMALLINT:
		return LogicalType::USMALLINT;
	case B2A_T_UINTEGER:
		return LogicalType::UINTEGER;
	case B2A_T_UBIGINT:
		return LogicalType::UBIGINT;
	case B2A_T_HUGEINT:
		return LogicalType::HUGEINT;
	case B2A_T_UHUGEINT:
		return LogicalType::UHUGEINT;
	case B2A_T_FLOAT:
		return LogicalType::FLOAT;
	case B2A_T_DOUBLE:
		return LogicalType::DOUBLE;
	case B2A_T_DECIMAL:
		return LogicalType::DECIMAL(NumericCast<uint8_t>(f.decimal_width), NumericCast<uint8_t>(f.decimal_scale));
	case B2A_T_VARCHAR:
		return LogicalType::VARCHAR;
	case B2A_T_BLOB:
		return LogicalType::BLOB;
	case B2A_T_DATE:
		return LogicalType::DATE;
	case B2A_T_TIME:
		return LogicalType::TIME;
	case B2A_T_TIMESTAMP:
		return LogicalType::TIMESTAMP;
	case B2A_T_TIMESTAMP_SEC:
		return LogicalType::TIMESTAMP_S;
	case B2A_T_TIMESTAMP_MS:
		return LogicalType::TIMESTAMP_MS;
	case B2A_T_TIMESTAMP_NS:
		return LogicalType::TIMESTAMP_NS;
	case B2A_T_TIMESTAMP_TZ:
		return LogicalType::TIMESTAMP_TZ;
	case B2A_T_INTERVAL:
		return LogicalType::INTERVAL;
	case B2A_T_UUID:
		return LogicalType::UUID;
	case B2A_T_TIME_TZ:
		return LogicalType::TIME_TZ;
	case B2A_T_BIT:
		return LogicalType::BIT;
	case B2A_T_VARINT:
		return LogicalType::VARINT;
	case B2A_T_JSON:
		return LogicalType::JSON();
	case B2A_T_LIST:
		return LogicalType::LIST(TypeOf(NumericCast<uint32_t>(f.first_child)));
	case B2A_T_ARRAY:
		return LogicalType::ARRAY(TypeOf(NumericCast<uint32_t>(f.first_child)), NumericCast<idx_t>(f.param));
	case B2A_T_MAP: {
		const b2a_field_info &entries = fields[NumericCast<uint32_t>(f.first_child)];
		const auto key = NumericCast<uint32_t>(entries.first_child);
		const auto value = NumericCast<uint32_t>(fields[key].next_sibling);
		return LogicalType::MAP(TypeOf(key), TypeOf(value));
	}
	case B2A_T_STRUCT: {
		child_list_t<LogicalType> children;
		for (int32_t c = f.first_child; c >= 0; c = fields[NumericCast<uint32_t>(c)].next_sibling) {
			children.emplace_back(fields[NumericCast<uint32_t>(c)].name, TypeOf(NumericCast<uint32_t>(c)));
		}
		return LogicalType::STRUCT(std::move(children));
	}
	case B2A_T_UNION: { // arrow_duck_schema.cpp:235-259
		child_list_t<LogicalType> members;
		for (int32_t c = f.first_child; c >= 0; c = fields[NumericCast<uint32_t>(c)].next_sibling) {
			members.emplace_back(fields[NumericCast<uint32_t>(c)].name, TypeOf(NumericCast<uint32_t>(c)));
		}
		return LogicalType::UNION(std::move(members));
	}
	default:
		throw NotImplementedException("Unsupported Internal Arrow Type %s", f.format);
	}
}

B2AFileScan::B2AFileScan(ClientContext &context, const OpenFileInfo &file_p, int device_p)
    : BaseFileReader(file_p), device(device_p) {
	// ArrowFileScan::ArrowFileScan (arrow_file_scan.cpp:11-26): open, read the schema, names + types, de-duplicate
	int rc = b2a_reader_open_file(file.path.c_str(), &reader);
	if (rc) {
		Throw(rc);
	}
	uint32_t n_top = 0, n_flat = 0;
	b2a_reader_num_fields(reader, &n_top, &n_flat);
	fields.resize(n_flat);
	for (uint32_t i = 0; i < n_flat; i++) {
		b2a_reader_field(reader, i, &fields[i]);
	}
	for (uint32_t i = 0; i < n_flat; i++) {
		if (fields[i].parent < 0) {
			top_level.push_back(i);
			names.emplace_back(fields[i].name);
			types.push_back(TypeOf(i));
		}
	}
	QueryResult::DeduplicateColumns(names);
	if (types.empty()) {
		throw InvalidInputException("Provided table/dataframe must have at least one column");
	}
	columns = MultiFileColumnDefinition::ColumnsFromNamesAndTypes(names, types);
	const void *img = nullptr;
	uint64_t img_size = 0;
	b2a_reader_image(reader, &img, &img_size);
	image = const_data_ptr_cast(img);
}

B2AFileScan::~B2AFileScan() {
	if (scan) {
		b2a_scan_destroy(scan);
	}
	if (reader) {
		b2a_reader_close(reader);
	}
}

shared_ptr<BaseUnionData> B2AFileScan::GetUnionData(idx_t file_idx) {
	auto data = make_shared_ptr<BaseUnionData>(file);
	data->names = names;
	data->types = types;
	return data;
}

// ---------------------------------------------------------------------------------------------- scan
bool B2AFileScan::TryInitializeScan(ClientContext &context, GlobalTableFunctionState &gstate,
                                    LocalTableFunctionState &lstate) {
	// one scan per file, like the reference (arrow_file_scan.cpp:35-41); a second thread is turned away
	bool expected = false;
	if (!claimed.compare_exchange_strong(expected, true)) {
		return false;
	}
	// projection pushdown: IPCStreamReader::SetColumnProjection (base_stream_reader.cpp:146-212)
	vector<const char *> proj;
	projected_flat.clear();
	for (auto &index : column_indexes) {
		const auto col = index.GetPrimaryIndex();
		if (col >= top_level.size()) {
			throw InternalException("B2AFileScan: column index %llu out of range", col);
		}
		proj.push_back(names[col].c_str());
		projected_flat.push_back(top_level[col]);
	}
	int rc;
	if (!proj.empty()) {
		rc = b2a_reader_set_projection(reader, proj.data(), NumericCast<uint32_t>(proj.size()));
		if (rc) {
			Throw(rc);
		}
	}
	uint64_t n_rows = 0;
	b2a_reader_num_batches(reader, &n_batches, &n_rows);
	// DirectConversion columns alias the file image (zero copy, as in the reference); everything that needs conversion
	// goes host -> device -> kernel -> host inside b2a_read_arrow_host
	rc = b2a_scan_create_ex(reader, device, 0, n_batches, B2A_SCAN_HOST_ALIAS, &scan);
	if (rc) {
		Throw(rc);
	}
	uint64_t image_lo, image_bytes, out_bytes, algorithmic, n_tasks, n_tiles;
	b2a_scan_sizes(scan, &image_lo, &image_bytes, &out_bytes, &algorithmic, &n_tasks, &n_tiles);
	slab = make_shared_ptr<B2ASlab>(NumericCast<idx_t>(out_bytes));
	rc = b2a_read_arrow_host(scan, slab->data, 0);
	if (rc) {
		Throw(rc);
	}
	batch = 0;
	chunk_offset = 0;
	return true;
}

void B2AFileScan::PointVector(Vector &vec, uint32_t flat, idx_t offset, idx_t count) {
	const b2a_field_info &f = fields[flat];
	b2a_column_out co;
	int rc = b2a_scan_column(scan, batch, flat, &co);
	if (rc) {
		Throw(rc);
	}
	// validity: an absent mask means the reference would not have touched it either (all valid)
	if (co.validity_off != B2A_NONE && f.dict_id < 0) {
		if (offset % 64 != 0) {
			throw InternalException("B2AFileScan: vectors start on multiples of 64 rows");
		}
		FlatVector::Validity(vec).Initialize(
		    reinterpret_cast<validity_t *>(slab->data + co.validity_off) + offset / 64, count);
	}
	if (f.dict_id >= 0) {
		// ColumnArrowToDuckDBDictionary (arrow_conversion.cpp:1360-1419): the dictionary values were decoded once per scan
		// (+1 NULL slot at index dict_rows), the kernel already widened the indices to sel_t with NULL -> dict_rows
		auto entry = dictionaries.find(flat);
		if (entry == dictionaries.end()) {
			b2a_column_out dco;
			uint32_t dict_flat = 0;
			rc = b2a_scan_dictionary(scan, flat, &dco, &dict_flat);
			if (rc) {
				Throw(rc);
			}
			auto base = make_uniq<Vector>(vec.GetType(), NumericCast<idx_t>(dco.rows) + 1);
			FlatVector::SetData(*base, slab->data + dco.data_off);
			if (dco.validity_off != B2A_NONE) {
				FlatVector::Validity(*base).Initialize(reinterpret_cast<validity_t *>(slab->data + dco.validity_off),
				                                       NumericCast<idx_t>(dco.rows) + 1);
			}
			base->GetBuffer()->SetAuxiliaryData(make_uniq<B2ASlabAuxData>(slab));
			entry = dictionaries.emplace(flat, std::move(base)).first;
		}
		SelectionVector sel(reinterpret_cast<sel_t *>(slab->data + co.data_off) + offset);
		vec.Slice(*entry->second, sel, count);
		return;
	}
	switch (f.type_id) {
	case B2A_T_SQLNULL:
		vec.SetVectorType(VectorType::CONSTANT_VECTOR);
		ConstantVector::SetNull(vec, true);
		return;
	case B2A_T_STRUCT: {
		auto &entries = StructVector::GetEntries(vec);
		idx_t i = 0;
		for (int32_t c = f.first_child; c >= 0; c = fields[NumericCast<uint32_t>(c)].next_sibling) {
			PointVector(*entries[i++], NumericCast<uint32_t>(c), offset, count);
		}
		return;
	}
	case B2A_T_UNION: {
		// DuckDB's UNION vector is STRUCT{tag UTINYINT, members...}: the node's own data are the tags and its mask is
		// also the tag child's (Vector::SetValue sets both, vector.cpp:442-517)
		auto &entries = StructVector::GetEntries(vec);
		Vector &tag = *entries[0];
		FlatVector::SetData(tag, slab->data + co.data_off + offset);
		FlatVector::Validity(tag).Initialize(reinterpret_cast<validity_t *>(slab->data + co.validity_off) + offset / 64,
		                                     count);
		tag.GetBuffer()->SetAuxiliaryData(make_uniq<B2ASlabAuxData>(slab));
		idx_t i = 1;
		for (int32_t c = f.first_child; c >= 0; c = fields[NumericCast<uint32_t>(c)].next_sibling) {
			PointVector(*entries[i++], NumericCast<uint32_t>(c), offset, count);
		}
		return;
	}
	case B2A_T_LIST:
	case B2A_T_MAP: {
		// list_entry_t{offset,length} are already rebased per 2 048-row chunk; aux = child start row of each chunk
		FlatVector::SetData(vec, slab->data + co.data_off + offset * sizeof(list_entry_t));
		const auto starts = reinterpret_cast<const uint64_t *>(slab->data + co.aux_off);
		const idx_t k = offset / STANDARD_VECTOR_SIZE;
		idx_t child_lo, child_hi;
		if (f.op == B2A_OP_LISTVIEW32 || f.op == B2A_OP_LISTVIEW64) {
			// list views: {lowest offset, sum of sizes} per chunk (ConvertArrowListViewOffsetsTemplated's child window)
			child_lo = NumericCast<idx_t>(starts[2 * k]);
			child_hi = child_lo + NumericCast<idx_t>(starts[2 * k + 1]);
		} else {
			child_lo = NumericCast<idx_t>(starts[k]);
			child_hi = NumericCast<idx_t>(starts[k + 1]);
		}
		auto &child = ListVector::GetEntry(vec);
		PointVector(child, NumericCast<uint32_t>(f.first_child), child_lo, child_hi - child_lo);
		ListVector::SetListSize(vec, child_hi - child_lo);
		break;
	}
	case B2A_T_ARRAY: {
		auto &child = ArrayVector::GetEntry(vec);
		const auto k = NumericCast<idx_t>(f.param);
		PointVector(child, NumericCast<uint32_t>(f.first_child), offset * k, count * k);
		return;
	}
	default: {
		const idx_t width = NumericCast<idx_t>(f.out_width);
		const_data_ptr_t base = co.data_off != B2A_NONE ? slab->data + co.data_off : image + co.image_off;
		FlatVector::SetData(vec, const_cast<data_ptr_t>(base) + offset * width); // NOLINT: DuckDB's own SetData idiom
		break;
	}
	}
	// keep the slab (and, through the reader, the file image the string_t pointers refer to) alive with the vector
	vec.GetBuffer()->SetAuxiliaryData(make_uniq<B2ASlabAuxData>(slab));
}

void B2AFileScan::Scan(ClientContext &context, GlobalTableFunctionState &global_state,
                       LocalTableFunctionState &local_state, DataChunk &chunk) {
	// ArrowTableFunction::ArrowScanFunction (duckdb/src/function/table/arrow.cpp:172-201): at most 2 048 rows per call,
	// next record batch when the current one is exhausted.  No per-chunk conversion is left: vectors are re-pointed.
	while (batch < n_batches) {
		b2a_batch_info bi;
		b2a_reader_batch_info(reader, batch, &bi);
		if (chunk_offset < bi.rows) {
			const idx_t n = MinValue<idx_t>(STANDARD_VECTOR_SIZE, NumericCast<idx_t>(bi.rows) - chunk_offset);
			for (idx_t c = 0; c < chunk.ColumnCount() && c < projected_flat.size(); c++) {
				PointVector(chunk.data[c], projected_flat[c], chunk_offset, n);
			}
			chunk.SetCardinality(n);
			chunk_offset += n;
			return;
		}
		batch++;
		chunk_offset = 0;
	}
	chunk.SetCardinality(0);
}

} // namespace ext_nanoarrow
} // namespace duckdb
