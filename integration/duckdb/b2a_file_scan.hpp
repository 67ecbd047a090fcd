//===----------------------------------------------------------------------===//
// B200 variant of the nanoarrow extension's per-file reader.
//
// Drop-in for ext_nanoarrow::ArrowFileScan (reference: src/include/file_scanner/arrow_file_scan.hpp:19-50,
// src/file_scanner/arrow_file_scan.cpp): same base class (duckdb::BaseFileReader), same three overrides, same place in
// MultiFileFunction<ArrowMultiFileInfo> — but the IPC framing / slicing and every per-value conversion go through
// libb2arrow.so (include/b2arrow.h) instead of nanoarrow + ArrowTableFunction::ArrowToDuckDB.
//
// This file is compiled against the DuckDB headers of the reference checkout (tests/test_integration_syntax.py runs
// `g++ -fsyntax-only` over it where /root/reference exists); it cannot be linked here because DuckDB itself is not built
// in this repository.
//===----------------------------------------------------------------------===//
#pragma once

#include "b2arrow.h"

#include "duckdb/common/multi_file/base_file_reader.hpp"
#include "duckdb/common/multi_file/multi_file_data.hpp"
#include "duckdb/common/types/data_chunk.hpp"

namespace duckdb {
namespace ext_nanoarrow {

//! Owner of the decoded vectors of one file (pinned host memory in production); attached to every vector that points
//! into it, the same keep-alive the reference uses for Arrow buffers (ArrowAuxiliaryData, arrow_conversion.cpp:862-864)
struct B2ASlab {
	explicit B2ASlab(idx_t bytes);
	~B2ASlab();
	data_ptr_t data;
	idx_t size;
};

struct B2ASlabAuxData : public VectorAuxiliaryData {
	explicit B2ASlabAuxData(shared_ptr<B2ASlab> slab_p)
	    : VectorAuxiliaryData(VectorAuxiliaryDataType::ARROW_AUXILIARY), slab(std::move(slab_p)) {
	}
	shared_ptr<B2ASlab> slab;
};

class B2AFileScan : public BaseFileReader {
public:
	B2AFileScan(ClientContext &context, const OpenFileInfo &file, int device);
	~B2AFileScan() override;

	string GetReaderType() const override {
		return "ARROW";
	}
	bool TryInitializeScan(ClientContext &context, GlobalTableFunctionState &gstate,
	                       LocalTableFunctionState &lstate) override;
	void Scan(ClientContext &context, GlobalTableFunctionState &global_state, LocalTableFunctionState &local_state,
	          DataChunk &chunk) override;
	shared_ptr<BaseUnionData> GetUnionData(idx_t file_idx) override;

	const vector<string> &GetNames() const {
		return names;
	}
	const vector<LogicalType> &GetTypes() const {
		return types;
	}

private:
	//! DuckDB type of flat field `flat` (b2a_field_info carries what arrow_duck_schema.cpp:59-297 derives)
	LogicalType TypeOf(uint32_t flat) const;
	//! points `vec` at rows [offset, offset + count) of field `flat` of the current batch
	void PointVector(Vector &vec, uint32_t flat, idx_t offset, idx_t count);
	[[noreturn]] void Throw(int status) const;

	int device;
	b2a_reader *reader = nullptr;
	b2a_scan *scan = nullptr;
	vector<b2a_field_info> fields;
	vector<uint32_t> top_level; // flat index of every top-level column of the file
	vector<string> names;
	vector<LogicalType> types;
	vector<uint32_t> projected_flat; // flat index of the column behind every entry of column_indexes

	unordered_map<uint32_t, unique_ptr<Vector>> dictionaries; // flat field -> decoded dictionary values
	shared_ptr<B2ASlab> slab;      // the decoded vectors
	const_data_ptr_t image = nullptr; // the file image (DirectConversion columns alias it)
	uint64_t n_batches = 0, batch = 0;
	idx_t chunk_offset = 0;
	atomic<bool> claimed {false};
};

} // namespace ext_nanoarrow
} // namespace duckdb
