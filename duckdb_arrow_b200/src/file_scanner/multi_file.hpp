// Multi-file scan binding (host): file list / glob expansion, the global schema (first file, or the union of all
// files by name), the per-file column mapping and the constant columns (filename, hive partitions).
//
// In the reference this is DuckDB's MultiFileReader framework driven by the extension's ArrowMultiFileInfo /
// ArrowFileScan (src/file_scanner/arrow_multi_file_info.cpp:57-72, arrow_file_scan.cpp:11-26,75-80):
//   bind      : MultiFileReader::BindReader / BindUnionReader (duckdb/src/common/multi_file/multi_file_reader.cpp:528-580),
//               UnionByName::CombineUnionTypes (:653-672), BindOptions (:220-286)
//   mapping   : MultiFileColumnMapper (duckdb/src/common/multi_file/multi_file_column_mapper.cpp:134-192,494-585):
//               columns are matched BY NAME (case-insensitive); a type difference becomes a cast expression, a column
//               missing from a file is an error unless union_by_name, where it becomes a constant NULL
//   constants : filename / hive partition values (multi_file_reader.cpp:300-345, hive_partitioning.cpp:95-124)
// The expressions this binding produces are executed per chunk by MultiFileReader::FinalizeChunk
// (multi_file_reader.cpp:435-458); here the casts run on the device (kernels/cast.cu), constants stay constants.
#pragma once
#include "../ipc/ipc_reader.hpp"

#include <map>
#include <memory>
#include <string>
#include <vector>

struct b2a_reader;

namespace b2a {

struct MfType {
	int type_id = B2A_T_INVALID;
	int decimal_width = 0, decimal_scale = 0;
	std::string name; // LogicalType::ToString(): "BIGINT", "DECIMAL(15,2)", "STRUCT(a BIGINT, b VARCHAR)", "BIGINT[]"
	bool operator==(const MfType &o) const {
		return name == o.name;
	}
	bool operator!=(const MfType &o) const {
		return name != o.name;
	}
};

enum MfColumnKind { MF_COL_DATA = 0, MF_COL_FILENAME = 1, MF_COL_HIVE = 2 };
enum MfMapKind { MF_MAP_LOCAL = 0, MF_MAP_CAST = 1, MF_MAP_NULL = 2, MF_MAP_CONSTANT = 3 };

struct MfColumn {
	std::string name;
	MfType type;
	int kind = MF_COL_DATA;
};

struct MfMap {
	int kind = MF_MAP_NULL;
	int local_top = -1;  // top-level column of the file
	int local_flat = -1; // its flat (pre-order) field index
	MfType from;         // local type (MF_MAP_CAST)
	std::string constant;
	bool constant_is_null = false;
};

struct MfOptions {
	bool union_by_name = false;
	bool filename = false;
	bool hive_partitioning = false;
	bool hive_types_autocast = true;
	std::string filename_column = "filename";
};

struct MfFile {
	std::string path;
	b2a_reader *reader = nullptr; // owned (closed by MultiFile)
	std::vector<std::string> names;
	std::vector<MfType> types;
	std::vector<MfMap> maps; // one per global column
};

class MultiFile {
public:
	MultiFile() {
	}
	~MultiFile();
	MultiFile(const MultiFile &) = delete;
	MultiFile &operator=(const MultiFile &) = delete;

	void Open(const std::vector<std::string> &patterns, const MfOptions &opt);
	// restricts the scan to these global columns (in this order): every reader gets the projection of the local
	// columns it has to deliver
	void SetProjection(const std::vector<std::string> &names);

	std::vector<MfFile> files;
	std::vector<MfColumn> columns;     // global schema (bind order)
	std::vector<int> projection;       // indices into columns, in output order
	MfOptions options;

private:
	void Bind();
	void MapFile(size_t file_idx);
	void ApplyReaderProjection();
};

// LogicalType of a field (arrow_duck_schema.cpp rules already applied by MapFieldToDuckDB)
MfType TypeOfField(const Field &f);
// LogicalType::ForceMaxLogicalType (duckdb/src/common/types.cpp:936-1010,1302-1315) for the type pairs this path supports
MfType ForceMaxType(const MfType &l, const MfType &r);
// can the device cast kernel convert `from` into `to`? (kernels/cast.cu)
bool CastSupported(const MfType &from, const MfType &to);
// HivePartitioning::Parse (duckdb/src/common/hive_partitioning.cpp:95-124)
std::map<std::string, std::string> ParseHivePartitions(const std::string &filename);
// expands a path or glob pattern into a sorted file list (FileSystem::GlobFiles semantics for * ? [..])
std::vector<std::string> ExpandGlob(const std::string &pattern);

} // namespace b2a
