// See multi_file.hpp.  Host only: nothing here touches buffer contents.
#include "multi_file.hpp"

#include "../capi_internal.hpp"

#include <glob.h>

#include <algorithm>
#include <cctype>
#include <cstring>
#include <unordered_map>

namespace b2a {

static std::string Lower(const std::string &s) {
	std::string r = s;
	std::transform(r.begin(), r.end(), r.begin(), [](unsigned char c) { return std::tolower(c); });
	return r;
}

// ------------------------------------------------------------------------------------------------ types
static const char *ScalarTypeName(int t) {
	switch (t) {
	case B2A_T_SQLNULL:
		return "\"NULL\"";
	case B2A_T_BOOLEAN:
		return "BOOLEAN";
	case B2A_T_TINYINT:
		return "TINYINT";
	case B2A_T_SMALLINT:
		return "SMALLINT";
	case B2A_T_INTEGER:
		return "INTEGER";
	case B2A_T_BIGINT:
		return "BIGINT";
	case B2A_T_UTINYINT:
		return "UTINYINT";
	case B2A_T_USMALLINT:
		return "USMALLINT";
	case B2A_T_UINTEGER:
		return "UINTEGER";
	case B2A_T_UBIGINT:
		return "UBIGINT";
	case B2A_T_HUGEINT:
		return "HUGEINT";
	case B2A_T_UHUGEINT:
		return "UHUGEINT";
	case B2A_T_FLOAT:
		return "FLOAT";
	case B2A_T_DOUBLE:
		return "DOUBLE";
	case B2A_T_VARCHAR:
		return "VARCHAR";
	case B2A_T_BLOB:
		return "BLOB";
	case B2A_T_DATE:
		return "DATE";
	case B2A_T_TIME:
		return "TIME";
	case B2A_T_TIMESTAMP:
		return "TIMESTAMP";
	case B2A_T_TIMESTAMP_SEC:
		return "TIMESTAMP_S";
	case B2A_T_TIMESTAMP_MS:
		return "TIMESTAMP_MS";
	case B2A_T_TIMESTAMP_NS:
		return "TIMESTAMP_NS";
	case B2A_T_TIMESTAMP_TZ:
		return "TIMESTAMP WITH TIME ZONE";
	case B2A_T_INTERVAL:
		return "INTERVAL";
	case B2A_T_UUID:
		return "UUID";
	case B2A_T_TIME_TZ:
		return "TIME WITH TIME ZONE";
	case B2A_T_BIT:
		return "BIT";
	case B2A_T_VARINT:
		return "VARINT";
	case B2A_T_JSON:
		return "JSON";
	default:
		return "INVALID";
	}
}

static MfType Scalar(int type_id) {
	MfType t;
	t.type_id = type_id;
	t.name = ScalarTypeName(type_id);
	return t;
}

static MfType Decimal(int width, int scale) {
	MfType t;
	t.type_id = B2A_T_DECIMAL;
	t.decimal_width = width;
	t.decimal_scale = scale;
	t.name = "DECIMAL(" + std::to_string(width) + "," + std::to_string(scale) + ")";
	return t;
}

// KeywordHelper::WriteOptionallyQuoted, reduced: quote anything that is not a plain lower-case identifier
static std::string FieldName(const std::string &n) {
	bool plain = !n.empty() && !std::isdigit(static_cast<unsigned char>(n[0]));
	for (unsigned char c : n) {
		plain = plain && (std::islower(c) || std::isdigit(c) || c == '_');
	}
	return plain ? n : "\"" + n + "\"";
}

MfType TypeOfField(const Field &f) {
	MfType t;
	t.type_id = f.type_id;
	switch (f.type_id) {
	case B2A_T_DECIMAL:
		t.decimal_width = f.decimal_width;
		t.decimal_scale = f.decimal_scale;
		t.name = "DECIMAL(" + std::to_string(f.decimal_width) + "," + std::to_string(f.decimal_scale) + ")";
		break;
	case B2A_T_LIST:
		t.name = TypeOfField(*f.children[0]).name + "[]";
		break;
	case B2A_T_ARRAY:
		t.name = TypeOfField(*f.children[0]).name + "[" + std::to_string(f.param) + "]";
		break;
	case B2A_T_MAP: {
		const Field &entries = *f.children[0];
		t.name = "MAP(" + TypeOfField(*entries.children[0]).name + ", " + TypeOfField(*entries.children[1]).name + ")";
		break;
	}
	case B2A_T_UNION:
	case B2A_T_STRUCT: {
		t.name = f.type_id == B2A_T_UNION ? "UNION(" : "STRUCT(";
		for (size_t i = 0; i < f.children.size(); i++) {
			t.name += (i ? ", " : "") + FieldName(f.children[i]->name) + " " + TypeOfField(*f.children[i]).name;
		}
		t.name += ")";
		break;
	}
	default:
		t.name = ScalarTypeName(f.type_id);
	}
	return t;
}

static bool IsSignedInt(int t) {
	return t == B2A_T_TINYINT || t == B2A_T_SMALLINT || t == B2A_T_INTEGER || t == B2A_T_BIGINT;
}
static bool IsUnsignedInt(int t) {
	return t == B2A_T_UTINYINT || t == B2A_T_USMALLINT || t == B2A_T_UINTEGER || t == B2A_T_UBIGINT;
}
static bool IsInt(int t) {
	return IsSignedInt(t) || IsUnsignedInt(t);
}
static int IntBytes(int t) {
	switch (t) {
	case B2A_T_TINYINT:
	case B2A_T_UTINYINT:
		return 1;
	case B2A_T_SMALLINT:
	case B2A_T_USMALLINT:
		return 2;
	case B2A_T_INTEGER:
	case B2A_T_UINTEGER:
		return 4;
	default:
		return 8;
	}
}
static int SignedOfBytes(int b) {
	return b == 1 ? B2A_T_TINYINT : b == 2 ? B2A_T_SMALLINT : b == 4 ? B2A_T_INTEGER : B2A_T_BIGINT;
}

// GetLogicalTypeScore (types.cpp:1203-1290)
static int TypeScore(int t) {
	switch (t) {
	case B2A_T_SQLNULL:
		return 0;
	case B2A_T_BOOLEAN:
		return 10;
	case B2A_T_UTINYINT:
		return 11;
	case B2A_T_TINYINT:
		return 12;
	case B2A_T_USMALLINT:
		return 13;
	case B2A_T_SMALLINT:
		return 14;
	case B2A_T_UINTEGER:
		return 15;
	case B2A_T_INTEGER:
		return 16;
	case B2A_T_UBIGINT:
		return 17;
	case B2A_T_BIGINT:
		return 18;
	case B2A_T_UHUGEINT:
		return 19;
	case B2A_T_HUGEINT:
		return 20;
	case B2A_T_DECIMAL:
		return 21;
	case B2A_T_FLOAT:
		return 22;
	case B2A_T_DOUBLE:
		return 23;
	case B2A_T_TIME:
		return 50;
	case B2A_T_DATE:
		return 51;
	case B2A_T_TIMESTAMP_SEC:
		return 52;
	case B2A_T_TIMESTAMP_MS:
		return 53;
	case B2A_T_TIMESTAMP:
	case B2A_T_TIMESTAMP_TZ:
		return 54;
	case B2A_T_TIMESTAMP_NS:
		return 55;
	case B2A_T_INTERVAL:
		return 56;
	case B2A_T_VARCHAR:
		return 77;
	case B2A_T_BLOB:
		return 101;
	case B2A_T_UUID:
		return 102;
	case B2A_T_STRUCT:
		return 125;
	case B2A_T_LIST:
	case B2A_T_ARRAY:
		return 126;
	case B2A_T_MAP:
		return 127;
	default:
		return 1000;
	}
}

MfType ForceMaxType(const MfType &l, const MfType &r) {
	if (l == r) {
		return l;
	}
	const int a = l.type_id, b = r.type_id;
	if (a == B2A_T_SQLNULL) {
		return r;
	}
	if (b == B2A_T_SQLNULL) {
		return l;
	}
	// everything casts implicitly to VARCHAR and VARCHAR to nothing (CastRules::ImplicitCast): VARCHAR wins
	if (a == B2A_T_VARCHAR || b == B2A_T_VARCHAR) {
		return Scalar(B2A_T_VARCHAR);
	}
	if (IsInt(a) && IsInt(b)) {
		if (IsSignedInt(a) == IsSignedInt(b)) {
			return IntBytes(a) >= IntBytes(b) ? l : r;
		}
		const MfType &s = IsSignedInt(a) ? l : r, &u = IsSignedInt(a) ? r : l;
		if (IntBytes(s.type_id) > IntBytes(u.type_id)) {
			return s; // unsigned -> wider signed is an implicit cast
		}
		// CombineNumericTypes (types.cpp:861-934): the next wider signed type
		if (IntBytes(u.type_id) >= 8) {
			return Scalar(B2A_T_HUGEINT);
		}
		return Scalar(SignedOfBytes(IntBytes(u.type_id) * 2));
	}
	const bool fa = a == B2A_T_FLOAT || a == B2A_T_DOUBLE, fb = b == B2A_T_FLOAT || b == B2A_T_DOUBLE;
	if ((fa || IsInt(a)) && (fb || IsInt(b))) {
		if (a == B2A_T_DOUBLE || b == B2A_T_DOUBLE) {
			return Scalar(B2A_T_DOUBLE);
		}
		return Scalar(B2A_T_FLOAT);
	}
	if ((a == B2A_T_DATE && b == B2A_T_TIMESTAMP) || (a == B2A_T_TIMESTAMP && b == B2A_T_DATE)) {
		return Scalar(B2A_T_TIMESTAMP);
	}
	if (a == B2A_T_DECIMAL && b == B2A_T_DECIMAL) {
		// CombineEqualTypes (types.cpp): the digits in front of the point and behind it both have to fit; past 38 digits
		// the scale gives way
		const int extra = std::max(l.decimal_width - l.decimal_scale, r.decimal_width - r.decimal_scale);
		int scale = std::max(l.decimal_scale, r.decimal_scale);
		int width = extra + scale;
		if (width > 38) {
			width = 38;
			scale = width - extra;
		}
		return Decimal(width, scale);
	}
	if (a == B2A_T_DECIMAL || b == B2A_T_DECIMAL) {
		const MfType &d = a == B2A_T_DECIMAL ? l : r;
		const int o = a == B2A_T_DECIMAL ? b : a;
		if (o == B2A_T_FLOAT || o == B2A_T_DOUBLE) {
			return Scalar(o); // DECIMAL casts implicitly to FLOAT / DOUBLE, not the other way (CastRules::ImplicitCast)
		}
		if (IsInt(o) || o == B2A_T_HUGEINT) {
			// DecimalSizeCheck (types.cpp:859-887) over LogicalType::GetDecimalProperties of the integer type
			int other_width = 38;
			switch (o) {
			case B2A_T_TINYINT:
			case B2A_T_UTINYINT:
				other_width = 3;
				break;
			case B2A_T_SMALLINT:
			case B2A_T_USMALLINT:
				other_width = 5;
				break;
			case B2A_T_INTEGER:
			case B2A_T_UINTEGER:
				other_width = 10;
				break;
			case B2A_T_BIGINT:
				other_width = 19;
				break;
			case B2A_T_UBIGINT:
				other_width = 20;
				break;
			default:
				break;
			}
			if (other_width > d.decimal_width - d.decimal_scale) {
				return Decimal(std::min(other_width + d.decimal_scale, 38), d.decimal_scale);
			}
			return d;
		}
	}
	if (a == B2A_T_DECIMAL || b == B2A_T_DECIMAL || TypeScore(a) >= 100 || TypeScore(b) >= 100) {
		throw Error(B2A_ENOTIMPL, "union_by_name: combining the column types " + l.name + " and " + r.name +
		                              " is not implemented");
	}
	return TypeScore(a) < TypeScore(b) ? r : l;
}

bool CastSupported(const MfType &from, const MfType &to) {
	const int a = from.type_id, b = to.type_id;
	if (from == to) {
		return true;
	}
	if (a == B2A_T_SQLNULL) {
		return true; // becomes a constant NULL
	}
	const bool fa = a == B2A_T_FLOAT || a == B2A_T_DOUBLE;
	if (IsInt(a) && (IsInt(b) || b == B2A_T_FLOAT || b == B2A_T_DOUBLE || b == B2A_T_VARCHAR)) {
		return true;
	}
	if (fa && (b == B2A_T_DOUBLE || b == B2A_T_FLOAT || IsInt(b) || b == B2A_T_VARCHAR)) {
		return true;
	}
	if (a == B2A_T_VARCHAR && (b == B2A_T_DOUBLE || b == B2A_T_BIGINT || b == B2A_T_DATE || b == B2A_T_TIMESTAMP ||
	                           b == B2A_T_DECIMAL)) {
		return true;
	}
	if (a == B2A_T_DATE && b == B2A_T_TIMESTAMP) {
		return true;
	}
	if (IsInt(a) && (b == B2A_T_DECIMAL || b == B2A_T_HUGEINT)) {
		return true; // kernels/cast_decimal.cuh; widening to 128 bits
	}
	if (a == B2A_T_DECIMAL && (b == B2A_T_DECIMAL || b == B2A_T_DOUBLE || b == B2A_T_FLOAT || IsInt(b))) {
		return true;
	}
	if (fa && b == B2A_T_DECIMAL) {
		return true;
	}
	if (b == B2A_T_VARCHAR && (a == B2A_T_BOOLEAN || a == B2A_T_DATE || a == B2A_T_TIME || a == B2A_T_TIMESTAMP ||
	                           a == B2A_T_DECIMAL || a == B2A_T_HUGEINT)) {
		return true; // kernels/cast_format.cuh
	}
	return false;
}

// ------------------------------------------------------------------------------------------------ paths
std::map<std::string, std::string> ParseHivePartitions(const std::string &filename) {
	size_t partition_start = 0, equality_sign = 0;
	bool candidate = true;
	std::map<std::string, std::string> result;
	for (size_t c = 0; c < filename.size(); c++) {
		if (filename[c] == '?' || filename[c] == '\n') {
			candidate = false;
		}
		if (filename[c] == '\\' || filename[c] == '/') {
			if (candidate && equality_sign > partition_start) {
				result.insert({filename.substr(partition_start, equality_sign - partition_start),
				               filename.substr(equality_sign + 1, c - equality_sign - 1)});
			}
			partition_start = c + 1;
			candidate = true;
		} else if (filename[c] == '=') {
			if (equality_sign > partition_start) {
				candidate = false;
			}
			equality_sign = c;
		}
	}
	return result;
}

std::vector<std::string> ExpandGlob(const std::string &pattern) {
	std::vector<std::string> out;
	if (pattern.find_first_of("*?[") == std::string::npos) {
		out.push_back(pattern);
		return out;
	}
	glob_t g;
	std::memset(&g, 0, sizeof(g));
	int rc = glob(pattern.c_str(), 0, nullptr, &g);
	if (rc == 0) {
		for (size_t i = 0; i < g.gl_pathc; i++) {
			out.emplace_back(g.gl_pathv[i]);
		}
	}
	globfree(&g);
	std::sort(out.begin(), out.end());
	return out;
}

// MultiFileOptions::AutoDetectHiveTypesInternal (multi_file_reader.cpp:694-736): candidates DATE, TIMESTAMP, BIGINT.
// Detected here: BIGINT and ISO DATE; TIMESTAMP-looking values stay VARCHAR (documented gap).
static int DetectHiveType(const std::string &v) {
	if (v.empty()) {
		return B2A_T_VARCHAR;
	}
	if (v.size() == 10 && v[4] == '-' && v[7] == '-') {
		bool ok = true;
		for (size_t i = 0; i < 10; i++) {
			ok = ok && (i == 4 || i == 7 || std::isdigit(static_cast<unsigned char>(v[i])));
		}
		if (ok) {
			int m = (v[5] - '0') * 10 + (v[6] - '0'), d = (v[8] - '0') * 10 + (v[9] - '0');
			if (m >= 1 && m <= 12 && d >= 1 && d <= 31) {
				return B2A_T_DATE;
			}
		}
	}
	size_t i = (v[0] == '-' || v[0] == '+') ? 1 : 0;
	if (i == v.size() || v.size() - i > 18) {
		return B2A_T_VARCHAR;
	}
	for (; i < v.size(); i++) {
		if (!std::isdigit(static_cast<unsigned char>(v[i]))) {
			return B2A_T_VARCHAR;
		}
	}
	return B2A_T_BIGINT;
}

// ------------------------------------------------------------------------------------------------ MultiFile
MultiFile::~MultiFile() {
	for (auto &f : files) {
		delete f.reader;
	}
}

void MultiFile::Open(const std::vector<std::string> &patterns, const MfOptions &opt) {
	options = opt;
	for (auto &p : patterns) {
		auto expanded = ExpandGlob(p);
		if (expanded.empty()) {
			// MultiFileReader::CreateFileList -> "No files found that match the pattern"
			throw Error(B2A_EIO, "No files found that match the pattern \"" + p + "\"");
		}
		for (auto &e : expanded) {
			MfFile f;
			f.path = e;
			files.push_back(std::move(f));
		}
	}
	if (files.empty()) {
		throw Error(B2A_EIO, "read_arrow needs at least one file to read");
	}
	// ArrowFileScan::ArrowFileScan (arrow_file_scan.cpp:11-26): schema -> names/types, names de-duplicated
	for (auto &f : files) {
		std::unique_ptr<b2a_reader> r(new b2a_reader());
		r->reader.OpenFile(f.path);
		const Schema &s = r->reader.BaseSchema();
		for (auto &fld : s.fields) {
			f.names.push_back(fld->name);
			f.types.push_back(TypeOfField(*fld));
		}
		DeduplicateColumns(f.names);
		if (f.types.empty()) {
			throw Error(B2A_EINVAL, "Provided table/dataframe must have at least one column");
		}
		f.reader = r.release();
	}
	Bind();
}

void MultiFile::Bind() {
	columns.clear();
	if (!options.union_by_name) {
		// BindReader: the schema of the FIRST file is the global schema (multi_file_reader.cpp:554-580)
		for (size_t i = 0; i < files[0].names.size(); i++) {
			MfColumn c;
			c.name = files[0].names[i];
			c.type = files[0].types[i];
			columns.push_back(std::move(c));
		}
	} else {
		// BindUnionReader -> UnionByName::CombineUnionTypes (:653-672): case-insensitive names, first-seen order
		std::unordered_map<std::string, size_t> seen;
		for (auto &f : files) {
			for (size_t i = 0; i < f.names.size(); i++) {
				auto key = Lower(f.names[i]);
				auto it = seen.find(key);
				if (it != seen.end()) {
					columns[it->second].type = ForceMaxType(columns[it->second].type, f.types[i]);
				} else {
					seen[key] = columns.size();
					MfColumn c;
					c.name = f.names[i];
					c.type = f.types[i];
					columns.push_back(std::move(c));
				}
			}
		}
	}
	// BindOptions (:220-286)
	if (options.filename) {
		for (auto &c : columns) {
			if (c.name == options.filename_column) {
				throw Error(B2A_EINVAL, "Option filename adds column \"" + options.filename_column +
				                            "\", but a column with this name is also in the file. Try setting a different "
				                            "name: filename='<filename column name>'");
			}
		}
		MfColumn c;
		c.name = options.filename_column;
		c.type = Scalar(B2A_T_VARCHAR);
		c.kind = MF_COL_FILENAME;
		columns.push_back(std::move(c));
	}
	if (options.hive_partitioning) {
		auto partitions = ParseHivePartitions(files[0].path);
		std::map<std::string, int> hive_types;
		for (auto &f : files) {
			auto fp = ParseHivePartitions(f.path);
			for (auto &part : partitions) {
				if (fp.find(part.first) == fp.end()) {
					throw Error(B2A_EINVAL, "Hive partition mismatch between file \"" + files[0].path + "\" and \"" +
					                            f.path + "\": key \"" + part.first + "\" not found");
				}
			}
			if (fp.size() != partitions.size()) {
				throw Error(B2A_EINVAL,
				            "Hive partition mismatch between file \"" + files[0].path + "\" and \"" + f.path + "\"");
			}
			for (auto &part : fp) {
				int t = options.hive_types_autocast ? DetectHiveType(part.second) : B2A_T_VARCHAR;
				auto it = hive_types.find(part.first);
				if (it == hive_types.end()) {
					hive_types[part.first] = t;
				} else if (it->second != t) {
					it->second = B2A_T_VARCHAR;
				}
			}
		}
		for (auto &part : partitions) {
			MfType t = Scalar(hive_types[part.first]);
			bool found = false;
			for (auto &c : columns) {
				if (Lower(c.name) == Lower(part.first)) {
					// the partition column also exists in the file: the hive value overrides it
					c.type = t;
					c.kind = MF_COL_HIVE;
					found = true;
					break;
				}
			}
			if (!found) {
				MfColumn c;
				c.name = part.first;
				c.type = t;
				c.kind = MF_COL_HIVE;
				columns.push_back(std::move(c));
			}
		}
	}
	projection.resize(columns.size());
	for (size_t i = 0; i < columns.size(); i++) {
		projection[i] = static_cast<int>(i);
	}
	for (size_t i = 0; i < files.size(); i++) {
		MapFile(i);
	}
	ApplyReaderProjection();
}

// MultiFileColumnMapper::CreateColumnMapping restricted to the NameMapper (the only mapping mode the Arrow reader
// uses: it has no field ids), multi_file_column_mapper.cpp:134-192,494-585
void MultiFile::MapFile(size_t file_idx) {
	MfFile &f = files[file_idx];
	const Schema &schema = f.reader->reader.BaseSchema();
	std::unordered_map<std::string, int> name_map; // case_insensitive_map_t: first occurrence wins
	for (size_t i = 0; i < f.names.size(); i++) {
		name_map.insert({Lower(f.names[i]), static_cast<int>(i)});
	}
	f.maps.assign(columns.size(), MfMap());
	auto hive = options.hive_partitioning ? ParseHivePartitions(f.path) : std::map<std::string, std::string>();
	for (size_t c = 0; c < columns.size(); c++) {
		const MfColumn &col = columns[c];
		MfMap &m = f.maps[c];
		if (col.kind == MF_COL_FILENAME) {
			m.kind = MF_MAP_CONSTANT;
			m.constant = f.path;
			continue;
		}
		if (col.kind == MF_COL_HIVE) {
			// HivePartitioning::GetValue (hive_partitioning.cpp:126-150): 'NULL' (any case) is a NULL, else the value
			m.kind = MF_MAP_CONSTANT;
			for (auto &part : hive) {
				if (Lower(part.first) == Lower(col.name)) {
					m.constant = part.second;
				}
			}
			m.constant_is_null = Lower(m.constant) == "null";
			continue;
		}
		auto it = name_map.find(Lower(col.name));
		if (it == name_map.end()) {
			if (!options.union_by_name) {
				// MultiFileColumnMapper::ThrowColumnNotFoundError (:173-192)
				std::string candidates;
				for (auto &n : f.names) {
					candidates += (candidates.empty() ? "" : ", ") + n;
				}
				throw Error(B2A_EINVAL, "Failed to read file \"" + f.path + "\": schema mismatch in glob: column \"" +
				                            col.name + "\" was read from the original file \"" + files[0].path +
				                            "\", but could not be found in file \"" + f.path + "\".\nCandidate names: " +
				                            candidates + "\nIf you are trying to read files with different schemas, try "
				                                         "setting union_by_name=True");
			}
			m.kind = MF_MAP_NULL;
			m.constant_is_null = true;
			continue;
		}
		m.local_top = it->second;
		m.local_flat = schema.fields[static_cast<size_t>(it->second)]->flat_index;
		m.from = f.types[static_cast<size_t>(it->second)];
		if (m.from == col.type) {
			m.kind = MF_MAP_LOCAL;
		} else if (m.from.type_id == B2A_T_SQLNULL) {
			m.kind = MF_MAP_NULL;
			m.constant_is_null = true;
		} else {
			if (m.from.type_id == B2A_T_STRUCT && col.type.type_id == B2A_T_STRUCT) {
				// remap_struct of unnamed / differently named structs (test/sql/arrow_testing.test:28-31)
				throw Error(B2A_EINVAL, "Struct remap can only remap named structs");
			}
			if (!CastSupported(m.from, col.type)) {
				throw Error(B2A_ENOTIMPL, "column \"" + col.name + "\" of file \"" + f.path + "\": the cast from " +
				                              m.from.name + " to " + col.type.name +
				                              " is not implemented on the device");
			}
			if (schema.fields[static_cast<size_t>(it->second)]->dict_id >= 0) {
				throw Error(B2A_ENOTIMPL, "column \"" + col.name + "\" of file \"" + f.path +
				                              "\": casting a dictionary-encoded column is not implemented");
			}
			m.kind = MF_MAP_CAST;
		}
	}
}

void MultiFile::SetProjection(const std::vector<std::string> &names) {
	std::vector<int> proj;
	for (auto &n : names) {
		int found = -1;
		for (size_t c = 0; c < columns.size(); c++) {
			if (Lower(columns[c].name) == Lower(n)) {
				found = static_cast<int>(c);
				break;
			}
		}
		if (found < 0) {
			std::string candidates;
			for (auto &c : columns) {
				candidates += (candidates.empty() ? "" : ", ") + c.name;
			}
			throw Error(B2A_EINVAL, "Referenced column \"" + n + "\" not found in FROM clause!\nCandidate bindings: " +
			                            candidates);
		}
		proj.push_back(found);
	}
	projection = std::move(proj);
	ApplyReaderProjection();
}

// every reader decodes only the local columns the projected global columns map to (IPCStreamReader::SetColumnProjection,
// base_stream_reader.cpp:146-212); a file that contributes constants only still reads its first column for the row counts
void MultiFile::ApplyReaderProjection() {
	for (auto &f : files) {
		std::vector<std::string> local;
		std::vector<bool> used(f.names.size(), false);
		for (int c : projection) {
			const MfMap &m = f.maps[static_cast<size_t>(c)];
			if ((m.kind == MF_MAP_LOCAL || m.kind == MF_MAP_CAST) && !used[static_cast<size_t>(m.local_top)]) {
				used[static_cast<size_t>(m.local_top)] = true;
			}
		}
		for (size_t i = 0; i < f.names.size(); i++) {
			if (used[i]) {
				local.push_back(f.names[i]);
			}
		}
		if (local.empty()) {
			local.push_back(f.names[0]);
		}
		f.reader->reader.SetProjection(local);
	}
}

} // namespace b2a
