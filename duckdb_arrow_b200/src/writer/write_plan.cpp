#include "write_plan.hpp"

#include <algorithm>
#include <cstring>

namespace b2a {

static const uint64_t kCopyTileBytes = 256 * 1024;

static uint64_t Align8(uint64_t v) {
	return (v + 7) & ~uint64_t(7);
}

int WritePlan::AddVec(const b2a_vector &v, WriteField &wf, int depth) {
	if (depth > 32) {
		throw Error(B2A_EINVAL, "writer: type nesting too deep");
	}
	WVec w;
	w.type_id = v.type_id;
	w.param = v.param;
	w.data = reinterpret_cast<uint64_t>(v.data);
	w.validity = reinterpret_cast<uint64_t>(v.validity);
	w.rows = v.rows;
	wf.name = v.name ? v.name : "";
	wf.type_id = v.type_id;
	wf.decimal_width = v.decimal_width;
	wf.decimal_scale = v.decimal_scale;
	wf.param = v.param;
	wf.timezone = v.timezone ? v.timezone : "";
	auto fixed = [&](int width) {
		w.op = B2A_ENC_COPY;
		w.src_width = w.dst_width = width;
		w.n_buffers = 2;
	};
	wf.lossless = lossless;
	switch (v.type_id) {
	case B2A_T_SQLNULL:
		w.n_buffers = 0;
		break;
	case B2A_T_BOOLEAN:
		if (lossless && depth == 0) {
			// arrow.bool8: ArrowBool8::DuckToArrow copies the byte of every row, then ArrowScalarData<int8_t>.  ROOT columns
			// only — ArrowAppender::Append runs duckdb_to_arrow on root_data[i] (arrow_appender.cpp:40-48) while a nested
			// BOOLEAN keeps ArrowBoolData (:147-149).  The reference still labels the nested field arrow.bool8 / "c" over
			// those bit-packed bytes; here the buffers are the same and the nested field says what they are: "b".
			fixed(1);
			break;
		}
		wf.lossless = false;
		w.op = B2A_ENC_BOOL;
		w.src_width = 1;
		w.n_buffers = 2;
		break;
	case B2A_T_TINYINT:
	case B2A_T_UTINYINT:
		fixed(1);
		break;
	case B2A_T_SMALLINT:
	case B2A_T_USMALLINT:
		fixed(2);
		break;
	case B2A_T_INTEGER:
	case B2A_T_UINTEGER:
	case B2A_T_FLOAT:
	case B2A_T_DATE:
		fixed(4);
		break;
	case B2A_T_BIGINT:
	case B2A_T_UBIGINT:
	case B2A_T_DOUBLE:
	case B2A_T_TIME:
	case B2A_T_TIMESTAMP:
	case B2A_T_TIMESTAMP_SEC:
	case B2A_T_TIMESTAMP_MS:
	case B2A_T_TIMESTAMP_NS:
	case B2A_T_TIMESTAMP_TZ:
		fixed(8);
		break;
	case B2A_T_TIME_TZ: // ArrowScalarData<int64_t, dtime_tz_t, ArrowTimeTzConverter>, arrow_appender.cpp:164-171
		if (lossless) { // ArrowScalarData<int64_t>: the dtime_tz_t bits as they are (w:8)
			fixed(8);
			break;
		}
		w.op = B2A_ENC_TIMETZ;
		w.src_width = w.dst_width = 8;
		w.n_buffers = 2;
		break;
	case B2A_T_UUID: // ArrowVarcharData<hugeint_t, ArrowUUIDConverter, int32_t>, arrow_appender.cpp:180-191
		if (lossless) { // ArrowScalarData<hugeint_t, hugeint_t, ArrowUUIDBlobConverter>: 16 big-endian bytes, NULLs skipped
			w.op = B2A_ENC_UUID_BLOB;
			w.src_width = w.dst_width = 16;
			w.n_buffers = 2;
			break;
		}
		if (string_view) {
			// the reference declares "vu" for a UUID column (arrow_converter.cpp:176-177) but appends offsets + characters
			// (arrow_appender.cpp:180-191): schema and buffers disagree there, nothing to be compatible with
			throw Error(B2A_ENOTIMPL, "writer: UUID columns under produce_arrow_string_view are not supported");
		}
		w.op = large_offsets ? B2A_ENC_UUID64 : B2A_ENC_UUID32;
		w.src_width = 16;
		w.dst_width = large_offsets ? 8 : 4;
		wf.large_offsets = large_offsets;
		w.n_buffers = 3;
		break;
	case B2A_T_UNION: // ArrowUnionData, arrow_appender.cpp:276-278: one buffer (type ids), members as children
		if (v.n_children < 1 || v.n_children > 256) {
			throw Error(B2A_EINVAL, "writer: UNION needs 1 to 256 members");
		}
		w.is_union = true;
		w.src_width = w.dst_width = 1;
		w.n_buffers = 1;
		break;
	case B2A_T_HUGEINT:
		fixed(16);
		break;
	case B2A_T_ENUM: { // ArrowEnumData<int8_t / int16_t / int32_t>, arrow_appender.cpp:253-268: the codes are copied as they are
		if (v.n_children != 1 || !v.children || v.children[0].type_id != B2A_T_VARCHAR) {
			throw Error(B2A_EINVAL, "writer: ENUM '" + wf.name + "' needs one VARCHAR child (its dictionary values)");
		}
		const uint64_t n_values = v.children[0].rows;
		if (n_values > 0xFFFFFFFFull) {
			throw Error(B2A_EINVAL, "writer: ENUM dictionary too large");
		}
		const int width = n_values <= 0xFF ? 1 : n_values <= 0xFFFF ? 2 : 4; // EnumTypeInfo::DictType
		fixed(width);
		wf.index_bits = width * 8;
		wf.dictionary_id = static_cast<int64_t>(dict_vecs.size());
		dict_vecs.push_back(-1); // the id is taken in field pre-order, before any ENUM nested deeper could take it
		break;
	}
	case B2A_T_UHUGEINT: // no switch case in SetArrowFormat: always the registered extension (arrow_converter.cpp:387-393)
		fixed(16);
		break;
	case B2A_T_DECIMAL:
		if (v.decimal_width <= 4) { // ArrowScalarData<int32_t, int16_t>, arrow_appender.cpp:218-220
			w.op = B2A_ENC_WIDEN16;
			w.src_width = 2;
			w.dst_width = 4;
			w.n_buffers = 2;
		} else {
			fixed(v.decimal_width <= 9 ? 4 : v.decimal_width <= 18 ? 8 : 16);
		}
		break;
	case B2A_T_INTERVAL:
		w.op = B2A_ENC_INTERVAL;
		w.src_width = w.dst_width = 16;
		w.n_buffers = 2;
		break;
	case B2A_T_VARCHAR:
	case B2A_T_JSON: // VARCHAR with the JSON alias: same appender; arrow.json in the schema under arrow_lossless_conversion
		if (string_view) { // ArrowVarcharToStringViewData, arrow_appender.cpp:234-235: validity, views, one data buffer
			w.op = B2A_ENC_STRINGVIEW;
			w.src_width = w.dst_width = 16;
			w.n_buffers = 3;
			wf.string_view = true;
			break;
		}
		/* fall through */
	case B2A_T_BLOB:
	case B2A_T_VARINT: // always arrow.opaque{DuckDB, varint} (arrow_converter.cpp:387-393), appended like a BLOB
	case B2A_T_BIT:    // arrow_appender.cpp:239-247: the internal bit string is appended like a BLOB
		w.op = large_offsets ? B2A_ENC_STRING64 : B2A_ENC_STRING32;
		w.src_width = 16;
		w.dst_width = large_offsets ? 8 : 4;
		wf.large_offsets = large_offsets;
		w.n_buffers = 3;
		break;
	case B2A_T_LIST:
	case B2A_T_MAP: { // no large map in Arrow: ArrowMapData<int32_t> whatever the offset size, arrow_appender.cpp:296-299
		const bool large = large_offsets && v.type_id == B2A_T_LIST;
		const bool view = list_view && v.type_id == B2A_T_LIST; // arrow_appender.cpp:280-299
		w.op = view ? (large ? B2A_ENC_LISTVIEW64 : B2A_ENC_LISTVIEW32) : (large ? B2A_ENC_LIST64 : B2A_ENC_LIST32);
		w.src_width = 16;
		w.dst_width = large ? 8 : 4;
		wf.large_offsets = large;
		wf.list_view = view;
		w.n_buffers = view ? 3 : 2; // ArrowListViewData::Finalize: validity, offsets, sizes
		break;
	}
	case B2A_T_STRUCT:
	case B2A_T_ARRAY:
		w.n_buffers = 1;
		break;
	default:
		throw Error(B2A_ENOTIMPL, "Unsupported type in DuckDB -> Arrow Conversion: type id " +
		                              std::to_string(v.type_id));
	}
	if ((w.op != B2A_ENC_NONE || w.is_union) && v.rows > 0 && !v.data) {
		throw Error(B2A_EINVAL, "writer: vector '" + wf.name + "' has no data pointer");
	}
	const int idx = static_cast<int>(vecs.size());
	vecs.push_back(w);
	if (v.n_children < 0 || (v.n_children > 0 && !v.children)) {
		throw Error(B2A_EINVAL, "writer: bad children of '" + wf.name + "'");
	}
	if (v.type_id == B2A_T_ENUM) {
		// the dictionary: a utf8 array with int32 offsets whatever the settings (EnumAppendVector, enum_data.hpp:34-65,
		// child format "u", arrow_converter.cpp:384); not a child of the field
		const bool sv = string_view, lo = large_offsets, ll = lossless;
		string_view = large_offsets = lossless = false;
		WriteField values;
		const int dv = AddVec(v.children[0], values, depth + 1);
		string_view = sv;
		large_offsets = lo;
		lossless = ll;
		dict_vecs[static_cast<size_t>(wf.dictionary_id)] = dv;
		return idx;
	}
	wf.children.resize(static_cast<size_t>(v.n_children));
	for (int c = 0; c < v.n_children; c++) {
		if (v.type_id == B2A_T_UNION) {
			// the appender boxes every row into a Value and SetValue()s it into fresh member vectors; what that leaves
			// in nested members (NULLs pushed down recursively) is not restated here: flat members only
			switch (v.children[c].type_id) {
			case B2A_T_LIST:
			case B2A_T_MAP:
			case B2A_T_STRUCT:
			case B2A_T_ARRAY:
			case B2A_T_UNION:
			case B2A_T_SQLNULL:
				throw Error(B2A_ENOTIMPL, "writer: UNION member '" +
				                              std::string(v.children[c].name ? v.children[c].name : "") +
				                              "' has a nested or NULL type (flat members only)");
			default:
				break;
			}
			if (v.children[c].rows != v.rows) {
				throw Error(B2A_EINVAL, "writer: UNION member shorter than the union");
			}
		}
		int ci = AddVec(v.children[c], wf.children[static_cast<size_t>(c)], depth + 1);
		vecs[static_cast<size_t>(idx)].children.push_back(ci);
		if (v.type_id == B2A_T_UNION) {
			vecs[static_cast<size_t>(ci)].union_vec = idx;
			vecs[static_cast<size_t>(ci)].union_member = c;
		}
	}
	if (v.type_id == B2A_T_MAP && v.n_children == 1) {
		vecs[static_cast<size_t>(vecs[static_cast<size_t>(idx)].children[0])].no_validity_buffer = true;
	}
	return idx;
}

void WritePlan::Build(const b2a_vector *columns, uint32_t n_columns, uint64_t rows, const b2a_write_options *opt) {
	if (n_columns == 0 || !columns) {
		throw Error(B2A_EINVAL, "writer: no columns");
	}
	row_group_size = (opt && opt->row_group_size) ? opt->row_group_size : 122880;
	large_offsets = opt && (opt->flags & B2A_WRITE_LARGE_OFFSETS);
	list_view = opt && (opt->flags & B2A_WRITE_LIST_VIEW);
	string_view = opt && (opt->flags & B2A_WRITE_STRING_VIEW);
	lossless = opt && (opt->flags & B2A_WRITE_LOSSLESS);
	if (opt) {
		for (uint32_t i = 0; i < opt->n_kv_metadata; i++) {
			kv_metadata.emplace_back(opt->kv_keys[i], opt->kv_values[i]);
		}
	}
	total_rows = rows;
	fields.resize(n_columns);
	for (uint32_t c = 0; c < n_columns; c++) {
		if (columns[c].rows != rows) {
			throw Error(B2A_EINVAL, "writer: column " + std::to_string(c) + " has " + std::to_string(columns[c].rows) +
			                            " rows, expected " + std::to_string(rows));
		}
		top_vecs.push_back(AddVec(columns[c], fields[c], 0));
		ResolveArrowFormat(fields[c]);
	}
	schema_message = EncodeSchemaMessage(fields, kv_metadata);
	for (size_t d = 0; d < dict_vecs.size(); d++) {
		WBatch b;
		b.dictionary_id = static_cast<int64_t>(d);
		b.rows = vecs[static_cast<size_t>(dict_vecs[d])].rows;
		b.top.push_back(AddNode(dict_vecs[d], b.rows, 0, 0));
		dict_batches.push_back(b);
	}
	for (uint64_t r0 = 0; r0 < rows; r0 += row_group_size) {
		WBatch b;
		b.row0 = r0;
		b.rows = std::min(row_group_size, rows - r0);
		batches.push_back(b);
	}
	for (auto &b : batches) {
		for (int tv : top_vecs) {
			int n = AddNode(tv, b.rows, b.row0, 0);
			b.top.push_back(n);
		}
	}
}

int WritePlan::AddNode(int vec, uint64_t rows, uint64_t row0, uint64_t sel) {
	const WVec v = vecs[static_cast<size_t>(vec)];
	if (!sel && row0 + rows > v.rows) {
		throw Error(B2A_EINVAL, "writer: vector shorter than the rows it must provide");
	}
	WNode n;
	n.vec = vec;
	n.rows = rows;
	n.row0 = row0;
	n.sel = sel;
	std::memset(n.buf, 0, sizeof(n.buf));
	const int idx = static_cast<int>(nodes.size());
	nodes.push_back(n);
	if (v.type_id == B2A_T_STRUCT || v.type_id == B2A_T_UNION) {
		for (int c : v.children) {
			int ci = AddNode(c, rows, row0, sel);
			nodes[static_cast<size_t>(idx)].children.push_back(ci);
		}
	} else if (v.type_id == B2A_T_ARRAY) {
		const uint64_t k = static_cast<uint64_t>(v.param);
		if (!sel) { // fixed_size_list_data.cpp:16-28: child rows [from*size, to*size)
			int ci = AddNode(v.children.at(0), rows * k, row0 * k, 0);
			nodes[static_cast<size_t>(idx)].children.push_back(ci);
		} else {
			nodes[static_cast<size_t>(idx)].needs_expand_array = true;
		}
	}
	return idx;
}

std::vector<int> WritePlan::PendingScanNodes() const {
	std::vector<int> out;
	for (size_t i = 0; i < nodes.size(); i++) {
		if (!nodes[i].measured && IsScanOp(vecs[static_cast<size_t>(nodes[i].vec)].op)) {
			out.push_back(static_cast<int>(i));
		}
	}
	return out;
}

void WritePlan::Preorder(int node, std::vector<int> &out) const {
	out.push_back(node);
	for (int c : nodes[static_cast<size_t>(node)].children) {
		Preorder(c, out);
	}
}

void WritePlan::Layout() {
	size_t n_nodes = 0, n_buffers = 0;
	uint64_t cursor = schema_message.size();
	body_bytes = 0;
	algorithmic_in = algorithmic_out = 0;
	dict_header_bytes = dict_batches.empty() ? 0 : DictionaryBatchMessageSize(1, 3);
	const size_t n_dicts = dict_batches.size();
	for (size_t bi_all = 0; bi_all < n_dicts + batches.size(); bi_all++) {
		const bool is_dict = bi_all < n_dicts;
		const size_t bi = is_dict ? bi_all : bi_all - n_dicts;
		WBatch &b = is_dict ? dict_batches[bi] : batches[bi];
		if (!is_dict && bi == 0) {
			preamble_bytes = cursor;
		}
		b.preorder.clear();
		for (int t : b.top) {
			Preorder(t, b.preorder);
		}
		uint64_t off = 0;
		size_t nb = 0;
		for (int ni : b.preorder) {
			WNode &n = nodes[static_cast<size_t>(ni)];
			const WVec &v = vecs[static_cast<size_t>(n.vec)];
			nb += static_cast<size_t>(v.n_buffers);
			if (v.is_union) { // type ids, one byte per row; no validity buffer
				n.buf[0] = BufferSlice {off, n.rows};
				off += Align8(n.rows);
				algorithmic_in += n.rows + (v.validity ? (n.rows + 7) / 8 : 0);
				continue;
			}
			if (v.n_buffers >= 1) { // validity
				uint64_t len = v.no_validity_buffer ? 0 : (n.rows + 7) / 8;
				n.buf[0] = BufferSlice {off, len};
				off += Align8(len);
				if (v.validity) {
					algorithmic_in += len;
				}
			}
			if (v.n_buffers >= 2) {
				uint64_t len = 0;
				switch (v.op) {
				case B2A_ENC_BOOL:
					len = (n.rows + 7) / 8;
					break;
				case B2A_ENC_STRING32:
				case B2A_ENC_LIST32:
				case B2A_ENC_UUID32:
					len = (n.rows + 1) * 4;
					break;
				case B2A_ENC_STRING64:
				case B2A_ENC_LIST64:
				case B2A_ENC_UUID64:
					len = (n.rows + 1) * 8;
					break;
				default:
					len = n.rows * static_cast<uint64_t>(v.dst_width);
					break;
				}
				n.buf[1] = BufferSlice {off, len};
				off += Align8(len);
				algorithmic_in += n.rows * static_cast<uint64_t>(v.src_width);
			}
			if (v.n_buffers >= 3) {
				// strings: the characters; list views: the sizes buffer
				const uint64_t len = IsListViewOp(v.op) ? n.rows * static_cast<uint64_t>(v.dst_width) : n.total;
				n.buf[2] = BufferSlice {off, len};
				off += Align8(len);
			}
		}
		if (!is_dict && bi == 0) {
			n_nodes = b.preorder.size();
			n_buffers = nb;
			n_variadic = 0;
			for (int ni : b.preorder) {
				n_variadic += vecs[static_cast<size_t>(nodes[static_cast<size_t>(ni)].vec)].op == B2A_ENC_STRINGVIEW;
			}
			batch_header_bytes = RecordBatchMessageSize(n_nodes, n_buffers, n_variadic);
		}
		b.meta_offset = cursor;
		b.body_offset = cursor + (is_dict ? dict_header_bytes : batch_header_bytes);
		b.body_length = off;
		cursor = b.body_offset + off;
		body_bytes += off;
	}
	if (batches.empty()) {
		preamble_bytes = cursor;
	}
	// the vectors' own bytes + what the string appenders fetch from the heap (a non-inlined string_t holds a pointer)
	algorithmic_in += heap_bytes;
	algorithmic_out = body_bytes;
	stream_bytes = cursor + 8; // end-of-stream marker
	laid_out = true;
}

b2a_enc_task WritePlan::BaseTask(const WNode &n, const WVec &v) const {
	b2a_enc_task t;
	std::memset(&t, 0, sizeof(t));
	t.src = v.data;
	t.validity = v.validity;
	t.sel = n.sel;
	t.row0 = n.row0;
	t.rows = n.rows;
	t.param = v.param;
	t.op = static_cast<uint16_t>(v.op);
	t.src_width = static_cast<uint8_t>(v.src_width);
	t.dst_width = static_cast<uint8_t>(v.dst_width);
	t.scratch = n.scratch;
	t.counter = n.counter;
	return t;
}

static void PushTask(std::vector<b2a_enc_task> &tasks, std::vector<uint32_t> &tile_task, b2a_enc_task t,
                     uint64_t n_tiles) {
	if (tile_task.size() + n_tiles > 0xFFFFFFF0ull) {
		throw Error(B2A_EINVAL, "writer: table too large for one launch");
	}
	t.tile0 = static_cast<uint32_t>(tile_task.size());
	tile_task.insert(tile_task.end(), n_tiles, static_cast<uint32_t>(tasks.size()));
	tasks.push_back(t);
}

void WritePlan::ScanTasks(const std::vector<int> &node_ids, std::vector<b2a_enc_task> &tasks,
                          std::vector<uint32_t> &tile_task) const {
	uint64_t n_tiles = 0;
	for (int ni : node_ids) {
		n_tiles += ScanTiles(nodes[static_cast<size_t>(ni)].rows);
	}
	tasks.reserve(tasks.size() + node_ids.size());
	tile_task.reserve(tile_task.size() + n_tiles);
	for (int ni : node_ids) {
		const WNode &n = nodes[static_cast<size_t>(ni)];
		b2a_enc_task t = BaseTask(n, vecs[static_cast<size_t>(n.vec)]);
		t.tile_rows = B2A_ENC_STR_TILE;
		PushTask(tasks, tile_task, t, ScanTiles(n.rows));
	}
}

// One walk for the three users below: emit(task, n_tiles) is called once per encode task, in launch order.
template <class EMIT>
void WritePlan::WalkEncodeTasks(EMIT &&emit) const {
	for (size_t bi = 0; bi < dict_batches.size() + batches.size(); bi++) {
		const WBatch &b = bi < dict_batches.size() ? dict_batches[bi] : batches[bi - dict_batches.size()];
		for (int ni : b.preorder) {
			const WNode &n = nodes[static_cast<size_t>(ni)];
			const WVec &v = vecs[static_cast<size_t>(n.vec)];
			if (v.is_union) {
				if (n.rows > 0) {
					b2a_enc_task t = BaseTask(n, v);
					t.op = B2A_ENC_UNION_TAG;
					t.dst = b.body_offset + n.buf[0].offset;
					t.dst_len = n.buf[0].length;
					t.tile_rows = 1u << 16;
					emit(t, (n.rows + t.tile_rows - 1) / t.tile_rows);
				}
				continue;
			}
			if (v.n_buffers >= 1 && !v.no_validity_buffer && n.rows > 0) {
				b2a_enc_task t = BaseTask(n, v);
				t.op = B2A_ENC_VALIDITY;
				t.dst = b.body_offset + n.buf[0].offset;
				t.dst_len = n.buf[0].length;
				t.counter = n.null_counter;
				t.tile_rows = 1u << 20;
				emit(t, (n.rows + t.tile_rows - 1) / t.tile_rows);
			}
			if (v.n_buffers < 2) {
				continue;
			}
			b2a_enc_task t = BaseTask(n, v);
			t.dst = b.body_offset + n.buf[1].offset;
			t.dst_len = n.buf[1].length;
			if (IsScanOp(v.op)) {
				t.tile_rows = B2A_ENC_STR_TILE;
				if (v.n_buffers >= 3) {
					t.dst2 = b.body_offset + n.buf[2].offset;
				}
				emit(t, ScanTiles(n.rows));
				continue;
			}
			if (n.rows == 0) {
				continue;
			}
			uint64_t tile_rows;
			if (v.op == B2A_ENC_BOOL) {
				tile_rows = 1u << 18;
			} else {
				tile_rows = kCopyTileBytes / static_cast<uint64_t>(v.src_width + v.dst_width);
				tile_rows = std::max<uint64_t>(2048, tile_rows / 2048 * 2048);
			}
			t.tile_rows = static_cast<uint32_t>(tile_rows);
			emit(t, (n.rows + tile_rows - 1) / tile_rows);
		}
	}
}

void WritePlan::EncodeTasks(std::vector<b2a_enc_task> &tasks, std::vector<uint32_t> &tile_task) const {
	WalkEncodeTasks([&](const b2a_enc_task &t, uint64_t n_tiles) { PushTask(tasks, tile_task, t, n_tiles); });
}

void WritePlan::BuildPreorders() {
	for (auto *set : {&dict_batches, &batches}) {
		for (auto &b : *set) {
			b.preorder.clear();
			for (int t : b.top) {
				Preorder(t, b.preorder);
			}
		}
	}
}

void WritePlan::EncodeTaskSkeleton(std::vector<b2a_enc_task> &tasks, std::vector<uint32_t> &tile_task) const {
	uint64_t n_tasks = 0, n_tiles = 0;
	WalkEncodeTasks([&](const b2a_enc_task &, uint64_t k) {
		n_tasks++;
		n_tiles += k;
	});
	tasks.reserve(n_tasks);
	tile_task.reserve(n_tiles);
	EncodeTasks(tasks, tile_task);
}

void WritePlan::PatchEncodeTasks(std::vector<b2a_enc_task> &tasks) const {
	size_t k = 0;
	WalkEncodeTasks([&](const b2a_enc_task &t, uint64_t) {
		if (k >= tasks.size() || tasks[k].op != t.op || tasks[k].rows != t.rows) {
			throw Error(B2A_EINTERNAL, "writer: encode task skeleton does not match the laid-out plan");
		}
		tasks[k].dst = t.dst;
		tasks[k].dst2 = t.dst2;
		tasks[k].dst_len = t.dst_len;
		tasks[k].counter = t.counter;
		k++;
	});
	if (k != tasks.size()) {
		throw Error(B2A_EINTERNAL, "writer: encode task skeleton does not match the laid-out plan");
	}
}

} // namespace b2a
