// see b2a_scanner.hpp
#include "b2a_scanner.hpp"

#include <chrono>
#include <cstdio>

#include "duckdb/common/exception.hpp"
#include "duckdb/common/exception/conversion_exception.hpp"
#include "duckdb/main/query_result.hpp"

#include <cstdlib>

namespace duckdb {
namespace ext_nanoarrow {

// ---------------------------------------------------------------------------------------------- errors / device
void B2AThrow(int status) {
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
	case B2A_ECUDA: // no device / a CUDA failure: an I/O-class error, not a reason to invalidate the database
		throw IOException(msg);
	default:
		throw InternalException(msg);
	}
}

int B2ADevice() {
	const char *e = std::getenv("B2A_DEVICE");
	return e ? std::atoi(e) : 0;
}

// ---------------------------------------------------------------------------------------------- pinned arena pool
namespace {
struct SlabPool {
	std::mutex mu;
	vector<std::pair<data_ptr_t, idx_t>> free_list;
	idx_t cached_bytes = 0;
	static constexpr idx_t kMaxCached = idx_t(4) << 30;
	~SlabPool() {
		for (auto &e : free_list) {
			b2a_pinned_free(e.first);
		}
	}
};
SlabPool &Pool() {
	static SlabPool pool;
	return pool;
}
} // namespace

shared_ptr<B2ASlab> B2ASlab::Acquire(idx_t bytes) {
	bytes = (bytes + 255) & ~idx_t(255);
	auto &pool = Pool();
	{
		std::lock_guard<std::mutex> guard(pool.mu);
		idx_t best = pool.free_list.size();
		for (idx_t i = 0; i < pool.free_list.size(); i++) { // smallest arena that fits
			if (pool.free_list[i].second >= bytes &&
			    (best == pool.free_list.size() || pool.free_list[i].second < pool.free_list[best].second)) {
				best = i;
			}
		}
		if (best != pool.free_list.size() && pool.free_list[best].second <= 2 * bytes + (idx_t(1) << 20)) {
			auto e = pool.free_list[best];
			pool.free_list.erase(pool.free_list.begin() + NumericCast<int64_t>(best));
			pool.cached_bytes -= e.second;
			return make_shared_ptr<B2ASlab>(e.first, e.second);
		}
	}
	void *p = nullptr;
	int rc = b2a_pinned_alloc(&p, bytes);
	if (rc) {
		B2AThrow(rc);
	}
	return make_shared_ptr<B2ASlab>(data_ptr_cast(p), bytes);
}

B2ASlab::~B2ASlab() {
	auto &pool = Pool();
	std::lock_guard<std::mutex> guard(pool.mu);
	if (pool.cached_bytes + capacity <= SlabPool::kMaxCached) {
		pool.free_list.emplace_back(data, capacity);
		pool.cached_bytes += capacity;
	} else {
		b2a_pinned_free(data);
	}
}

// ---------------------------------------------------------------------------------------------- schema
LogicalType B2AScanner::TypeOf(uint32_t flat) const {
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


B2AScanner::B2AScanner(b2a_reader *reader_p, int device_p)
    : device(device_p), reader(reader_p), bin(make_shared_ptr<B2AScanBin>()) {
	// ArrowFileScan::ArrowFileScan (arrow_file_scan.cpp:11-26) / ScanArrowIPCBind (scan_arrow_ipc.cpp:20-48): read the
	// schema, names + types (ArrowTableFunction::PopulateArrowTableType), de-duplicate the names
	uint32_t n_top = 0, n_flat = 0;
	b2a_reader_num_fields(reader, &n_top, &n_flat);
	fields.resize(n_flat);
	for (uint32_t i = 0; i < n_flat; i++) {
		b2a_reader_field(reader, i, &fields[i]);
	}
	for (uint32_t i = 0; i < n_flat; i++) {
		if (fields[i].parent < 0) {
			top_level.push_back(i);
			string name = fields[i].name;
			if (name.empty()) { // arrow.cpp:31-34
				name = "v" + std::to_string(names.size());
			}
			names.push_back(std::move(name));
			types.push_back(TypeOf(i));
		}
	}
	QueryResult::DeduplicateColumns(names);
	if (types.empty()) {
		throw InvalidInputException("Provided table/dataframe must have at least one column");
	}
	const void *img = nullptr;
	uint64_t img_size = 0;
	b2a_reader_image(reader, &img, &img_size);
	image = const_data_ptr_cast(img);
}

B2AScanBin::~B2AScanBin() {
	for (auto scan : scans) {
		b2a_scan_destroy(scan);
	}
}

void B2AScanBin::Add(b2a_scan *scan) {
	std::lock_guard<std::mutex> guard(mu);
	scans.push_back(scan);
}

B2AScanner::Slice::~Slice() {
	dictionaries.clear();
	if (scan) {
		if (bin) {
			bin->Add(scan);
		} else {
			b2a_scan_destroy(scan);
		}
	}
}

B2AScanner::~B2AScanner() {
	if (!workers.empty()) {
		{
			std::lock_guard<std::mutex> guard(mu);
			stop = true;
		}
		cv.notify_all();
		for (auto &w : workers) {
			w.join();
		}
	}
	out_of_order.clear();
	own.slice.reset();
	ready.clear();
	bin.reset(); // (the scans go now unless a cursor of another thread still holds a slice)
	if (reader) {
		b2a_reader_close(reader);
	}
}

// ---------------------------------------------------------------------------------------------- slices
// B2A_SCANNER_TRACE=1: wall clock of the producer's steps per slice, on stderr
static bool ScannerTrace() {
	static const bool on = std::getenv("B2A_SCANNER_TRACE") != nullptr;
	return on;
}
static double NowMs() {
	return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

unique_ptr<B2AScanner::Slice> B2AScanner::Load(uint64_t lo, uint64_t hi) {
	const double t0 = NowMs();
	double t1 = t0, t2 = t0;
	auto s = make_uniq<Slice>();
	s->bin = bin;
	s->batch_lo = lo;
	s->batch_hi = hi;
	// DirectConversion columns alias the stream image (zero copy, as in the reference); everything that needs
	// conversion goes host -> device -> kernels -> host inside b2a_read_arrow_host
	// (compressed bodies have nothing to alias: every buffer is decompressed on the device, base_stream_reader.cpp:11-50)
	b2a_batch_info first;
	bool compressed = lo < hi && b2a_reader_batch_info(reader, lo, &first) == B2A_OK && first.compressed != 0;
	int rc = b2a_scan_create_ex(reader, device, lo, hi, compressed ? 0 : B2A_SCAN_HOST_ALIAS, &s->scan);
	t1 = NowMs();
	if (rc == B2A_OK) {
		uint64_t image_lo, image_bytes, out_bytes, algorithmic, n_tasks, n_tiles;
		b2a_scan_sizes(s->scan, &image_lo, &image_bytes, &out_bytes, &algorithmic, &n_tasks, &n_tiles);
		uint64_t scratch_offset = 0, scratch_bytes = 0, n_jobs = 0;
		b2a_scan_scratch(s->scan, &scratch_offset, &scratch_bytes, &n_jobs);
		try {
			// vectors first, the host copy of the decompressed character buffers behind them in the same arena
			const idx_t out_padded = (NumericCast<idx_t>(out_bytes) + 255) & ~idx_t(255);
			s->slab = B2ASlab::Acquire(out_padded + NumericCast<idx_t>(scratch_bytes));
			t2 = NowMs();
			// sub-slices of the H2D / decode / D2H pipeline inside this slice (0 = the library's default)
			static const uint32_t pipe_slices = [] {
				const char *e = std::getenv("B2A_PIPE_SLICES");
				return e ? NumericCast<uint32_t>(std::atoll(e)) : 0u;
			}();
			rc = b2a_read_arrow_host_scratch(s->scan, s->slab->data, scratch_bytes ? s->slab->data + out_padded : nullptr,
			                                 pipe_slices);
		} catch (std::exception &e) {
			s->status = B2A_EOOM;
			s->error = e.what();
			return s;
		}
	}
	if (rc != B2A_OK) {
		s->status = rc;
		s->error = b2a_last_error();
	}
	if (ScannerTrace()) {
		std::fprintf(stderr, "[b2a scanner] batches %llu-%llu: scan_create %.2f ms, slab %.2f ms, read_arrow_host %.2f ms\n",
		             static_cast<unsigned long long>(lo), static_cast<unsigned long long>(hi), t1 - t0, t2 - t1,
		             NowMs() - t2);
	}
	return s;
}

B2AScanner::Cursor::Cursor() = default;
B2AScanner::Cursor::~Cursor() = default;
B2AScanner::Cursor::Cursor(Cursor &&) noexcept = default;
B2AScanner::Cursor &B2AScanner::Cursor::operator=(Cursor &&) noexcept = default;

void B2AScanner::WorkerLoop() {
	while (true) {
		idx_t k;
		{
			std::unique_lock<std::mutex> lock(mu);
			// a slice is started while fewer than `depth` are started-but-not-drained (drained = popped from `ready`)
			cv.wait(lock, [&] {
				return stop || produced == n_slices || claimed == n_slices || claimed < handed_out + depth;
			});
			if (stop || produced == n_slices || claimed == n_slices) {
				return;
			}
			k = claimed++;
		}
		shared_ptr<Slice> s = Load(slice_bounds[k], slice_bounds[k + 1]);
		s->next_batch = s->batch_lo;
		{
			std::lock_guard<std::mutex> guard(mu);
			out_of_order[k] = std::move(s);
			while (produced < n_slices) { // move what is now in order
				auto it = out_of_order.find(produced);
				if (it == out_of_order.end()) {
					break;
				}
				const bool failed = it->second->status != B2A_OK;
				ready.push_back(std::move(it->second));
				out_of_order.erase(it);
				produced = failed ? n_slices : produced + 1; // nothing is handed out behind a failed slice
			}
		}
		cv.notify_all();
	}
}

bool B2AScanner::TakeSlice(Cursor &cursor) {
	cursor.slice.reset();
	shared_ptr<Slice> s;
	uint64_t lo = 0, hi = 0;
	bool first = false, drained = false;
	{
		std::unique_lock<std::mutex> lock(mu);
		cv.wait(lock, [&] { return !ready.empty() || produced == n_slices; });
		if (ready.empty()) {
			return false;
		}
		s = ready.front();
		first = s->next_batch == s->batch_lo;
		lo = s->next_batch;
		hi = s->status != B2A_OK ? s->batch_hi : MinValue<uint64_t>(lo + 1, s->batch_hi);
		s->next_batch = hi;
		if (hi >= s->batch_hi) {
			ready.pop_front();
			handed_out++;
			drained = true;
		}
	}
	if (drained) {
		cv.notify_all(); // room for the producer
	}
	if (s->status != B2A_OK) {
		// re-raise on the executor's thread with the class the reference would have used
		const string msg = s->error;
		switch (s->status) {
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
		case B2A_ECUDA:
			throw IOException(msg);
		default:
			throw InternalException(msg);
		}
	}
	if (first) {
		gpu_launches += b2a_scan_gpu_launches(s->scan);
	}
	cursor.batch = lo;
	cursor.batch_end = hi;
	cursor.chunk_offset = 0;
	cursor.slice = std::move(s);
	return true;
}

void B2AScanner::Begin(const vector<idx_t> &columns) {
	if (begun) {
		throw InternalException("B2AScanner::Begin called twice");
	}
	vector<const char *> proj;
	projected_flat.clear();
	for (auto col : columns) {
		if (col >= top_level.size()) {
			throw InternalException("B2AScanner: column index %llu out of range", col);
		}
		proj.push_back(names[col].c_str());
		projected_flat.push_back(top_level[col]);
	}
	if (!proj.empty()) {
		int rc = b2a_reader_set_projection(reader, proj.data(), NumericCast<uint32_t>(proj.size()));
		if (rc) {
			B2AThrow(rc);
		}
	}
	b2a_reader_num_batches(reader, &n_batches, &total_rows);
	// slices of whole record batches, ~512 MiB of message bodies each (measured on the box, tools/bench_sql.py: 128 MiB ...
	// 1 GiB within 15 % of each other once a scan costs 0.1 ms to create; smaller slices start the consumers earlier)
	static const uint64_t target = [] {
		const char *e = std::getenv("B2A_SLICE_MB");
		return (e ? NumericCast<uint64_t>(std::atoll(e)) : 512) << 20;
	}();
	slice_bounds.clear();
	slice_bounds.push_back(0);
	uint64_t acc = 0;
	for (uint64_t b = 0; b < n_batches; b++) {
		b2a_batch_info bi;
		b2a_reader_batch_info(reader, b, &bi);
		acc += bi.body_length;
		if (acc >= target && b + 1 < n_batches) {
			slice_bounds.push_back(b + 1);
			acc = 0;
		}
	}
	if (n_batches > 0) {
		slice_bounds.push_back(n_batches);
	}
	n_slices = slice_bounds.size() > 1 ? slice_bounds.size() - 1 : 0;
	rows_done = 0;
	begun = true;
	if (const char *e = std::getenv("B2A_SLICES_AHEAD")) {
		depth = MaxValue<idx_t>(1, NumericCast<idx_t>(std::atoll(e)));
	}
	idx_t producers = 2;
	if (const char *e = std::getenv("B2A_PRODUCERS")) {
		producers = MaxValue<idx_t>(1, NumericCast<idx_t>(std::atoll(e)));
	}
	for (idx_t i = 0; i < MinValue<idx_t>(producers, n_slices); i++) {
		workers.emplace_back([this] { WorkerLoop(); });
	}
}

// ---------------------------------------------------------------------------------------------- vectors
void B2AScanner::PointVector(Slice &slice, Vector &vec, uint32_t flat, uint64_t batch, idx_t offset, idx_t count) {
	b2a_scan *scan = slice.scan;
	const shared_ptr<B2ASlab> &slab = slice.slab;
	auto &dictionaries = slice.dictionaries;
	const b2a_field_info &f = fields[flat];
	b2a_column_out co;
	int rc = b2a_scan_column(scan, batch, flat, &co);
	if (rc) {
		B2AThrow(rc);
	}
	// validity: an absent mask means the reference would not have touched it either (all valid)
	if (co.validity_off != B2A_NONE && f.dict_id < 0) {
		if (offset % 64 != 0) {
			throw InternalException("B2AScanner: vectors start on multiples of 64 rows");
		}
		FlatVector::Validity(vec).Initialize(
		    reinterpret_cast<validity_t *>(slab->data + co.validity_off) + offset / 64, count);
	}
	if (f.dict_id >= 0) {
		// ColumnArrowToDuckDBDictionary (arrow_conversion.cpp:1360-1419): the dictionary values were decoded once per scan
		// (+1 NULL slot at index dict_rows), the kernel already widened the indices to sel_t with NULL -> dict_rows
		std::lock_guard<std::mutex> dict_guard(slice.dict_mu);
		auto entry = dictionaries.find(flat);
		if (entry == dictionaries.end()) {
			b2a_column_out dco;
			uint32_t dict_flat = 0;
			rc = b2a_scan_dictionary(scan, flat, &dco, &dict_flat);
			if (rc) {
				B2AThrow(rc);
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
			PointVector(slice, *entries[i++], NumericCast<uint32_t>(c), batch, offset, count);
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
			PointVector(slice, *entries[i++], NumericCast<uint32_t>(c), batch, offset, count);
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
		PointVector(slice, child, NumericCast<uint32_t>(f.first_child), batch, child_lo, child_hi - child_lo);
		ListVector::SetListSize(vec, child_hi - child_lo);
		break;
	}
	case B2A_T_ARRAY: {
		auto &child = ArrayVector::GetEntry(vec);
		const auto k = NumericCast<idx_t>(f.param);
		PointVector(slice, child, NumericCast<uint32_t>(f.first_child), batch, offset * k, count * k);
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

void B2AScanner::NextFrom(Cursor &cur, DataChunk &chunk) {
	// ArrowTableFunction::ArrowScanFunction (duckdb/src/function/table/arrow.cpp:172-201): at most 2 048 rows per call,
	// next record batch when the current one is exhausted.  No per-chunk conversion is left: vectors are re-pointed.
	if (cur.slice) {
		while (cur.batch < cur.batch_end) {
			b2a_batch_info bi;
			b2a_reader_batch_info(reader, cur.batch, &bi);
			if (cur.chunk_offset < bi.rows) {
				const idx_t n = MinValue<idx_t>(STANDARD_VECTOR_SIZE, NumericCast<idx_t>(bi.rows) - cur.chunk_offset);
				for (idx_t c = 0; c < chunk.ColumnCount() && c < projected_flat.size(); c++) {
					PointVector(*cur.slice, chunk.data[c], projected_flat[c], cur.batch, cur.chunk_offset, n);
				}
				chunk.SetCardinality(n);
				cur.chunk_offset += n;
				rows_done += n;
				return;
			}
			cur.batch++;
			cur.chunk_offset = 0;
		}
		cur.slice.reset(); // the slab lives on for as long as a vector points into it
	}
	chunk.SetCardinality(0);
}

void B2AScanner::Next(DataChunk &chunk) {
	while (true) {
		if (!own.slice && (n_slices == 0 || !TakeSlice(own))) {
			chunk.SetCardinality(0);
			return;
		}
		NextFrom(own, chunk);
		if (chunk.size() > 0) {
			return;
		}
	}
}

double B2AScanner::Progress() const {
	return total_rows ? 100.0 * static_cast<double>(rows_done.load()) / static_cast<double>(total_rows) : 100.0;
}

} // namespace ext_nanoarrow
} // namespace duckdb
