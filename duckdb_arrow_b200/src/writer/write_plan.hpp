// Write planner (host): a table of flat DuckDB vectors x row_group_size -> record batches, their field nodes, the
// encode tasks of every buffer and the layout of the produced IPC stream.  This is the part of
// ColumnDataCollectionSerializer::Serialize / ArrowAppender (src/writer/column_data_collection_serializer.cpp:80-115,
// duckdb/src/common/arrow/arrow_appender.cpp:37-131) that decides WHICH appender runs on WHICH vector and where its
// buffers land in the message body (nanoarrow encoder: buffers in pre-order, each padded to 8 bytes); the appenders
// themselves are the kernels in src/kernels/encode.cu.
#pragma once
#include "../ipc/ipc_writer.hpp"
#include "../kernels/enc_task.h"

#include <cstdint>
#include <string>
#include <vector>

namespace b2a {

// one node of the input vector tree (the whole table, device pointers)
struct WVec {
	int type_id = B2A_T_INVALID;
	int op = B2A_ENC_NONE; // op of the main buffer
	int src_width = 0, dst_width = 0;
	int n_buffers = 0; // IPC buffers of the node: 0 null, 1 struct/array, 2 fixed/bool/list/map, 3 string
	int64_t param = 0;
	uint64_t data = 0, validity = 0; // device addresses
	uint64_t rows = 0;
	bool no_validity_buffer = false; // the entries struct of a MAP: its validity buffer is never materialised
	// UNION (ArrowUnionData, union_data.cpp): the node's single buffer holds the type ids (no validity buffer); a
	// member is appended with the mask "union row valid AND tag == member AND member value valid", which
	// b2a_writer_measure computes once per member vector (LaunchEncUnionMask) and stores in `validity`
	bool is_union = false;
	int union_vec = -1;    // member: index of the UNION vector it belongs to
	int union_member = -1; // member: its tag
	bool union_mask_ready = false;
	std::vector<int> children;       // indices into WritePlan::vecs
};

// one (record batch, field) node
struct WNode {
	int vec = -1;
	uint64_t rows = 0;
	uint64_t row0 = 0;     // identity mapping: first input row
	uint64_t sel = 0;      // device address of the uint32 selection, 0 = identity
	uint64_t total = 0;    // scan ops: character bytes / child count (known after measure)
	uint64_t scratch = 0;  // scan ops: device address of the per-tile sums / bases
	uint64_t counter = 0;  // scan ops: device address of `total`
	uint64_t null_counter = 0; // device address of the null count (after layout)
	int64_t null_count = 0;    // read back by finish
	bool measured = false;
	bool needs_expand_array = false; // ARRAY under a selection: child sel must be built from this node's sel
	std::vector<int> children;       // node indices
	BufferSlice buf[3];              // body-relative offsets + unpadded lengths
};

struct WBatch {
	uint64_t rows = 0;
	uint64_t row0 = 0;
	int64_t dictionary_id = -1; // >= 0: a DictionaryBatch (one utf8 column: the values of that ENUM dictionary)
	std::vector<int> top; // node index per column
	uint64_t meta_offset = 0, body_offset = 0, body_length = 0;
	std::vector<int> preorder; // nodes in RecordBatch.nodes order
};

struct WritePlan {
	std::vector<WriteField> fields; // schema (names, formats)
	std::vector<std::pair<std::string, std::string>> kv_metadata;
	std::vector<WVec> vecs;
	std::vector<int> top_vecs;
	std::vector<WNode> nodes;
	std::vector<WBatch> batches;
	// one per ENUM field, in field pre-order = dictionary id order; laid out between the schema message and the first
	// record batch (a dictionary must precede the first batch that refers to it)
	std::vector<WBatch> dict_batches;
	std::vector<int> dict_vecs; // dictionary id -> index of its VARCHAR vector in `vecs`
	size_t dict_header_bytes = 0;
	uint64_t preamble_bytes = 0; // schema message + dictionary batches
	uint64_t row_group_size = 122880;
	bool large_offsets = false; // B2A_WRITE_LARGE_OFFSETS
	bool list_view = false;     // B2A_WRITE_LIST_VIEW
	bool string_view = false;   // B2A_WRITE_STRING_VIEW
	bool lossless = false;      // B2A_WRITE_LOSSLESS
	size_t n_variadic = 0;      // string-view nodes per batch: RecordBatch.variadicBufferCounts = [1] * n_variadic
	uint64_t total_rows = 0;

	std::vector<uint8_t> schema_message;
	size_t batch_header_bytes = 0;
	uint64_t stream_bytes = 0, body_bytes = 0;
	uint64_t algorithmic_in = 0, algorithmic_out = 0;
	uint64_t heap_bytes = 0; // characters of the strings longer than 12 bytes (summed by the measure kernel)
	bool laid_out = false;

	// builds vecs / fields from the C-ABI description and creates the nodes that need no device-side measurement
	void Build(const b2a_vector *columns, uint32_t n_columns, uint64_t rows, const b2a_write_options *opt);
	// creates node `vec` (+ its struct / identity-array descendants); list children are created after measuring
	int AddNode(int vec, uint64_t rows, uint64_t row0, uint64_t sel);
	// nodes whose op is a scan op and that have not been measured yet
	std::vector<int> PendingScanNodes() const;
	// after every scan node is measured: message offsets and per-buffer offsets
	void Layout();
	// encode tasks for all nodes (dst = absolute stream offsets); needs Layout()
	void EncodeTasks(std::vector<b2a_enc_task> &tasks, std::vector<uint32_t> &tile_task) const;
	// The same walk split in two, so that the host builds the 10^4 tasks / 10^6 tile entries of a large table WHILE the
	// measure kernels run: the skeleton (everything but the stream offsets; needs the final node set and
	// BuildPreorders()) before the totals are known, the offsets (dst / dst2 / dst_len) after Layout().
	void BuildPreorders();
	void EncodeTaskSkeleton(std::vector<b2a_enc_task> &tasks, std::vector<uint32_t> &tile_task) const;
	void PatchEncodeTasks(std::vector<b2a_enc_task> &tasks) const;
	// measure-phase tasks for the given scan nodes
	void ScanTasks(const std::vector<int> &node_ids, std::vector<b2a_enc_task> &tasks,
	               std::vector<uint32_t> &tile_task) const;

	static bool IsScanOp(int op) {
		return op == B2A_ENC_STRING32 || op == B2A_ENC_STRING64 || op == B2A_ENC_LIST32 || op == B2A_ENC_LIST64 ||
		       op == B2A_ENC_UUID32 || op == B2A_ENC_UUID64 || op == B2A_ENC_LISTVIEW32 || op == B2A_ENC_LISTVIEW64 ||
		       op == B2A_ENC_STRINGVIEW;
	}
	static bool IsListViewOp(int op) {
		return op == B2A_ENC_LISTVIEW32 || op == B2A_ENC_LISTVIEW64;
	}
	static bool IsListOp(int op) { // ops whose node has a child appended through a selection
		return op == B2A_ENC_LIST32 || op == B2A_ENC_LIST64 || IsListViewOp(op);
	}
	static uint64_t ScanTiles(uint64_t rows) {
		uint64_t n = (rows + B2A_ENC_STR_TILE - 1) / B2A_ENC_STR_TILE;
		return n ? n : 1; // a zero-row node still writes offsets[0] = 0
	}

private:
	int AddVec(const b2a_vector &v, WriteField &wf, int depth);
	void Preorder(int node, std::vector<int> &out) const;
	b2a_enc_task BaseTask(const WNode &n, const WVec &v) const;
	template <class EMIT>
	void WalkEncodeTasks(EMIT &&emit) const;
};

} // namespace b2a
