// Host-side Arrow IPC stream reader: message framing + RecordBatch slicing into per-buffer
// descriptors (K0).  Mirrors IPCStreamReader / IPCFileStreamReader / IPCBufferStreamReader of the
// reference (src/ipc/stream_reader/*.cpp) but never touches buffer contents: bodies stay where they
// are (mmap / caller memory) and are described by (offset, length) pairs for the kernels.
#pragma once
#include "ipc_schema.hpp"

#include <string>
#include <unordered_map>
#include <vector>

namespace b2a {

struct BufferSlice {
	uint64_t offset; // relative to the stream image
	uint64_t length; // Buffer.length (unpadded); 0 = absent
};

struct NodeSlice {
	int64_t length;
	int64_t null_count;
};

struct RecordBatch {
	uint64_t rows = 0;
	uint64_t meta_offset = 0; // offset of the 8-byte prefix in the image
	uint64_t body_offset = 0;
	uint64_t body_length = 0;
	int compression = -1; // -1 none, 0 LZ4_FRAME, 1 ZSTD
	std::vector<NodeSlice> nodes;     // pre-order over the FULL schema (or the dictionary value type)
	std::vector<BufferSlice> buffers; // absolute image offsets
	std::vector<int64_t> variadic_counts;
};

struct DictionaryBatch {
	int64_t id = -1;
	bool is_delta = false;
	RecordBatch data;
};

// QueryResult::DeduplicateColumns (duckdb/src/main/query_result.cpp:71-93)
void DeduplicateColumns(std::vector<std::string> &names);

class Reader {
public:
	Reader() {
	}
	~Reader();
	Reader(const Reader &) = delete;
	Reader &operator=(const Reader &) = delete;

	// FileIPCStreamFactory::InitReader (src/ipc/stream_factory.cpp:57-63)
	void OpenFile(const std::string &path);
	// BufferIPCStreamFactory::InitReader (:45-55): a list of {ptr,size}; one buffer is used in
	// place (zero copy), several are laid end to end (8-byte aligned) in an owned image.
	void OpenBuffers(const b2a_ipc_buffer *bufs, uint32_t n);

	// IPCStreamReader::SetColumnProjection (base_stream_reader.cpp:146-212)
	void SetProjection(const std::vector<std::string> &names);
	bool HasProjection() const {
		return !projected_top.empty();
	}

	const Schema &BaseSchema() const {
		return schema;
	}
	// top-level columns after projection (indices into schema.fields)
	std::vector<int> OutputColumns() const;

	const uint8_t *Image() const {
		return image;
	}
	uint64_t ImageSize() const {
		return image_size;
	}
	const std::vector<RecordBatch> &Batches() const {
		return batches;
	}
	const std::unordered_map<int64_t, DictionaryBatch> &Dictionaries() const {
		return dictionaries;
	}
	uint64_t TotalRows() const {
		return total_rows;
	}
	// first buffer index (into RecordBatch.buffers) and node index of every flat field
	const std::vector<int> &FlatBufferStart() const {
		return flat_buffer_start;
	}
	bool HasVariadic() const {
		return has_variadic;
	}

	// the message index came from the Arrow file footer (Block{offset, metaDataLength, bodyLength}) instead of the walk
	bool UsedFooter() const {
		return used_footer;
	}
	// file descriptor of a file-backed image (-1: caller buffers): staged uploads pread() from it
	int FileDescriptor() const {
		return fd;
	}
	// the same file opened with O_DIRECT (-1: not a file, or the file system refuses it, e.g. tmpfs); opened on first use
	int DirectFileDescriptor();

private:
	void Index();
	bool IndexFromFooter();
	bool used_footer = false;
	bool ReadNextMessage(uint64_t &pos); // false at end of stream
	void ParseRecordBatch(const uint8_t *tbl_buf, size_t tbl_len, size_t table_pos, uint64_t body_offset,
	                      uint64_t body_length, RecordBatch &out, bool is_dictionary);

	const uint8_t *image = nullptr;
	uint64_t image_size = 0;
	// ownership
	void *mmap_base = nullptr;
	uint64_t mmap_len = 0;
	int fd = -1;
	int fd_direct = -2; // -2: not tried yet
	std::string path;
	std::vector<uint8_t> owned;

	Schema schema;
	bool have_schema = false;
	std::vector<RecordBatch> batches;
	std::unordered_map<int64_t, DictionaryBatch> dictionaries;
	uint64_t total_rows = 0;
	std::vector<int> projected_top;
	std::vector<int> flat_buffer_start;
	bool has_variadic = false;
};

} // namespace b2a
