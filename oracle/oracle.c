/* See oracle.h.  TEST INFRASTRUCTURE ONLY — never linked into the product.
 * "ref:" comments cite duckdb/src/function/table/arrow_conversion.cpp unless another file is named. */
#include "oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef uint64_t idx_t;

/* ---------------------------------------------------------------- a minimal DuckDB flat vector */
typedef struct OVector {
	int32_t type_id;
	int32_t width;
	idx_t capacity;
	uint8_t *data;
	uint64_t *validity; /* NULL = all valid (validity_mask.hpp:82-84) */
	idx_t validity_cap; /* in rows */
	int n_children;
	struct OVector **children;
	idx_t list_size;
	int is_constant_null;
	int borrowed_data; /* FlatVector::SetData: the vector aliases the Arrow buffer (DirectConversion) */
	/* dictionary vector (Vector::Slice) */
	uint32_t *sel;
	struct OVector *dict;
	idx_t dict_len;
} OVector;

#define O_MAX_DICTS 64
typedef struct OState {
	idx_t chunk_offset; /* ArrowScanLocalState::chunk_offset */
	/* ArrowArrayScanState dictionary cache (arrow_array_scan_state.cpp:29-52), keyed by ArrowArray* */
	const void *dict_key[O_MAX_DICTS];
	struct OVector *dict_vec[O_MAX_DICTS];
	int n_dicts;
	/* ArrowRunEndEncodingState (arrow_array_scan_state.hpp): run_ends + values vectors, decoded once per array */
	const void *ree_key[O_MAX_DICTS];
	struct OVector *ree_runs[O_MAX_DICTS];
	struct OVector *ree_values[O_MAX_DICTS];
	int n_rees;
	int alias_direct; /* 1: DirectConversion re-points the vector like the reference (timed baseline); 0: copies */
	char *err;
	int errlen;
} OState;

#define O_OK 0
#define O_ECONV 1
#define O_ENOTIMPL 2
#define O_EINVAL 3
#define TRY(x)                                                                                                         \
	do {                                                                                                               \
		int _rc = (x);                                                                                                 \
		if (_rc) {                                                                                                     \
			return _rc;                                                                                                \
		}                                                                                                              \
	} while (0)

static int fail(OState *st, int code, const char *msg) {
	if (st->err && st->errlen > 0) {
		snprintf(st->err, (size_t)st->errlen, "%s", msg);
	}
	return code;
}

static OVector *vector_new(const OType *type, idx_t capacity) {
	OVector *v = (OVector *)calloc(1, sizeof(OVector));
	v->type_id = type->id;
	v->width = type->width;
	v->capacity = capacity;
	if (type->width > 0) {
		v->data = (uint8_t *)calloc(capacity ? capacity : 1, (size_t)type->width);
	}
	if (type->id == O_STRUCT || type->id == O_UNION) {
		v->n_children = type->n_children;
		v->children = (OVector **)calloc((size_t)type->n_children, sizeof(OVector *));
		for (int i = 0; i < type->n_children; i++) {
			v->children[i] = vector_new(type->children[i], capacity);
		}
	} else if (type->id == O_LIST || type->id == O_MAP) {
		v->n_children = 1;
		v->children = (OVector **)calloc(1, sizeof(OVector *));
		v->children[0] = NULL; /* created by list_reserve */
	} else if (type->id == O_ARRAY) {
		v->n_children = 1;
		v->children = (OVector **)calloc(1, sizeof(OVector *));
		v->children[0] = vector_new(type->children[0], capacity * (idx_t)type->fixed_size);
	}
	return v;
}

static void vector_free(OVector *v) {
	if (!v) {
		return;
	}
	for (int i = 0; i < v->n_children; i++) {
		vector_free(v->children[i]);
	}
	free(v->children);
	if (!v->borrowed_data) {
		free(v->data);
	}
	free(v->validity);
	free(v->sel);
	/* v->dict is owned by the scan state */
	free(v);
}

/* ValidityMask::EnsureWritable / Resize: fresh masks are all ones (validity_mask.hpp:320-324) */
static void mask_ensure(OVector *v, idx_t rows) {
	idx_t cap = v->capacity > rows ? v->capacity : rows;
	if (cap < O_VECTOR_SIZE) {
		cap = O_VECTOR_SIZE;
	}
	idx_t words = (cap + 63) / 64;
	if (!v->validity) {
		v->validity = (uint64_t *)malloc(words * 8);
		memset(v->validity, 0xFF, words * 8);
		v->validity_cap = words * 64;
	} else if (v->validity_cap < rows) {
		idx_t old_words = v->validity_cap / 64;
		v->validity = (uint64_t *)realloc(v->validity, words * 8);
		memset(v->validity + old_words, 0xFF, (words - old_words) * 8);
		v->validity_cap = words * 64;
	}
}
static int row_is_valid(const OVector *v, idx_t i) {
	return !v->validity || ((v->validity[i >> 6] >> (i & 63)) & 1);
}
static void set_invalid(OVector *v, idx_t i) {
	mask_ensure(v, i + 1);
	v->validity[i >> 6] &= ~((uint64_t)1 << (i & 63));
}
static void set_valid(OVector *v, idx_t i) { /* ValidityMask::Set(i, true): nothing to do without a buffer */
	if (v->validity) {
		v->validity[i >> 6] |= (uint64_t)1 << (i & 63);
	}
}
static int all_valid(const OVector *v) {
	return v->validity == NULL;
}

/* ref :33-42 */
static void ShiftRight(unsigned char *ar, int size, int shift) {
	int carry = 0;
	while (shift--) {
		for (int i = size - 1; i >= 0; --i) {
			int next = (ar[i] & 1) ? 0x80 : 0;
			ar[i] = (unsigned char)(carry | (ar[i] >> 1));
			carry = next;
		}
	}
}

/* ref :45-56 */
static idx_t GetEffectiveOffset(const OArray *array, int64_t parent_offset, const OState *state, int64_t nested_offset) {
	if (nested_offset != -1) {
		return (idx_t)(array->offset + nested_offset);
	}
	return (idx_t)(array->offset + parent_offset) + state->chunk_offset;
}

/* ref :63-108 (STANDARD_VECTOR_SIZE > 64 branch) */
static void GetValidityMask(OVector *mask_owner, const OArray *array, const OState *state, idx_t size,
                            int64_t parent_offset, int64_t nested_offset, int add_null) {
	if (array->null_count != 0 && array->n_buffers > 0 && array->buffers[0]) {
		idx_t bit_offset = GetEffectiveOffset(array, parent_offset, state, nested_offset);
		mask_ensure(mask_owner, size);
		idx_t n_bitmask_bytes = (size + 8 - 1) / 8;
		const uint8_t *src = (const uint8_t *)array->buffers[0];
		if (bit_offset % 8 == 0) {
			memcpy((void *)mask_owner->validity, src + bit_offset / 8, n_bitmask_bytes);
		} else {
			unsigned char *tmp = (unsigned char *)malloc(n_bitmask_bytes + 1);
			memcpy(tmp, src + bit_offset / 8, n_bitmask_bytes + 1);
			ShiftRight(tmp, (int)(n_bitmask_bytes + 1), (int)(bit_offset % 8ull));
			memcpy((void *)mask_owner->validity, tmp, n_bitmask_bytes);
			free(tmp);
		}
	}
	if (add_null) {
		mask_ensure(mask_owner, size + 1);
		set_invalid(mask_owner, size);
	}
}

/* ref :139-165 */
typedef struct {
	idx_t list_size;
	idx_t start_offset;
} ListOffsetData;

static ListOffsetData ConvertArrowListOffsets(OVector *vector, const OArray *array, idx_t size, const OType *type,
                                              idx_t effective_offset) {
	ListOffsetData result = {0, 0};
	if (type->list_view) { /* ConvertArrowListViewOffsetsTemplated :166-203 */
		uint64_t *le = (uint64_t *)vector->data;
		const int wide = type->size_type != O_SIZE_NORMAL;
		const uint32_t *o32 = (const uint32_t *)array->buffers[1] + effective_offset;
		const uint32_t *s32 = (const uint32_t *)array->buffers[2] + effective_offset;
		const uint64_t *o64 = (const uint64_t *)array->buffers[1] + effective_offset;
		const uint64_t *s64 = (const uint64_t *)array->buffers[2] + effective_offset;
		uint64_t lowest_offset = size ? (wide ? o64[0] : o32[0]) : 0;
		for (idx_t i = 0; i < size; i++) {
			uint64_t off = wide ? o64[i] : o32[i], len = wide ? s64[i] : s32[i];
			le[2 * i] = off;
			le[2 * i + 1] = len;
			result.list_size += len;
			if (len != 0 && off < lowest_offset) {
				lowest_offset = off;
			}
		}
		result.start_offset = lowest_offset;
		if (result.start_offset) {
			for (idx_t i = 0; i < size; i++) {
				le[2 * i] = le[2 * i] <= result.start_offset ? 0 : le[2 * i] - result.start_offset;
			}
		}
		return result;
	}
	if (size == 0) {
		return result;
	}
	idx_t cur_offset = 0;
	uint64_t *list_data = (uint64_t *)vector->data; /* list_entry_t {offset,length} */
	if (type->size_type == O_SIZE_NORMAL) {
		const uint32_t *offsets = (const uint32_t *)array->buffers[1] + effective_offset;
		result.start_offset = offsets[0];
		for (idx_t i = 0; i < size; i++) {
			list_data[2 * i] = cur_offset;
			list_data[2 * i + 1] = (uint32_t)(offsets[i + 1] - offsets[i]);
			cur_offset += list_data[2 * i + 1];
		}
		result.list_size = offsets[size];
	} else {
		const uint64_t *offsets = (const uint64_t *)array->buffers[1] + effective_offset;
		result.start_offset = offsets[0];
		for (idx_t i = 0; i < size; i++) {
			list_data[2 * i] = cur_offset;
			list_data[2 * i + 1] = offsets[i + 1] - offsets[i];
			cur_offset += list_data[2 * i + 1];
		}
		result.list_size = offsets[size];
	}
	result.list_size -= result.start_offset;
	return result;
}

/* string_t(const char*, uint32_t): duckdb/src/include/duckdb/common/types/string_type.hpp:43-64, layout :229-239 */
static void make_string_t(uint8_t *dst, const char *data, uint32_t len) {
	memcpy(dst, &len, 4);
	if (len <= 12) {
		memset(dst + 4, 0, 12);
		if (len) {
			memcpy(dst + 4, data, len);
		}
	} else {
		memcpy(dst + 4, data, 4);
		uint64_t p = (uint64_t)(uintptr_t)data;
		memcpy(dst + 8, &p, 8);
	}
}

/* ref :409-423 */
static int SetVectorString32(OState *st, OVector *vector, idx_t size, const char *cdata, const uint32_t *offsets) {
	(void)st;
	for (idx_t row = 0; row < size; row++) {
		if (!row_is_valid(vector, row)) {
			continue;
		}
		const char *cptr = cdata + offsets[row];
		uint32_t str_len = offsets[row + 1] - offsets[row];
		make_string_t(vector->data + 16 * row, cptr, str_len);
	}
	return O_OK;
}
static int SetVectorString64(OState *st, OVector *vector, idx_t size, const char *cdata, const uint64_t *offsets) {
	for (idx_t row = 0; row < size; row++) {
		if (!row_is_valid(vector, row)) {
			continue;
		}
		const char *cptr = cdata + offsets[row];
		uint64_t str_len = offsets[row + 1] - offsets[row];
		if (str_len > 0xFFFFFFFFull) {
			return fail(st, O_ECONV, "DuckDB does not support Strings over 4GB");
		}
		make_string_t(vector->data + 16 * row, cptr, (uint32_t)str_len);
	}
	return O_OK;
}

/* ref :425-452; view layout duckdb/src/include/duckdb/common/types/arrow_string_view_type.hpp:24-85 */
static void SetVectorStringView(OVector *vector, idx_t size, const OArray *array, idx_t current_pos) {
	const uint8_t *views = (const uint8_t *)array->buffers[1] + 16 * current_pos;
	for (idx_t row = 0; row < size; row++) {
		if (!row_is_valid(vector, row)) {
			continue;
		}
		const uint8_t *v = views + 16 * row;
		int32_t length;
		memcpy(&length, v, 4);
		if (length <= 12) {
			make_string_t(vector->data + 16 * row, (const char *)v + 4, (uint32_t)length);
		} else {
			int32_t buffer_index, offset;
			memcpy(&buffer_index, v + 8, 4);
			memcpy(&offset, v + 12, 4);
			const char *c_data = (const char *)array->buffers[2 + buffer_index];
			make_string_t(vector->data + 16 * row, c_data + offset, (uint32_t)length);
		}
	}
}

/* ref :342-390 — AddStringOrBlob copies into the vector's heap; the resulting string_t has the same length/content,
 * so the oracle points at the Arrow data instead (canonical form compares content, not the address space). */
static int ArrowToDuckDBBlob(OState *st, OVector *vector, const OArray *array, idx_t size, const OType *type,
                             int64_t nested_offset, int64_t parent_offset) {
	GetValidityMask(vector, array, st, size, parent_offset, nested_offset, 0);
	if (type->size_type == O_SIZE_FIXED) {
		idx_t fixed_size = (idx_t)type->fixed_size;
		idx_t offset = GetEffectiveOffset(array, parent_offset, st, nested_offset) * fixed_size;
		const char *cdata = (const char *)array->buffers[1];
		for (idx_t row = 0; row < size; row++) {
			if (!row_is_valid(vector, row)) {
				continue; /* NB: the reference does not advance `offset` for NULL rows (:352-360) */
			}
			make_string_t(vector->data + 16 * row, cdata + offset, (uint32_t)fixed_size);
			offset += fixed_size;
		}
	} else if (type->size_type == O_SIZE_NORMAL) {
		const uint32_t *offsets =
		    (const uint32_t *)array->buffers[1] + GetEffectiveOffset(array, parent_offset, st, nested_offset);
		return SetVectorString32(st, vector, size, (const char *)array->buffers[2], offsets);
	} else {
		if (((const uint64_t *)array->buffers[1])[array->length] > 0xFFFFFFFFull) {
			return fail(st, O_ECONV, "DuckDB does not support Blobs over 4GB");
		}
		const uint64_t *offsets =
		    (const uint64_t *)array->buffers[1] + GetEffectiveOffset(array, parent_offset, st, nested_offset);
		return SetVectorString64(st, vector, size, (const char *)array->buffers[2], offsets);
	}
	return O_OK;
}

/* ref :454-461 — the reference re-points the vector at the Arrow buffer; same bytes, so copy them */
static void DirectConversion(OVector *vector, const OArray *array, const OState *st, int64_t nested_offset,
                             uint64_t parent_offset, idx_t size) {
	idx_t w = (idx_t)vector->width;
	const uint8_t *p = (const uint8_t *)array->buffers[1] + w * GetEffectiveOffset(array, (int64_t)parent_offset, st, nested_offset);
	if (st->alias_direct) {
		if (!vector->borrowed_data) {
			free(vector->data);
		}
		vector->data = (uint8_t *)(uintptr_t)p; /* FlatVector::SetData(vector, data_ptr) */
		vector->borrowed_data = 1;
		return;
	}
	memcpy(vector->data, p, w * size);
}

/* TryMultiplyOperator int64: duckdb/src/function/scalar/operator/multiply.cpp:190-194 */
static int try_mul(int64_t a, int64_t b, int64_t *r) {
	return !__builtin_mul_overflow(a, b, r);
}

/* ref :463-478 */
static int TimeConversion(OState *st, OVector *vector, const OArray *array, int64_t nested_offset, int64_t parent_offset,
                          idx_t size, int64_t conversion, int src_is_32) {
	int64_t *tgt = (int64_t *)vector->data;
	idx_t off = GetEffectiveOffset(array, parent_offset, st, nested_offset);
	for (idx_t row = 0; row < size; row++) {
		if (!row_is_valid(vector, row)) {
			continue;
		}
		int64_t v = src_is_32 ? (int64_t)((const int32_t *)array->buffers[1])[off + row]
		                      : ((const int64_t *)array->buffers[1])[off + row];
		if (!try_mul(v, conversion, &tgt[row])) {
			return fail(st, O_ECONV, "Could not convert Time to Microsecond");
		}
	}
	return O_OK;
}

static uint64_t bswap64(uint64_t v) {
	return __builtin_bswap64(v);
}

/* ref :480-495 */
static void UUIDConversion(OVector *vector, const OArray *array, const OState *st, int64_t nested_offset,
                           int64_t parent_offset, idx_t size) {
	uint64_t *tgt = (uint64_t *)vector->data; /* hugeint_t {uint64 lower; int64 upper} */
	const uint64_t *src = (const uint64_t *)array->buffers[1] + 2 * GetEffectiveOffset(array, parent_offset, st, nested_offset);
	for (idx_t row = 0; row < size; row++) {
		if (!row_is_valid(vector, row)) {
			continue;
		}
		uint64_t src_lower = src[2 * row], src_upper = src[2 * row + 1];
		tgt[2 * row] = bswap64(src_upper);
		tgt[2 * row + 1] = bswap64(src_lower) ^ ((uint64_t)1 << 63);
	}
}

/* ref :513-525 — no validity check: garbage under NULL slots can raise */
static int IntervalConversionUs(OState *st, OVector *vector, const OArray *array, int64_t nested_offset,
                                int64_t parent_offset, idx_t size, int64_t conversion) {
	const int64_t *src = (const int64_t *)array->buffers[1] + GetEffectiveOffset(array, parent_offset, st, nested_offset);
	for (idx_t row = 0; row < size; row++) {
		int32_t zero = 0;
		int64_t micros;
		memcpy(vector->data + 16 * row, &zero, 4);     /* months */
		memcpy(vector->data + 16 * row + 4, &zero, 4); /* days */
		if (!try_mul(src[row], conversion, &micros)) {
			return fail(st, O_ECONV, "Could not convert Interval to Microsecond");
		}
		memcpy(vector->data + 16 * row + 8, &micros, 8);
	}
	return O_OK;
}

static void store_interval(OVector *vector, idx_t row, int32_t months, int32_t days, int64_t micros) {
	memcpy(vector->data + 16 * row, &months, 4);
	memcpy(vector->data + 16 * row + 4, &days, 4);
	memcpy(vector->data + 16 * row + 8, &micros, 8);
}

/* hugeint -> narrower integer: HugeintTryCastInteger, duckdb/src/common/types/hugeint.cpp:559-582 */
static int hugeint_try_cast(uint64_t lower, int64_t upper, int64_t max, int64_t *result) {
	if (upper == 0) {
		if (lower <= (uint64_t)max) {
			*result = (int64_t)lower;
			return 1;
		}
	} else if (upper == -1) {
		if (lower >= UINT64_MAX - (uint64_t)max) {
			*result = -(int64_t)(UINT64_MAX - lower) - 1;
			return 1;
		}
	}
	return 0;
}

/* ref :768-844 */
static int ConvertDecimal(OState *st, OVector *vector, const OArray *array, idx_t size, int64_t nested_offset,
                          uint64_t parent_offset, const OType *type) {
	idx_t off = GetEffectiveOffset(array, (int64_t)parent_offset, st, nested_offset);
	int src_bytes = type->bit_width / 8;
	if (src_bytes == vector->width) {
		DirectConversion(vector, array, st, nested_offset, parent_offset, size);
		return O_OK;
	}
	static const int64_t maxes[9] = {0, 0, 32767, 0, 2147483647ll, 0, 0, 0, INT64_MAX};
	for (idx_t row = 0; row < size; row++) {
		if (!row_is_valid(vector, row)) {
			continue;
		}
		uint64_t lower;
		int64_t upper;
		if (src_bytes == 4) {
			int32_t v = ((const int32_t *)array->buffers[1])[off + row];
			lower = (uint64_t)(int64_t)v;
			upper = v < 0 ? -1 : 0;
		} else if (src_bytes == 8) {
			int64_t v = ((const int64_t *)array->buffers[1])[off + row];
			lower = (uint64_t)v;
			upper = v < 0 ? -1 : 0;
		} else {
			lower = ((const uint64_t *)array->buffers[1])[2 * (off + row)];
			upper = ((const int64_t *)array->buffers[1])[2 * (off + row) + 1];
		}
		if (vector->width == 16) {
			memcpy(vector->data + 16 * row, &lower, 8);
			memcpy(vector->data + 16 * row + 8, &upper, 8);
			continue;
		}
		int64_t r;
		if (!hugeint_try_cast(lower, upper, maxes[vector->width], &r)) {
			continue; /* release build: slot left unwritten (:776-781) */
		}
		if (vector->width == 2) {
			int16_t x = (int16_t)r;
			memcpy(vector->data + 2 * row, &x, 2);
		} else if (vector->width == 4) {
			int32_t x = (int32_t)r;
			memcpy(vector->data + 4 * row, &x, 4);
		} else {
			memcpy(vector->data + 8 * row, &r, 8);
		}
	}
	return O_OK;
}

/* ------------------------------------------------------------------------------------------ */
static int ColumnArrowToDuckDB(OState *st, OVector *vector, const OArray *array, idx_t size, const OType *type,
                               int64_t nested_offset, OVector *parent_mask, uint64_t parent_offset);
static int ColumnArrowToDuckDBDictionary(OState *st, OVector *vector, const OArray *array, idx_t size,
                                         const OType *type, int64_t nested_offset, OVector *parent_mask,
                                         uint64_t parent_offset);

static int ColumnArrowToDuckDBRunEndEncoded(OState *st, OVector *vector, const OArray *array, idx_t size,
                                            const OType *type, int64_t nested_offset, OVector *parent_mask,
                                            uint64_t parent_offset);

static int ConvertChild(OState *st, OVector *child_vector, const OArray *child_array, idx_t size, const OType *child_type,
                        int64_t nested_offset, OVector *parent_mask, uint64_t parent_offset) {
	if (child_type->run_end_encoded) {
		return ColumnArrowToDuckDBRunEndEncoded(st, child_vector, child_array, size, child_type, nested_offset,
		                                        parent_mask, parent_offset);
	}
	if (child_type->has_dictionary) {
		return ColumnArrowToDuckDBDictionary(st, child_vector, child_array, size, child_type, nested_offset, parent_mask,
		                                     parent_offset);
	}
	return ColumnArrowToDuckDB(st, child_vector, child_array, size, child_type, nested_offset, parent_mask, parent_offset);
}

static void and_parent_mask(OVector *vector, const OVector *parent_mask, idx_t size) {
	if (parent_mask && !all_valid(parent_mask)) {
		for (idx_t i = 0; i < size; i++) {
			if (!row_is_valid(parent_mask, i)) {
				set_invalid(vector, i);
			}
		}
	}
}

/* ref :226-283 */
static int ArrowToDuckDBList(OState *st, OVector *vector, const OArray *array, idx_t size, const OType *type,
                             int64_t nested_offset, OVector *parent_mask, int64_t parent_offset) {
	GetValidityMask(vector, array, st, size, parent_offset, nested_offset, 0);
	idx_t effective_offset = GetEffectiveOffset(array, parent_offset, st, nested_offset);
	ListOffsetData ld = ConvertArrowListOffsets(vector, array, size, type, effective_offset);
	/* ListVector::Reserve + SetListSize */
	vector_free(vector->children[0]);
	vector->children[0] = vector_new(type->children[0], ld.list_size > 0 ? ld.list_size : 1);
	vector->list_size = ld.list_size;
	OVector *child_vector = vector->children[0];
	GetValidityMask(child_vector, array->children[0], st, ld.list_size, array->offset, (int64_t)ld.start_offset, 0);
	and_parent_mask(vector, parent_mask, size);
	const OArray *child_array = array->children[0];
	const OType *child_type = type->children[0];
	if (ld.list_size == 0 && ld.start_offset == 0) {
		return ColumnArrowToDuckDB(st, child_vector, child_array, ld.list_size, child_type, -1, NULL, 0);
	}
	return ConvertChild(st, child_vector, child_array, ld.list_size, child_type, (int64_t)ld.start_offset, NULL, 0);
}

/* ref :285-340 */
static int ArrowToDuckDBArray(OState *st, OVector *vector, const OArray *array, idx_t size, const OType *type,
                              int64_t nested_offset, OVector *parent_mask, int64_t parent_offset) {
	idx_t array_size = (idx_t)type->fixed_size;
	idx_t child_count = array_size * size;
	idx_t child_offset = GetEffectiveOffset(array, parent_offset, st, nested_offset) * array_size;
	GetValidityMask(vector, array, st, size, parent_offset, nested_offset, 0);
	OVector *child_vector = vector->children[0];
	if (child_vector->capacity < child_count) { /* ArrayVector child holds capacity*array_size */
		vector_free(child_vector);
		child_vector = vector->children[0] = vector_new(type->children[0], child_count);
	}
	GetValidityMask(child_vector, array->children[0], st, child_count, array->offset, (int64_t)child_offset, 0);
	and_parent_mask(vector, parent_mask, size);
	if (!all_valid(vector)) {
		for (idx_t i = 0; i < size; i++) {
			if (!row_is_valid(vector, i)) {
				for (idx_t j = 0; j < array_size; j++) {
					set_invalid(child_vector, i * array_size + j);
				}
			}
		}
	}
	const OArray *child_array = array->children[0];
	const OType *child_type = type->children[0];
	if (child_count == 0 && child_offset == 0) {
		return ColumnArrowToDuckDB(st, child_vector, child_array, child_count, child_type, -1, NULL, 0);
	}
	if (child_array->dictionary) {
		return ColumnArrowToDuckDBDictionary(st, child_vector, child_array, child_count, child_type, (int64_t)child_offset,
		                                     NULL, 0);
	}
	return ColumnArrowToDuckDB(st, child_vector, child_array, child_count, child_type, (int64_t)child_offset, NULL, 0);
}

/* ref :846-1225 */
static int ColumnArrowToDuckDB(OState *st, OVector *vector, const OArray *array, idx_t size, const OType *type,
                               int64_t nested_offset, OVector *parent_mask, uint64_t parent_offset) {
	const int64_t po = (int64_t)parent_offset;
	switch (type->id) {
	case O_SQLNULL:
		vector->is_constant_null = 1; /* vector.Reference(Value()) */
		break;
	case O_BOOLEAN: { /* :869-891 */
		if (type->bool8) {
			/* :851-860 extension with arrow_to_duckdb: the storage is converted into a TINYINT vector (DirectConversion),
			 * then ArrowBool8::ArrowToDuck (arrow_type_extension.cpp:344-350): result_ptr[i] = source_ptr[i] (int8 -> bool) */
			idx_t off = GetEffectiveOffset(array, po, st, nested_offset);
			const int8_t *source_ptr = (const int8_t *)array->buffers[1] + off;
			for (idx_t i = 0; i < size; i++) {
				vector->data[i] = source_ptr[i] != 0;
			}
			break;
		}
		idx_t effective_offset = GetEffectiveOffset(array, po, st, nested_offset);
		const uint8_t *src_ptr = (const uint8_t *)array->buffers[1] + effective_offset / 8;
		uint8_t *tgt_ptr = vector->data;
		int src_pos = 0;
		idx_t cur_bit = effective_offset % 8;
		for (idx_t row = 0; row < size; row++) {
			tgt_ptr[row] = (src_ptr[src_pos] & (1 << cur_bit)) == 0 ? 0 : 1;
			cur_bit++;
			if (cur_bit == 8) {
				src_pos++;
				cur_bit = 0;
			}
		}
		break;
	}
	case O_TINYINT: case O_SMALLINT: case O_INTEGER: case O_FLOAT: case O_DOUBLE: case O_UTINYINT: case O_USMALLINT:
	case O_UINTEGER: case O_UBIGINT: case O_BIGINT: case O_HUGEINT: case O_UHUGEINT: case O_TIMESTAMP:
	case O_TIMESTAMP_SEC: case O_TIMESTAMP_MS: case O_TIMESTAMP_NS: case O_TIME_TZ: /* :892-911 */
		DirectConversion(vector, array, st, nested_offset, parent_offset, size);
		break;
	case O_UUID:
		UUIDConversion(vector, array, st, nested_offset, po, size);
		break;
	case O_JSON: /* LogicalType::JSON() is VARCHAR with an alias */
	case O_VARCHAR: { /* :915-942 */
		idx_t off = GetEffectiveOffset(array, po, st, nested_offset);
		switch (type->size_type) {
		case O_SIZE_SUPER:
			TRY(SetVectorString64(st, vector, size, (const char *)array->buffers[2], (const uint64_t *)array->buffers[1] + off));
			break;
		case O_SIZE_NORMAL:
		case O_SIZE_FIXED:
			TRY(SetVectorString32(st, vector, size, (const char *)array->buffers[2], (const uint32_t *)array->buffers[1] + off));
			break;
		default:
			SetVectorStringView(vector, size, array, off);
			break;
		}
		break;
	}
	case O_DATE: /* :943-966 */
		if (type->datetime == O_DT_DAYS) {
			DirectConversion(vector, array, st, nested_offset, parent_offset, size);
		} else if (type->datetime == O_DT_MILLISECONDS) {
			const uint64_t *src = (const uint64_t *)array->buffers[1] + GetEffectiveOffset(array, po, st, nested_offset);
			int32_t *tgt = (int32_t *)vector->data;
			for (idx_t row = 0; row < size; row++) {
				tgt[row] = (int32_t)((int64_t)src[row] / (int64_t)(1000 * 60 * 60 * 24));
			}
		} else {
			return fail(st, O_ENOTIMPL, "Unsupported precision for Date Type ");
		}
		break;
	case O_TIME: /* :967-999 */
		switch (type->datetime) {
		case O_DT_SECONDS:
			TRY(TimeConversion(st, vector, array, nested_offset, po, size, 1000000, 1));
			break;
		case O_DT_MILLISECONDS:
			TRY(TimeConversion(st, vector, array, nested_offset, po, size, 1000, 1));
			break;
		case O_DT_MICROSECONDS:
			TRY(TimeConversion(st, vector, array, nested_offset, po, size, 1, 0));
			break;
		case O_DT_NANOSECONDS: {
			const int64_t *src = (const int64_t *)array->buffers[1] + GetEffectiveOffset(array, po, st, nested_offset);
			int64_t *tgt = (int64_t *)vector->data;
			for (idx_t row = 0; row < size; row++) {
				tgt[row] = src[row] / 1000;
			}
			break;
		}
		default:
			return fail(st, O_ENOTIMPL, "Unsupported precision for Time Type ");
		}
		break;
	case O_TIMESTAMP_TZ: /* :1000-1031; TimestampTZConversion :497-511 has the same body as TimeConversion<int64_t> */
		switch (type->datetime) {
		case O_DT_SECONDS:
			if (TimeConversion(st, vector, array, nested_offset, po, size, 1000000, 0)) {
				return fail(st, O_ECONV, "Could not convert TimestampTZ to Microsecond");
			}
			break;
		case O_DT_MILLISECONDS:
			if (TimeConversion(st, vector, array, nested_offset, po, size, 1000, 0)) {
				return fail(st, O_ECONV, "Could not convert TimestampTZ to Microsecond");
			}
			break;
		case O_DT_MICROSECONDS:
			DirectConversion(vector, array, st, nested_offset, parent_offset, size);
			break;
		case O_DT_NANOSECONDS: {
			const int64_t *src = (const int64_t *)array->buffers[1] + GetEffectiveOffset(array, po, st, nested_offset);
			int64_t *tgt = (int64_t *)vector->data;
			for (idx_t row = 0; row < size; row++) {
				tgt[row] = src[row] / 1000;
			}
			break;
		}
		default:
			return fail(st, O_ENOTIMPL, "Unsupported precision for TimestampTZ Type ");
		}
		break;
	case O_INTERVAL: { /* :1032-1077 */
		idx_t off = GetEffectiveOffset(array, po, st, nested_offset);
		switch (type->datetime) {
		case O_DT_SECONDS:
			TRY(IntervalConversionUs(st, vector, array, nested_offset, po, size, 1000000));
			break;
		case O_DT_DAYS:
		case O_DT_MILLISECONDS:
			TRY(IntervalConversionUs(st, vector, array, nested_offset, po, size, 1000));
			break;
		case O_DT_MICROSECONDS:
			TRY(IntervalConversionUs(st, vector, array, nested_offset, po, size, 1));
			break;
		case O_DT_NANOSECONDS: {
			const int64_t *src = (const int64_t *)array->buffers[1] + off;
			for (idx_t row = 0; row < size; row++) {
				store_interval(vector, row, 0, 0, src[row] / 1000);
			}
			break;
		}
		case O_DT_MONTHS: { /* :527-537 */
			const int32_t *src = (const int32_t *)array->buffers[1] + off;
			for (idx_t row = 0; row < size; row++) {
				store_interval(vector, row, src[row], 0, 0);
			}
			break;
		}
		case O_DT_MONTH_DAY_NANO: { /* :539-549, ArrowInterval duckdb/src/include/duckdb/function/table/arrow.hpp:23-31 */
			const uint8_t *src = (const uint8_t *)array->buffers[1] + 16 * off;
			for (idx_t row = 0; row < size; row++) {
				int32_t months, days;
				int64_t nanos;
				memcpy(&months, src + 16 * row, 4);
				memcpy(&days, src + 16 * row + 4, 4);
				memcpy(&nanos, src + 16 * row + 8, 8);
				store_interval(vector, row, months, days, nanos / 1000);
			}
			break;
		}
		default:
			return fail(st, O_ENOTIMPL, "Unsupported precision for Interval/Duration Type ");
		}
		break;
	}
	case O_DECIMAL: /* :1078-1108 */
		TRY(ConvertDecimal(st, vector, array, size, nested_offset, parent_offset, type));
		break;
	case O_BIT:
	case O_VARINT:
	case O_BLOB: /* :1109-1115 */
		TRY(ArrowToDuckDBBlob(st, vector, array, size, type, nested_offset, po));
		break;
	case O_LIST:
	case O_MAP: /* :1116-1131 (MapVerify is not restated: it only raises on duplicate / NULL keys) */
		TRY(ArrowToDuckDBList(st, vector, array, size, type, nested_offset, parent_mask, po));
		break;
	case O_ARRAY:
		TRY(ArrowToDuckDBArray(st, vector, array, size, type, nested_offset, parent_mask, po));
		break;
	case O_STRUCT: /* :1132-1172 */
		for (idx_t c = 0; c < (idx_t)array->n_children; c++) {
			OVector *child_entry = vector->children[c];
			const OArray *child_array = array->children[c];
			GetValidityMask(child_entry, child_array, st, size, array->offset, nested_offset, 0);
			if (!all_valid(vector)) {
				for (idx_t i = 0; i < size; i++) {
					if (!row_is_valid(vector, i)) {
						set_invalid(child_entry, i);
					}
				}
			}
			TRY(ConvertChild(st, child_entry, child_array, size, type->children[c], nested_offset, vector,
			                 (uint64_t)array->offset));
		}
		break;
	case O_UNION: { /* :1173-1221 (sparse union; the vector here = DuckDB's STRUCT{tag, members...} with the tag child
	                   folded into the node's own data + validity) */
		const int8_t *type_ids = (const int8_t *)array->buffers[array->n_buffers == 1 ? 0 : 1];
		idx_t n = (idx_t)array->n_children;
		OVector **children = (OVector **)calloc(n, sizeof(OVector *));
		int rc = O_OK;
		for (idx_t c = 0; c < n && rc == O_OK; c++) {
			children[c] = vector_new(type->children[c], size);
			GetValidityMask(children[c], array->children[c], st, size, (int64_t)po, nested_offset, 0); /* :1187 */
			/* :1190-1203 — the member is converted with the union's mask as parent_mask and parent_offset = false (0) */
			rc = ConvertChild(st, children[c], array->children[c], size, type->children[c], nested_offset, vector, 0);
		}
		for (idx_t row_idx = 0; row_idx < size && rc == O_OK; row_idx++) {
			/* :1209 type_ids[row_idx]: NOT offset by the chunk / parent offset the members were converted at */
			int8_t raw = type_ids[row_idx];
			if (raw < 0) { /* NumericCast<uint8_t> */
				char msg[128];
				snprintf(msg, sizeof(msg), "Information loss on integer cast: value %d outside of target range [0, 255]", raw);
				rc = fail(st, O_EINVAL, msg);
				break;
			}
			uint8_t tag = (uint8_t)raw;
			if (tag >= n) {
				char msg[64];
				snprintf(msg, sizeof(msg), "Arrow union tag out of range: %d", tag);
				rc = fail(st, O_EINVAL, msg);
				break;
			}
			const OVector *src = children[tag];
			int is_null = src->is_constant_null || !row_is_valid(src, row_idx);
			/* :1216-1217 vector.SetValue(row, value.IsNull() ? Value() : Value::UNION(members, tag, value)); Vector::SetValue
			 * (vector.cpp:442-517): validity.Set(row, !null), then child by child */
			if (is_null) {
				set_invalid(vector, row_idx);
				for (idx_t c = 0; c < n; c++) {
					set_invalid(vector->children[c], row_idx);
				}
			} else {
				set_valid(vector, row_idx);
				vector->data[row_idx] = tag;
				for (idx_t c = 0; c < n; c++) {
					OVector *dst = vector->children[c];
					if (c == tag) {
						set_valid(dst, row_idx);
						memcpy(dst->data + row_idx * (idx_t)dst->width, src->data + row_idx * (idx_t)src->width,
						       (size_t)dst->width);
					} else {
						set_invalid(dst, row_idx);
					}
				}
			}
		}
		for (idx_t c = 0; c < n; c++) {
			vector_free(children[c]);
		}
		free(children);
		TRY(rc);
		break;
	}
	default:
		return fail(st, O_ENOTIMPL, "Unsupported type for arrow conversion");
	}
	return O_OK;
}

/* ref :1227-1348 — sel.set_index(row, UnsafeNumericCast<idx_t>(indices[row])), sel_t = uint32_t */
static int SetSelectionVector(OState *st, uint32_t *sel, const uint8_t *indices, int index_type, idx_t size,
                              const OVector *mask, idx_t last_element_pos) {
	for (idx_t row = 0; row < size; row++) {
		uint64_t v;
		switch (index_type) {
		case O_UTINYINT: v = ((const uint8_t *)indices)[row]; break;
		case O_TINYINT: v = (uint64_t)(int64_t)((const int8_t *)indices)[row]; break;
		case O_USMALLINT: v = ((const uint16_t *)indices)[row]; break;
		case O_SMALLINT: v = (uint64_t)(int64_t)((const int16_t *)indices)[row]; break;
		case O_UINTEGER: v = ((const uint32_t *)indices)[row]; break;
		case O_INTEGER: v = (uint64_t)(int64_t)((const int32_t *)indices)[row]; break;
		case O_UBIGINT: v = ((const uint64_t *)indices)[row]; break;
		case O_BIGINT: v = (uint64_t)((const int64_t *)indices)[row]; break;
		default:
			return fail(st, O_ENOTIMPL, "(Arrow) Unsupported type for selection vectors");
		}
		if (mask) {
			if ((index_type == O_UINTEGER || index_type == O_UBIGINT || index_type == O_BIGINT) &&
			    last_element_pos > 0xFFFFFFFFull) {
				return fail(st, O_ECONV, "DuckDB only supports indices that fit on an uint32");
			}
			sel[row] = row_is_valid(mask, row) ? (uint32_t)v : (uint32_t)last_element_pos;
		} else {
			if ((index_type == O_UBIGINT || index_type == O_BIGINT) && last_element_pos > 0xFFFFFFFFull &&
			    v > 0xFFFFFFFFull) {
				return fail(st, O_ECONV, "DuckDB only supports indices that fit on an uint32");
			}
			sel[row] = (uint32_t)v;
		}
	}
	return O_OK;
}

static int index_width(int index_type) {
	switch (index_type) {
	case O_UTINYINT: case O_TINYINT: return 1;
	case O_USMALLINT: case O_SMALLINT: return 2;
	case O_UINTEGER: case O_INTEGER: return 4;
	default: return 8;
	}
}

/* ref :1350-1419 */
static int ColumnArrowToDuckDBDictionary(OState *st, OVector *vector, const OArray *array, idx_t size,
                                         const OType *type, int64_t nested_offset, OVector *parent_mask,
                                         uint64_t parent_offset) {
	const int has_nulls = array->null_count > 0 || (parent_mask && !all_valid(parent_mask)); /* CanContainNull */
	OVector *base = NULL;
	for (int i = 0; i < st->n_dicts; i++) {
		if (st->dict_key[i] == (const void *)array->dictionary) {
			base = st->dict_vec[i];
		}
	}
	if (!base) { /* CacheOutdated */
		idx_t dl = (idx_t)array->dictionary->length;
		OType value_type = *type;
		value_type.has_dictionary = 0;
		base = vector_new(&value_type, dl + 1);
		OState dst = *st;
		dst.chunk_offset = 0; /* SetValidityMask(..., parent_offset=0, nested_offset=0, has_nulls) */
		GetValidityMask(base, array->dictionary, &dst, dl, 0, 0, has_nulls);
		/* ColumnArrowToDuckDB(*base_vector, *array.dictionary, array_state, length, dictionary_type): nested_offset -1,
		 * so the effective offset is dictionary.offset + state.chunk_offset — the reference converts the dictionary on
		 * the first chunk it meets, where chunk_offset is whatever the scan is at (0 for the first chunk). */
		OState cst = *st;
		TRY(ColumnArrowToDuckDB(&cst, base, array->dictionary, dl, &value_type, -1, NULL, 0));
		if (st->n_dicts >= O_MAX_DICTS) {
			return fail(st, O_ENOTIMPL, "too many dictionaries");
		}
		st->dict_key[st->n_dicts] = (const void *)array->dictionary;
		st->dict_vec[st->n_dicts++] = base;
	}
	const uint8_t *indices = (const uint8_t *)array->buffers[1] +
	                         (idx_t)index_width(type->index_type) *
	                             GetEffectiveOffset(array, (int64_t)parent_offset, st, nested_offset);
	free(vector->sel);
	vector->sel = (uint32_t *)calloc(size ? size : 1, sizeof(uint32_t));
	if (has_nulls) {
		OType dummy;
		memset(&dummy, 0, sizeof(dummy));
		OVector *indices_validity = vector_new(&dummy, size);
		GetValidityMask(indices_validity, array, st, size, (int64_t)parent_offset, -1, 0);
		and_parent_mask(indices_validity, parent_mask, size);
		int rc = SetSelectionVector(st, vector->sel, indices, type->index_type, size, indices_validity,
		                            (idx_t)array->dictionary->length);
		vector_free(indices_validity);
		TRY(rc);
	} else {
		TRY(SetSelectionVector(st, vector->sel, indices, type->index_type, size, NULL, 0));
	}
	vector->dict = base;
	vector->dict_len = (idx_t)array->dictionary->length;
	return O_OK;
}

/* ------------------------------------------------------------------------------------------ slabs */
static OSlab *slab_new(const OType *type) {
	OSlab *s = (OSlab *)calloc(1, sizeof(OSlab));
	s->width = type->has_dictionary ? 0 : type->width;
	s->type_id = type->id;
	if (!type->has_dictionary && (type->id == O_STRUCT || type->id == O_LIST || type->id == O_MAP || type->id == O_ARRAY ||
	                              type->id == O_UNION)) {
		s->n_children = type->n_children;
		s->children = (OSlab **)calloc((size_t)type->n_children, sizeof(OSlab *));
		for (int i = 0; i < type->n_children; i++) {
			s->children[i] = slab_new(type->children[i]);
		}
	}
	return s;
}

void oracle_slab_free(OSlab *s) {
	if (!s) {
		return;
	}
	for (int i = 0; i < s->n_children; i++) {
		oracle_slab_free(s->children[i]);
	}
	free(s->children);
	free(s->data);
	free(s->valid);
	free(s->sel);
	oracle_slab_free(s->dict);
	free(s);
}

static void slab_reserve(OSlab *s, uint64_t extra, int with_sel) {
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

static void slab_append(OSlab *s, const OVector *v, const OType *type, idx_t count) {
	slab_reserve(s, count, v->sel != NULL);
	for (idx_t i = 0; i < count; i++) {
		s->valid[s->rows + i] = (uint8_t)(v->is_constant_null ? 0 : row_is_valid(v, i));
	}
	if (v->is_constant_null) {
		s->all_null = 1;
	}
	if (v->sel) {
		memcpy(s->sel + s->rows, v->sel, count * 4);
		if (!s->dict) {
			OType value_type = *type;
			value_type.has_dictionary = 0;
			s->dict = slab_new(&value_type);
			/* dict_len values + the extra NULL slot if the mask was resized for it */
			idx_t n = v->dict_len + ((v->dict->validity && v->dict->validity_cap > v->dict_len) ? 1 : 0);
			if (n > v->dict->capacity) {
				n = v->dict->capacity;
			}
			slab_append(s->dict, v->dict, &value_type, n);
		}
	} else if (s->width) {
		memcpy(s->data + s->rows * (uint64_t)s->width, v->data, count * (uint64_t)s->width);
	}
	if (!v->sel) {
		if (type->id == O_STRUCT || type->id == O_UNION) {
			for (int c = 0; c < type->n_children; c++) {
				slab_append(s->children[c], v->children[c], type->children[c], count);
			}
		} else if (type->id == O_LIST || type->id == O_MAP) {
			if (v->children[0]) {
				slab_append(s->children[0], v->children[0], type->children[0], v->list_size);
			}
		} else if (type->id == O_ARRAY) {
			slab_append(s->children[0], v->children[0], type->children[0], count * (idx_t)type->fixed_size);
		}
	}
	s->rows += count;
}

/* ref :551-577 FindRunIndex: first run end strictly greater than offset, count if none */
static idx_t FindRunIndex(const uint8_t *run_ends, int width, idx_t count, idx_t offset) {
	idx_t begin = 0, end = count;
	while (begin < end) {
		idx_t middle = (begin + end) / 2;
		idx_t v = width == 2 ? (idx_t)((const int16_t *)run_ends)[middle] : (idx_t)((const int32_t *)run_ends)[middle];
		if (offset >= v) {
			begin = middle + 1;
		} else {
			end = middle;
		}
	}
	return begin;
}

/* ref :579-663 FlattenRunEnds<RUN_END_TYPE, VALUE_TYPE>: VALUE_TYPE is copied as `width` bytes */
static void FlattenRunEnds(OVector *result, const OVector *runs, const OVector *values, int run_width,
                           idx_t compressed_size, idx_t scan_offset, idx_t count) {
	const uint8_t *run_ends_data = runs->data;
	const int w = values->width;
	idx_t run = FindRunIndex(run_ends_data, run_width, compressed_size, scan_offset);
	idx_t logical_index = scan_offset, index = 0;
	const int values_all_valid = all_valid(values);
	for (; run < compressed_size; ++run) {
		idx_t run_end = run_width == 2 ? (idx_t)((const int16_t *)run_ends_data)[run]
		                               : (idx_t)((const int32_t *)run_ends_data)[run];
		idx_t to_scan = run_end - (logical_index + index);
		if (to_scan > count - index) {
			to_scan = count - index;
		}
		if (values_all_valid || row_is_valid(values, run)) {
			for (idx_t i = 0; i < to_scan; i++) {
				memcpy(result->data + (index + i) * (idx_t)w, values->data + run * (idx_t)w, (size_t)w);
				/* validity.SetValid on a mask that was never allocated is a no-op; an allocated one gets the bit */
				if (!values_all_valid && result->validity) {
					result->validity[(index + i) >> 6] |= (uint64_t)1 << ((index + i) & 63);
				}
			}
		} else {
			for (idx_t i = 0; i < to_scan; i++) {
				set_invalid(result, index + i);
			}
		}
		index += to_scan;
		if (index >= count) {
			break;
		}
	}
}

/* ref :721-767 ColumnArrowToDuckDBRunEndEncoded */
static int ColumnArrowToDuckDBRunEndEncoded(OState *st, OVector *vector, const OArray *array, idx_t size,
                                            const OType *type, int64_t nested_offset, OVector *parent_mask,
                                            uint64_t parent_offset) {
	(void)parent_mask; /* the reference does not use it either */
	const OArray *run_ends_array = array->children[0], *values_array = array->children[1];
	const OType *run_ends_type = type->children[0], *values_type = type->children[1];
	idx_t compressed_size = (idx_t)run_ends_array->length;
	int slot = -1;
	for (int i = 0; i < st->n_rees; i++) {
		if (st->ree_key[i] == (const void *)array) {
			slot = i;
		}
	}
	if (slot < 0) {
		if (st->n_rees >= O_MAX_DICTS) {
			return fail(st, O_ENOTIMPL, "oracle: too many run-end encoded arrays in one scan");
		}
		slot = st->n_rees++;
		st->ree_key[slot] = (const void *)array;
		st->ree_runs[slot] = vector_new(run_ends_type, compressed_size);
		st->ree_values[slot] = vector_new(values_type, compressed_size);
		const int save_alias = st->alias_direct;
		TRY(ColumnArrowToDuckDB(st, st->ree_runs[slot], run_ends_array, compressed_size, run_ends_type, -1, NULL, 0));
		GetValidityMask(st->ree_values[slot], values_array, st, compressed_size, (int64_t)parent_offset, nested_offset, 0);
		TRY(ColumnArrowToDuckDB(st, st->ree_values[slot], values_array, compressed_size, values_type, -1, NULL, 0));
		st->alias_direct = save_alias;
	}
	idx_t scan_offset = GetEffectiveOffset(array, (int64_t)parent_offset, st, nested_offset);
	int run_width;
	switch (run_ends_type->id) {
	case O_SMALLINT:
		run_width = 2;
		break;
	case O_INTEGER:
	case O_BIGINT: /* ref :760-762: INT64 run ends are flattened through FlattenRunEndsSwitch<int32_t> */
		run_width = 4;
		break;
	default:
		return fail(st, O_ENOTIMPL, "Type not implemented for RunEndEncoding");
	}
	switch (values_type->width) {
	case 1:
	case 2:
	case 4:
	case 8:
	case 16:
		break;
	default:
		return fail(st, O_ENOTIMPL, "RunEndEncoded value type not supported yet");
	}
	if (values_type->id == O_LIST || values_type->id == O_MAP || values_type->id == O_UHUGEINT ||
	    values_type->id == O_UUID) { /* (BLOB is physical VARCHAR: FlattenRunEnds<.., string_t>, :706-711) */
		return fail(st, O_ENOTIMPL, "RunEndEncoded value type not supported yet");
	}
	FlattenRunEnds(vector, st->ree_runs[slot], st->ree_values[slot], run_width, compressed_size, scan_offset, size);
	return O_OK;
}

/* The scan loop: ArrowScanFunction (duckdb/src/function/table/arrow.cpp:172-201) + ArrowToDuckDB (:1421-1484) */
int oracle_scan_column(const OArray *array, const OType *type, int64_t parent_offset, OSlab **out, char *errbuf,
                       int errlen) {
	OState st;
	memset(&st, 0, sizeof(st));
	st.err = errbuf;
	st.errlen = errlen;
	OSlab *slab = slab_new(type);
	slab_reserve(slab, 0, 0);
	int rc = O_OK;
	for (idx_t chunk_offset = 0; chunk_offset < (idx_t)array->length && rc == O_OK; chunk_offset += O_VECTOR_SIZE) {
		idx_t output_size = (idx_t)array->length - chunk_offset;
		if (output_size > O_VECTOR_SIZE) {
			output_size = O_VECTOR_SIZE;
		}
		st.chunk_offset = chunk_offset;
		OType vt = *type;
		vt.has_dictionary = 0;
		OVector *vec = vector_new(&vt, O_VECTOR_SIZE); /* DataChunk vector, Reset() between calls */
		if (type->has_dictionary) {
			rc = ColumnArrowToDuckDBDictionary(&st, vec, array, output_size, type, -1, NULL, 0);
		} else if (type->run_end_encoded) { /* ArrowToDuckDB :1466-1469 */
			rc = ColumnArrowToDuckDBRunEndEncoded(&st, vec, array, output_size, type, -1, NULL, 0);
		} else {
			GetValidityMask(vec, array, &st, output_size, parent_offset, -1, 0); /* SetValidityMask :1477 */
			rc = ColumnArrowToDuckDB(&st, vec, array, output_size, type, -1, NULL, 0);
		}
		if (rc == O_OK) {
			slab_append(slab, vec, type, output_size);
		}
		vector_free(vec);
	}
	for (int i = 0; i < st.n_dicts; i++) {
		vector_free(st.dict_vec[i]);
	}
	for (int i = 0; i < st.n_rees; i++) {
		vector_free(st.ree_runs[i]);
		vector_free(st.ree_values[i]);
	}
	if (rc != O_OK) {
		oracle_slab_free(slab);
		slab = NULL;
	}
	*out = slab;
	return rc;
}

uint64_t oracle_scan_batch_checksum(const OArray *batch, OType *const *types, int n_cols) {
	uint64_t sum = 0;
	OState st;
	memset(&st, 0, sizeof(st));
	/* one DataChunk, reused (Reset) for every 2 048-row step like the executor does */
	st.alias_direct = 1;
	OVector **vecs = (OVector **)calloc((size_t)n_cols, sizeof(OVector *));
	for (int c = 0; c < n_cols; c++) {
		OType vt = *types[c];
		vt.has_dictionary = 0;
		vecs[c] = vector_new(&vt, O_VECTOR_SIZE);
	}
	for (idx_t chunk_offset = 0; chunk_offset < (idx_t)batch->length; chunk_offset += O_VECTOR_SIZE) {
		idx_t output_size = (idx_t)batch->length - chunk_offset;
		if (output_size > O_VECTOR_SIZE) {
			output_size = O_VECTOR_SIZE;
		}
		st.chunk_offset = chunk_offset;
		for (int c = 0; c < n_cols; c++) {
			const OArray *array = batch->children[c];
			/* DataChunk::Reset: validity back to "all valid" */
			free(vecs[c]->validity);
			vecs[c]->validity = NULL;
			vecs[c]->validity_cap = 0;
			if (types[c]->has_dictionary) {
				ColumnArrowToDuckDBDictionary(&st, vecs[c], array, output_size, types[c], -1, NULL, 0);
				sum += vecs[c]->sel[output_size - 1];
			} else {
				GetValidityMask(vecs[c], array, &st, output_size, batch->offset, -1, 0);
				ColumnArrowToDuckDB(&st, vecs[c], array, output_size, types[c], -1, NULL, 0);
				if (vecs[c]->width && vecs[c]->data) {
					sum += vecs[c]->data[(output_size - 1) * (idx_t)vecs[c]->width];
				}
			}
		}
	}
	for (int c = 0; c < n_cols; c++) {
		vector_free(vecs[c]);
	}
	free(vecs);
	for (int i = 0; i < st.n_dicts; i++) {
		vector_free(st.dict_vec[i]);
	}
	return sum;
}

/* ------------------------------------------------------------------------------------------------
 * a1 + a2 of SURVEY §8: what the reference does to a record-batch message BEFORE the scan loop.
 * ---------------------------------------------------------------------------------------------- */

/* IPCFileStreamReader::DecodeBody (src/ipc/stream_reader/ipc_file_stream_reader.cpp:71-89): a fresh allocation of
 * body_size bytes filled by BufferedFileReader::ReadData (duckdb/src/common/serializer/buffered_file_reader.cpp:25-49),
 * i.e. through a FILE_BUFFER_SIZE = 4096 byte bounce buffer
 * (duckdb/src/include/duckdb/common/serializer/buffered_file_writer.hpp:16).  fs.Read() is restated as a memcpy out of
 * the in-memory file image (page-cache-warm file, no syscall cost: a lower bound of the reference's cost). */
#define O_FILE_BUFFER_SIZE 4096
uint8_t *oracle_decode_body(const uint8_t *file_image, uint64_t body_len) {
	uint8_t *target = (uint8_t *)malloc(body_len ? body_len : 1);
	uint8_t bounce[O_FILE_BUFFER_SIZE];
	uint64_t file_pos = 0;
	uint8_t *target_buffer = target, *end_ptr = target + body_len;
	idx_t offset = 0, read_data = 0;
	for (;;) {
		idx_t to_read = (idx_t)(end_ptr - target_buffer);
		if (to_read > read_data - offset) {
			to_read = read_data - offset;
		}
		if (to_read > 0) {
			memcpy(target_buffer, bounce + offset, to_read);
			offset += to_read;
			target_buffer += to_read;
		}
		if (target_buffer < end_ptr) {
			offset = 0;
			read_data = body_len - file_pos < O_FILE_BUFFER_SIZE ? body_len - file_pos : O_FILE_BUFFER_SIZE;
			memcpy(bounce, file_image + file_pos, read_data); /* fs.Read(*handle, data.get(), FILE_BUFFER_SIZE) */
			file_pos += read_data;
		} else {
			return target;
		}
	}
}

/* nanoarrow ArrowArrayViewValidate(NANOARROW_VALIDATION_LEVEL_FULL) as called by ArrowIpcDecoderDecodeArrayFromShared
 * (call sites: src/ipc/stream_reader/base_stream_reader.cpp:117,124,135,139).  nanoarrow (apache/arrow-nanoarrow
 * @4bf5a93, 0.7.0-SNAPSHOT) is not in the reference tree; its published FULL level walks every offsets buffer and
 * requires offsets[i] >= offsets[i-1] (ArrowAssertIncreasingInt32/Int64), first offset >= 0 and last offset <= size
 * of the data buffer; recursion over children and the dictionary.  Returns 0 when valid. */
int oracle_validate_full(const OArray *array, const OType *type) {
	if (type->has_dictionary) {
		OType vt = *type;
		vt.has_dictionary = 0;
		return array->dictionary ? oracle_validate_full(array->dictionary, &vt) : 0;
	}
	int is_var = (type->id == O_VARCHAR || type->id == O_BLOB || type->id == O_JSON || type->id == O_BIT ||
	              type->id == O_VARINT) &&
	             (type->size_type == O_SIZE_NORMAL || type->size_type == O_SIZE_SUPER);
	int is_list = type->id == O_LIST || type->id == O_MAP;
	if ((is_var || (is_list && !type->list_view)) && array->length > 0 && array->n_buffers > 1 && array->buffers[1]) {
		idx_t n = (idx_t)(array->offset + array->length) + 1;
		if (type->size_type == O_SIZE_SUPER) {
			const int64_t *o = (const int64_t *)array->buffers[1];
			if (o[0] < 0) {
				return 1;
			}
			for (idx_t i = 1; i < n; i++) {
				if (o[i] < o[i - 1]) {
					return 1;
				}
			}
		} else {
			const int32_t *o = (const int32_t *)array->buffers[1];
			if (o[0] < 0) {
				return 1;
			}
			for (idx_t i = 1; i < n; i++) {
				if (o[i] < o[i - 1]) {
					return 1;
				}
			}
		}
	}
	if (is_list && type->list_view && array->length > 0 && array->n_children > 0) {
		/* list views: every [offset, offset + size) lies inside the child */
		uint64_t child_len = (uint64_t)array->children[0]->length;
		for (idx_t i = (idx_t)array->offset; i < (idx_t)(array->offset + array->length); i++) {
			uint64_t o = type->size_type == O_SIZE_SUPER ? ((const uint64_t *)array->buffers[1])[i]
			                                            : ((const uint32_t *)array->buffers[1])[i];
			uint64_t z = type->size_type == O_SIZE_SUPER ? ((const uint64_t *)array->buffers[2])[i]
			                                            : ((const uint32_t *)array->buffers[2])[i];
			if (o > child_len || z > child_len - o) {
				return 1;
			}
		}
	}
	for (int c = 0; c < type->n_children && c < (int)array->n_children; c++) {
		int rc = oracle_validate_full(array->children[c], type->children[c]);
		if (rc) {
			return rc;
		}
	}
	return 0;
}

/* One record-batch message the way the reference's read_arrow handles it: DecodeBody (a1), FULL validation (a2), then
 * the scan loop (a4-a9).  `body` is the message body inside the in-memory file image; `batch` describes the same body
 * as Arrow C data (sliced by pyarrow, untimed).  Timed CPU baseline only. */
uint64_t oracle_read_arrow_batch_checksum(const uint8_t *body, uint64_t body_len, const OArray *batch,
                                          OType *const *types, int n_cols) {
	uint8_t *copy = oracle_decode_body(body, body_len);
	uint64_t sum = body_len ? copy[body_len - 1] : 0;
	for (int c = 0; c < n_cols; c++) {
		if (oracle_validate_full(batch->children[c], types[c])) {
			free(copy);
			return UINT64_MAX;
		}
	}
	sum += oracle_scan_batch_checksum(batch, types, n_cols);
	free(copy);
	return sum;
}

