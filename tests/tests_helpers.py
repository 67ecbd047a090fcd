"""Small shared helpers for the tests."""
import pyarrow as pa


def same_rows(a: pa.Table, b: pa.Table) -> bool:
    """logical row equality, independent of field names inside nested types and of decimal bit widths"""
    if a.num_rows != b.num_rows or a.num_columns != b.num_columns:
        return False
    for i in range(a.num_columns):
        if a.column(i).to_pylist() != b.column(i).to_pylist():
            return False
    return True


def pylist(arr) -> list:
    """to_pylist that does not slice sparse unions: pyarrow's own to_pylist reads the members of a SLICED sparse union
    (which is what the scalars of a list<union> are) at the unsliced positions"""
    if isinstance(arr, pa.ChunkedArray):
        return [v for c in arr.chunks for v in pylist(c)]
    if (pa.types.is_list_view(arr.type) or pa.types.is_large_list_view(arr.type)) and pa.types.is_union(arr.type.value_type):
        vals, offs, sizes = arr.values.to_pylist(), arr.offsets.to_pylist(), arr.sizes.to_pylist()
        valid = arr.is_valid().to_pylist()
        return [vals[offs[i]: offs[i] + sizes[i]] if valid[i] else None for i in range(len(arr))]
    if (pa.types.is_list(arr.type) or pa.types.is_large_list(arr.type)) and pa.types.is_union(arr.type.value_type):
        vals = arr.values.to_pylist()
        offs = arr.offsets.to_pylist()
        valid = arr.is_valid().to_pylist()
        return [vals[offs[i]: offs[i + 1]] if valid[i] else None for i in range(len(arr))]
    return arr.to_pylist()
