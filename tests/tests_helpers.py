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
