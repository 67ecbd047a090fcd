"""b2arrow — B200-native Arrow IPC <-> DuckDB vector hot path (drop-in for the nanoarrow extension's scan path).

Importing this package loads libb2arrow.so (hand-written sm_100a kernels behind the C ABI of include/b2arrow.h)
and fails loudly if it has not been built: there is no CPU fallback."""
from ._lib import B2AError, lib, LIB_PATH, VECTOR_SIZE, B2A_NONE  # noqa: F401
from .api import (ArrowIPCReader, ArrowScan, ArrowWriter, ColumnView, DecodedScan, DeviceVector, Field,  # noqa: F401
                  copy_to_arrows, read_arrow, to_arrow_ipc,
                  scan_arrow_ipc)

__version__ = lib.b2a_version().decode()
