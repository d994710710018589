"""datafusion_parallelism_b200 — B200-native (sm_100a) parallel hash join behind the operator surface of
jamesfer/datafusion-parallelism.  All compute is in libb200join.so (hand-written CUDA, C ABI in
include/b200join.h); importing this package without the built library raises — there is no fallback."""
from . import _native
from ._native import NativeLibraryMissing, build_library, LIB_PATH
from .join import (HashJoin, PairResult, ColumnRef, B200JoinError, JOIN_TYPES, LAYOUTS, MEM_HOST, MEM_DEVICE, OUT_PAIRS, OUT_HOST,
                   OUT_MATCHES_ONLY, OUT_COUNT_ONLY, OUT_BUILD_U32, OUT_GATHER_HINT, OUT_PROBE_ORDER, NULL_BUILD, NULL_PROBE, T_UTF8, device_count, column, make_batch, PinnedArray)

_native.lib()   # fail loudly at import time when the CUDA extension has not been built

__all__ = ["HashJoin", "PairResult", "ColumnRef", "B200JoinError", "JOIN_TYPES", "LAYOUTS", "MEM_HOST", "MEM_DEVICE", "OUT_PAIRS", "OUT_HOST",
           "OUT_MATCHES_ONLY", "OUT_COUNT_ONLY", "OUT_BUILD_U32", "OUT_GATHER_HINT", "OUT_PROBE_ORDER", "NULL_BUILD", "NULL_PROBE", "T_UTF8", "device_count", "column", "make_batch", "PinnedArray",
           "build_library", "NativeLibraryMissing", "LIB_PATH"]
