"""ctypes binding of libb200join.so (include/b200join.h). No compute happens in Python and there is no
fallback: if the shared library is missing the import of the product fails loudly."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
LIB_PATH = os.environ.get("B200J_LIB") or os.path.join(_HERE, "libb200join.so")   # B200J_LIB: tuning variants only
CSRC = os.path.join(_HERE, "csrc")
HEADER = os.path.join(_ROOT, "include", "b200join.h")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared", "-cudart", "static", "-diag-suppress", "177",
]


def sources():
    out = [HEADER]
    for f in sorted(os.listdir(CSRC)):
        if f.endswith((".cu", ".cuh", ".inl", ".h")):
            out.append(os.path.join(CSRC, f))
    return out


def build_library(force: bool = False, verbose: bool = False) -> str:
    """nvcc-compile csrc/b200join.cu for sm_100a into libb200join.so (in-tree, so it travels to the GPU box)."""
    stale = (not os.path.exists(LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in sources())
    if not (force or stale):
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + ["-o", LIB_PATH, os.path.join(CSRC, "b200join.cu")]
    if verbose:
        print(" ".join(cmd), file=sys.stderr)
    subprocess.run(cmd, check=True)
    return LIB_PATH


# ---- structs (mirror include/b200join.h) -------------------------------------------------------------

class Column(C.Structure):
    _fields_ = [("values", C.c_void_p), ("offsets", C.c_void_p), ("validity", C.c_void_p), ("bit_offset", C.c_int64)]


class Batch(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("mem", C.c_int32), ("n_rows", C.c_int64), ("n_keys", C.c_int32),
                ("n_payload", C.c_int32), ("keys", C.POINTER(Column)), ("payload", C.POINTER(Column))]


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("join_type", C.c_int32), ("n_partitions", C.c_int32), ("device", C.c_int32),
                ("n_keys", C.c_int32), ("key_widths", C.POINTER(C.c_int32)),
                ("n_build_payload", C.c_int32), ("build_payload_types", C.POINTER(C.c_int32)),
                ("n_probe_payload", C.c_int32), ("probe_payload_types", C.POINTER(C.c_int32)),
                ("hash_keys", C.c_uint64 * 4), ("hash_flags", C.c_int32), ("layout", C.c_int32), ("radix_bits", C.c_int32),
                ("load_factor", C.c_float), ("expected_build_rows", C.c_int64), ("l2_budget_bytes", C.c_int64),
                ("stream", C.c_void_p), ("flags", C.c_uint32)]


class Result(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("mem", C.c_int32), ("n", C.c_int64), ("build_idx", C.c_void_p),
                ("probe_idx", C.c_void_p), ("n_build_null", C.c_int64), ("n_probe_null", C.c_int64), ("_owner", C.c_void_p),
                ("build_idx_width", C.c_int32), ("probe_order", C.c_int32)]


class OutColumn(C.Structure):
    _fields_ = [("values", C.c_void_p), ("offsets", C.c_void_p), ("validity", C.c_void_p), ("null_count", C.c_int64),
                ("values_bytes", C.c_int64)]


class Gathered(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("mem", C.c_int32), ("n_rows", C.c_int64), ("n_cols", C.c_int32),
                ("cols", C.POINTER(OutColumn)), ("_owner", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("layout", C.c_int32), ("build_rows", C.c_int64), ("build_null_keys", C.c_int64),
                ("distinct_keys", C.c_int64), ("duplicate_rows", C.c_int64), ("table_slots", C.c_int64), ("table_bytes", C.c_int64),
                ("radix_bits", C.c_int32), ("_pad", C.c_int32), ("probe_rows", C.c_int64), ("result_rows", C.c_int64),
                ("ms_build_copy", C.c_double), ("ms_build_partition", C.c_double), ("ms_build_insert", C.c_double),
                ("ms_probe_partition", C.c_double), ("ms_probe", C.c_double), ("ms_gather", C.c_double),
                ("probe_launches", C.c_int64), ("kernel_launches", C.c_int64), ("probe_retries", C.c_int64),
                ("dense_probe_launches", C.c_int64), ("compact_launches", C.c_int64), ("ms_exchange_partition", C.c_double),
                ("seg_probe_launches", C.c_int64), ("narrow_probe_launches", C.c_int64), ("gather_seg_launches", C.c_int64),
                ("gather_hint_launches", C.c_int64), ("ms_slot_order", C.c_double), ("probe_order_results", C.c_int64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if not n.startswith("_") and n != "struct_size"}


# every symbol include/b200join.h declares (tests/test_boundary.py checks the header against this and the .so)
SIGNATURES = {
    "b200j_abi_version": (C.c_int32, []),
    "b200j_device_count": (C.c_int32, []),
    "b200j_create": (C.c_int32, [C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "b200j_destroy": (C.c_int32, [C.c_void_p]),
    "b200j_last_error": (C.c_char_p, [C.c_void_p]),
    "b200j_get_stats": (C.c_int32, [C.c_void_p, C.POINTER(Stats)]),
    "b200j_reset_stats": (C.c_int32, [C.c_void_p]),
    "b200j_synchronize": (C.c_int32, [C.c_void_p]),
    "b200j_reset": (C.c_int32, [C.c_void_p]),
    "b200j_trim": (C.c_int32, [C.c_void_p]),
    "b200j_build_push": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(Batch)]),
    "b200j_build_finish": (C.c_int32, [C.c_void_p, C.c_int32]),
    "b200j_probe_batch": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(Batch), C.c_uint32, C.POINTER(Result)]),
    "b200j_probe_pieces": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.c_int32, C.c_uint32, C.POINTER(Result)]),
    "b200j_apply_pairs": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                                      C.c_uint32, C.POINTER(Result)]),
    "b200j_probe_finish": (C.c_int32, [C.c_void_p, C.c_int32, C.c_uint32, C.POINTER(Result)]),
    "b200j_result_release": (C.c_int32, [C.c_void_p, C.POINTER(Result)]),
    "b200j_gather": (C.c_int32, [C.c_void_p, C.POINTER(Result), C.POINTER(Batch), C.c_uint32, C.POINTER(Gathered)]),
    "b200j_gathered_release": (C.c_int32, [C.c_void_p, C.POINTER(Gathered)]),
    "b200j_probe_push": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(Batch)]),
    "b200j_staged_rows": (C.c_int64, [C.c_void_p, C.c_int32]),
    "b200j_probe_flush": (C.c_int32, [C.c_void_p, C.c_int32, C.c_uint32, C.POINTER(Result)]),
    "b200j_gather_staged": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(Result), C.c_uint32, C.POINTER(Gathered)]),
    "b200j_hash_keys": (C.c_int32, [C.c_void_p, C.POINTER(Batch), C.c_void_p]),
    "b200j_shuffle_count": (C.c_int32, [C.c_void_p, C.POINTER(Batch), C.c_int32, C.POINTER(C.c_int64)]),
    "b200j_resolve_origin": (C.c_int32, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_void_p),
                                         C.POINTER(C.c_uint32), C.c_void_p, C.c_void_p]),
    "b200j_gather_pieces": (C.c_int32, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                        C.c_int32, C.c_void_p, C.c_void_p]),
    "b200j_gather_pieces_utf8_offsets": (C.c_int32, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_void_p),
                                                     C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]),
    "b200j_gather_pieces_utf8_data": (C.c_int32, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_void_p),
                                                  C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p]),
    "b200j_utf8_rows_equal": (C.c_int32, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200j_shuffle_count_static": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p]),
    "b200j_shuffle_set_hot": (C.c_int32, [C.c_void_p, C.POINTER(C.c_uint64), C.c_int32]),
    "b200j_shuffle_scatter": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.c_uint32]),
    "b200j_shuffle_local_scatter": (C.c_int32, [C.c_void_p, C.POINTER(Batch), C.c_int32, C.c_void_p, C.c_void_p, C.c_int64, C.c_uint32, C.c_int32, C.POINTER(C.c_int64)]),
    "b200j_copy_async": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int32]),
    "b200j_copy_wait": (C.c_int32, [C.c_void_p]),
    "b200j_ipc_alloc": (C.c_int32, [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]),
    "b200j_ipc_open": (C.c_int32, [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p)]),
    "b200j_ipc_close": (C.c_int32, [C.c_void_p, C.c_void_p]),
    "b200j_ipc_free": (C.c_int32, [C.c_void_p, C.c_void_p]),
    "b200j_host_alloc": (C.c_int32, [C.c_size_t, C.POINTER(C.c_void_p)]),
    "b200j_host_free": (C.c_int32, [C.c_void_p]),
}

_lib = None


class NativeLibraryMissing(ImportError):
    pass


def lib() -> C.CDLL:
    """The loaded libb200join.so. Raises NativeLibraryMissing if it has not been built — never falls back."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NativeLibraryMissing(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). datafusion_parallelism_b200 has no CPU or PyTorch fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib
