"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/b200join.h declares, the Python veneer's signatures cover the header, the product never touches
oracle/, and without a GPU the product fails loudly instead of falling back."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200join.h")
PKG = os.path.join(ROOT, "datafusion_parallelism_b200")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200j_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def native():
    import __graft_entry__ as g
    g.build()
    from datafusion_parallelism_b200 import _native
    return _native


def test_header_declares_something():
    syms = _declared_symbols()
    assert "b200j_probe_batch" in syms and "b200j_build_push" in syms and len(syms) >= 15


def test_library_exports_every_declared_symbol(native):
    L = C.CDLL(native.LIB_PATH)
    for s in _declared_symbols():
        assert hasattr(L, s), f"{s} declared in include/b200join.h but not exported by libb200join.so"
    out = subprocess.run(["nm", "-D", "--defined-only", native.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\sT\s+(b200j_[a-z0-9_]+)", out))
    assert exported == set(_declared_symbols())


def test_python_signatures_cover_the_header(native):
    assert sorted(native.SIGNATURES) == _declared_symbols()
    assert native.lib().b200j_abi_version() == 1


def test_struct_sizes_match_c_layout(native):
    """sizeof of every ctypes mirror equals the C compiler's sizeof for the header's structs."""
    prog = r'''
    #include "b200join.h"
    #include <stdio.h>
    int main(void){ printf("%zu %zu %zu %zu %zu %zu %zu\n", sizeof(b200j_column), sizeof(b200j_batch), sizeof(b200j_config),
        sizeof(b200j_result), sizeof(b200j_out_column), sizeof(b200j_gathered), sizeof(b200j_stats)); return 0; }'''
    d = os.path.join(ROOT, "oracle", "_build")
    os.makedirs(d, exist_ok=True)
    src, exe = os.path.join(d, "sizes.c"), os.path.join(d, "sizes")
    open(src, "w").write(prog)
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", exe, src], check=True)
    sizes = [int(x) for x in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    mirrors = [native.Column, native.Batch, native.Config, native.Result, native.OutColumn, native.Gathered, native.Stats]
    assert sizes == [C.sizeof(m) for m in mirrors]


def test_field_offsets_and_flag_values_match_the_header(native):
    """offsetof of every field of the result and stats structs (the two that grew this round) equals the ctypes mirror's, and the
    `want` flags of the Python veneer are the header's values."""
    import datafusion_parallelism_b200.join as J
    fields = {"b200j_result": [n for n, _ in native.Result._fields_], "b200j_stats": [n for n, _ in native.Stats._fields_]}
    flags = ["PAIRS", "HOST", "MATCHES_ONLY", "COUNT_ONLY", "BUILD_U32", "GATHER_HINT", "PROBE_ORDER"]
    body = "".join(f'printf("%zu ", offsetof({st}, {f}));' for st, fs in fields.items() for f in fs)
    body += "".join(f'printf("%u ", (unsigned)B200J_OUT_{f});' for f in flags)
    prog = '#include "b200join.h"\n#include <stdio.h>\n#include <stddef.h>\nint main(void){' + body + 'return 0;}'
    d = os.path.join(ROOT, "oracle", "_build")
    os.makedirs(d, exist_ok=True)
    src, exe = os.path.join(d, "offsets.c"), os.path.join(d, "offsets")
    open(src, "w").write(prog)
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", exe, src], check=True)
    got = [int(x) for x in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    want = [getattr(m, f).offset for m, st in ((native.Result, "b200j_result"), (native.Stats, "b200j_stats")) for f in fields[st]]
    want += [getattr(J, "OUT_" + f) for f in flags]
    assert got == want


def test_product_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under the package may import, link or open it."""
    bad = []
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                if re.search(r"^\s*(import|from)\s+oracle\b", txt, flags=re.M) or "liborc" in txt or "oracle/_" in txt:
                    bad.append(os.path.join(dirpath, f))
    assert not bad, bad


def test_no_gpu_means_loud_failure_not_fallback(native):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import datafusion_parallelism_b200 as d
    assert d.device_count() == 0
    with pytest.raises(d.B200JoinError) as e:
        d.HashJoin("inner")
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_create_rejects_bad_config(native):
    L = native.lib()
    cfg = native.Config()
    h = C.c_void_p()
    cfg.struct_size = 4            # ABI mismatch
    assert L.b200j_create(C.byref(cfg), C.byref(h)) == -1
    assert b"struct_size" in L.b200j_last_error(None)
    assert L.b200j_create(None, C.byref(h)) == -1
    assert L.b200j_destroy(None) == 0
