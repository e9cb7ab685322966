"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/ggml_b200_ops.h
declares, and fails loudly (no fallback) when no CUDA device is present.  No compute calls here."""
import ctypes as C
import os
import re
import subprocess

import pytest

import b200ops as B

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _built():
    if not os.path.exists(B.LIB_PATH):
        subprocess.check_call(["make", "-s", "-f", os.path.join(ROOT, "ggml-hexagon_b200", "Makefile")])
    return B.LIB_PATH


def test_header_symbols_match_binding_list():
    hdr = open(os.path.join(ROOT, "include", "ggml_b200_ops.h")).read()
    declared = set(re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", hdr))
    assert declared == set(B.SYMBOLS), declared ^ set(B.SYMBOLS)


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(_built())
    for s in B.SYMBOLS:
        assert hasattr(lib, s), s
    out = subprocess.run(["nm", "-D", "--defined-only", _built()], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (\w+)", out))
    assert set(B.SYMBOLS) <= exported
    # nothing but the C-ABI leaks out of the library
    assert all(n.startswith("b200_") for n in exported if not n.startswith("_")), exported


def test_block_geometry_matches_ggml_common():
    _built()
    L = B.lib()
    # ggml/src/ggml-common.h:167-334
    for t, (el, by) in B.BLOCK.items():
        assert L.b200_block_elems(t) == el and L.b200_block_bytes(t) == by
    assert L.b200_block_elems(999) == 0


def test_struct_layout_matches_header():
    # b200_tensor: 4 i32 + 4 i64 + 4 i64 + 16 i32 + ptr + i64 + ptr + 2 i64
    assert C.sizeof(B.Tensor) == 16 + 32 + 32 + 64 + 8 + 8 + 8 + 16
    assert C.sizeof(B.Node) == 4 * C.sizeof(B.Tensor)        # src0, src1, dst, src2


def test_no_cpu_fallback_without_gpu():
    _built()
    if B.device_count() > 0:
        pytest.skip("a CUDA device is visible")
    with pytest.raises(B.B200Error) as e:
        B.Session(0)
    assert "no CUDA device" in str(e.value)
    # a bad handle is rejected, not dereferenced
    assert B.lib().b200_synchronize(C.c_uint64(0)) == -2


def test_product_sources_never_touch_the_oracle():
    pkg = os.path.join(ROOT, "ggml-hexagon_b200")
    for dp, _, fns in os.walk(pkg):
        if os.path.basename(dp) == "lib":
            continue
        for fn in fns:
            if fn.endswith((".cu", ".cuh", ".cpp", ".h", ".py")) or fn == "Makefile":
                src = open(os.path.join(dp, fn)).read()
                assert "pyoracle" not in src and "liboracle" not in src and "ggml_oracle" not in src, fn
