"""CPU: the C-ABI libraries load and export every symbol include/*.h declares (no compute calls)."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
LIB = ROOT / "zyk-llama.cpp_b200" / "lib"


def _declared(header: str, marker: str):
    text = (ROOT / "include" / header).read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"#define[^\n]*", "", text)
    return sorted(set(re.findall(marker + r"[^;(]*?\b(\w+)\s*\(", text)))


def test_libb200mm_exports_every_declared_symbol():
    names = _declared("b200_mm.h", "B200_API")
    assert len(names) >= 40
    lib = ctypes.CDLL(str(LIB / "libb200mm.so"))
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/b200_mm.h but not exported: {missing}"


def test_format_tables_match_ggml_block_sizes():
    lib = ctypes.CDLL(str(LIB / "libb200mm.so"))
    lib.b200_row_size.restype = ctypes.c_size_t
    lib.b200_row_size.argtypes = [ctypes.c_int, ctypes.c_int64]
    # ggml-common.h:170-431 -- bytes per 256 elements
    want = {2: 18 * 8, 8: 34 * 8, 9: 36 * 8, 12: 144, 14: 210, 15: 292, 23: 136}
    for t, nbytes in want.items():
        assert lib.b200_row_size(t, 256) == nbytes
    assert [lib.b200_act_type(t) for t in (2, 8, 12, 14, 23)] == [8, 8, 15, 15, 15]  # vec_dot_type, ggml-cpu.c:217-370


def test_row_split_covers_all_rows_in_tile_multiples():
    lib = ctypes.CDLL(str(LIB / "libb200mm.so"))
    lo, hi = ctypes.c_int64(), ctypes.c_int64()
    for m in (1024, 4096, 14336, 128256, 100):
        for world in (1, 2, 4, 8):
            prev = 0
            for r in range(world):
                assert lib.b200_row_split(ctypes.c_int64(m), world, r, ctypes.byref(lo), ctypes.byref(hi)) == 0
                assert lo.value == prev and hi.value >= lo.value
                if r < world - 1:
                    assert hi.value % 128 == 0
                prev = hi.value
            assert prev == m


def test_no_device_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = ctypes.CDLL(str(LIB / "libb200mm.so"))
    lib.b200_malloc.restype = ctypes.c_void_p
    assert lib.b200_device_count() == 0
    assert lib.b200_malloc(0, ctypes.c_size_t(1024)) is None
    lib.b200_last_error.restype = ctypes.c_char_p
    assert lib.b200_last_error()


def test_plugin_exports_ggml_entry_points():
    so = LIB / "libggml-b200.so"
    if not so.exists():
        pytest.skip("plugin not built")
    names = _declared("ggml_b200.h", "GGML_B200_API")
    out = __import__("subprocess").run(["nm", "-D", "--defined-only", str(so)], capture_output=True, text=True).stdout
    exported = {line.split()[-1] for line in out.splitlines() if line.strip()}
    assert "ggml_backend_init" in names and "ggml_backend_score" in names
    missing = [n for n in names if n not in exported]
    assert not missing, f"declared in include/ggml_b200.h but not exported: {missing}"
