"""The C-ABI library loads without a GPU and exports exactly what include/event3dgs.h declares."""
import ctypes
import os
import re

from event_3dgs_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = []
    for fn in sorted(os.listdir(os.path.join(ROOT, "include"))):
        if fn.endswith(".h"):
            src = open(os.path.join(ROOT, "include", fn)).read()
            names += re.findall(r"E3_API\s+[\w\s\*]+?\b(e3_\w+)\s*\(", src)
    return names


def test_header_declares_entry_points():
    names = _declared()
    assert "e3_rasterize_forward" in names and "e3_rasterize_backward" in names and "e3_mark_visible" in names
    assert len(names) == len(set(names))


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "build libevent3dgs.so first (__graft_entry__.build())"
    L = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared():
        assert hasattr(L, name), name


def test_binding_table_matches_header():
    assert set(_lib.SIGNATURES) == set(_declared())


def test_host_only_calls():
    L = _lib.lib()
    assert L.e3_version() >= 100
    assert L.e3_error_string(0) == b"success"
    assert b"precomputed" in L.e3_error_string(-4)
    # state sizes grow with their argument and cover at least the visible arrays
    assert L.e3_geometry_bytes(1000) >= 1000 * (4 + 48 + 24)
    assert L.e3_geometry_bytes(2000) > L.e3_geometry_bytes(1000)
    assert L.e3_image_bytes(64, 64) >= 64 * 64 * 8 + 16 * 8
    assert L.e3_binning_bytes(10_000) >= 10_000 * 16
    assert L.e3_binning_bytes(0) >= 0
    # argument validation happens before any CUDA call
    assert L.e3_mark_visible(-1, None, None, None, None, None) == -1
    assert L.e3_mark_visible(0, None, None, None, None, None) == 0
