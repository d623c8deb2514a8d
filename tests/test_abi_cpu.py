"""CPU: libexg.so loads and exports every symbol include/exg.h declares (no compute calls)."""
import ctypes
import os
import re

from exoground_b200.ops import lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "exg.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(exg_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert _header_symbols() == lib.exported_symbols()


def test_library_exports_all_symbols():
    assert os.path.exists(lib.LIB_PATH), "build with: python -m exoground_b200.build"
    so = ctypes.CDLL(lib.LIB_PATH)
    for name in _header_symbols():
        assert hasattr(so, name), name
    assert lib.load().exg_version() >= 100


def test_argument_errors_are_reported_without_a_gpu():
    so = lib.load()
    rc = so.exg_cast(None, 0, None, 0, 16, None)
    assert rc == -1 and b"cast" in so.exg_last_error()
