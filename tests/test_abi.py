"""CPU: the C-ABI library loads and exports every symbol include/ydorb.h declares (no compute calls)."""
import ctypes
import pathlib
import re

ROOT = pathlib.Path(__file__).resolve().parents[1]


def declared_symbols():
    txt = (ROOT / "include" / "ydorb.h").read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(ydorb_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_header_symbols():
    from ydorbslam_b200 import build

    so = build.build()
    lib = ctypes.CDLL(str(so))
    syms = declared_symbols()
    assert len(syms) >= 25
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_error_reporting_without_device():
    from ydorbslam_b200._lib import lib

    l = lib()
    n = l.ydorb_device_count()
    assert n >= 0
    if n == 0:
        # no CPU fallback: creating a handle must fail loudly
        h = ctypes.c_void_p()
        rc = l.ydorb_matcher_create(0, ctypes.byref(h))
        assert rc != 0 and len(l.ydorb_last_error()) > 0
