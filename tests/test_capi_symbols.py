"""The C-ABI library loads and exports every symbol include/cofhe_b200.h declares (no compute without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared():
    txt = open(os.path.join(ROOT, "include", "cofhe_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(cofhe_b200_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported():
    from cofhe_b200 import build, capi
    build.build()
    lib = ctypes.CDLL(capi.LIB_PATH)
    names = declared()
    assert len(names) >= 10
    for n in names:
        assert hasattr(lib, n), "missing export " + n
    # and the Python binding covers the same set
    assert sorted(capi.SIGNATURES) == names
    assert b"sm_100a" in capi.load().cofhe_b200_version()


def test_no_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from cofhe_b200 import capi
    lib = capi.load()
    h = ctypes.c_void_p()
    d = (ctypes.c_uint32 * 2)(7, 1)
    assert lib.cofhe_b200_create(ctypes.byref(h), 0, d, 2) == capi.ECUDA
    from cofhe_b200.group import ClassGroupB200
    with pytest.raises(RuntimeError):
        ClassGroupB200(-23)
