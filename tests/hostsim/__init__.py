"""TEST INFRASTRUCTURE: builds and loads tests/_hostsim/libhostsim.so (the device headers compiled with g++)."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(ROOT, "tests", "_hostsim")
LIB = os.path.join(OUT, "libhostsim.so")
SRC = os.path.join(HERE, "hostsim.cpp")
CSRC = os.path.join(ROOT, "cofhe_b200", "csrc")

_lib = None
EXTRA_SIGS = {
    "hs_coop_compose": [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p,
                        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p],
    "hs_coop_pow": [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p,
                    ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p],
    "hs_codec_make": [ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p],
    "hs_codec_get": [ctypes.c_int, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p],
    "hs_fast_compose": [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p,
                        ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p],
    "hs_scheme": [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                  ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int,
                  ctypes.c_void_p],
    "hs_coop_tabprod": [ctypes.c_int] * 5 + [ctypes.c_uint] * 3 + [ctypes.c_int] * 3 + [ctypes.c_uint] * 2 + [ctypes.c_void_p, ctypes.c_void_p,
                        ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p],
    "hs_lift": [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p],
    "hs_fast_pow": [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p,
                    ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p],
}


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [SRC, os.path.join(HERE, "hostsim_extra.inc"), os.path.join(HERE, "coop_sim.inc")] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh")]
    return any(os.path.getmtime(d) > t for d in deps)


def load():
    global _lib
    if _lib is None:
        if _stale():
            os.makedirs(OUT, exist_ok=True)
            subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-Wall", "-Wno-unused-function",
                                   "-I" + CSRC, "-I" + HERE, "-o", LIB, SRC])
        _lib = ctypes.CDLL(LIB)
        vp, i, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
        sigs = {
            "hs_compose": [i, vp, i, i, i, sz, vp, vp, vp, i],
            "hs_square": [i, vp, i, i, i, sz, vp, vp],
            "hs_reduce": [i, vp, i, i, i, i, i, sz, vp, vp],
            "hs_pow": [i, vp, i, i, i, sz, vp, vp, vp, i, i, vp, i],
        }
        sigs.update(EXTRA_SIGS)
        for name, args in sigs.items():
            fn = getattr(_lib, name)
            fn.restype = i
            fn.argtypes = args
    return _lib


def P(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def cap_for(nd):
    need = 2 * nd + 4
    for c in (32, 48, 136, 216):
        if c >= need:
            return c
    raise ValueError("discriminant too large")
