"""ORACLE / TEST INFRASTRUCTURE ONLY -- ctypes loader for oracle/_ref/libcofhe_ref.so.

The library is the UNMODIFIED reference (CoFHE CPUCryptoSystem + BICYCL + system libgmp) behind the tiny
C ABI of oracle/ref_capi.cpp.  Only tests/, tools/gen_golden.py, __graft_entry__.build()/smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libcofhe_ref.so")


def available():
    if not os.path.exists(LIB_PATH):
        return False
    try:
        _load()
        return True
    except OSError:
        return False


_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(LIB_PATH)
        lib.ref_new.restype = ctypes.c_void_p
        lib.ref_new.argtypes = [ctypes.c_uint, ctypes.c_uint]
        lib.ref_free.argtypes = [ctypes.c_void_p]
        lib.ref_call.restype = ctypes.c_void_p
        lib.ref_call.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_char_p]
        lib.ref_free_str.argtypes = [ctypes.c_void_p]
        lib.ref_num_threads.restype = ctypes.c_int
        lib.ref_set_threads.argtypes = [ctypes.c_int]
        u32p = ctypes.c_void_p
        lib.ref_bulk_matadd.restype = ctypes.c_double
        lib.ref_bulk_matadd.argtypes = [ctypes.c_void_p, ctypes.c_uint, ctypes.c_uint, ctypes.c_uint,
                                        u32p, u32p, u32p, ctypes.c_uint, ctypes.c_uint]
        lib.ref_bulk_scal_matmul.restype = ctypes.c_double
        lib.ref_bulk_scal_matmul.argtypes = [ctypes.c_void_p, ctypes.c_uint, ctypes.c_uint, ctypes.c_uint,
                                             ctypes.c_uint, ctypes.c_char_p, u32p, u32p, ctypes.c_uint,
                                             ctypes.c_uint]
        lib.ref_bulk_encrypt.restype = ctypes.c_double
        lib.ref_bulk_encrypt.argtypes = [ctypes.c_void_p, ctypes.c_uint, u32p, ctypes.c_uint, u32p,
                                         ctypes.c_uint, u32p, ctypes.c_uint, ctypes.c_uint]
        lib.ref_bulk_part_decrypt.restype = ctypes.c_double
        lib.ref_bulk_part_decrypt.argtypes = [ctypes.c_void_p, ctypes.c_uint, ctypes.c_char_p, u32p, u32p,
                                              ctypes.c_uint, ctypes.c_uint]
        _lib = lib
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


class Reference:
    """One reference CPUCryptoSystem(nbits, k) (default-seeded RandGen, deterministic) plus a key pair."""

    def __init__(self, nbits=128, k=128, compact=False):
        self.lib = _load()
        if compact:
            self.lib.ref_new_compact.restype = ctypes.c_void_p
            self.lib.ref_new_compact.argtypes = [ctypes.c_uint, ctypes.c_uint, ctypes.c_int]
            self.h = self.lib.ref_new_compact(nbits, k, 1)
        else:
            self.h = self.lib.ref_new(nbits, k)
        if not self.h:
            raise RuntimeError("reference constructor failed")
        self.nbits, self.k = nbits, k

    def close(self):
        if self.h:
            self.lib.ref_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def call(self, op, *args):
        flat = []
        for a in args:
            if isinstance(a, (list, tuple)):
                flat.extend(_flatten(a))
            else:
                flat.append(a)
        s = " ".join(str(int(x)) for x in flat).encode()
        p = self.lib.ref_call(self.h, op.encode(), s)
        out = ctypes.string_at(p).decode()
        self.lib.ref_free_str(p)
        if not out.startswith("OK"):
            raise RuntimeError("reference %s: %s" % (op, out))
        return [int(t) for t in out[2:].split()]

    def params(self):
        v = self.call("params")
        names = ["N", "DeltaK", "Delta", "k", "exponent_bound", "d", "e"]
        p = dict(zip(names, v[:7]))
        p["h"] = tuple(v[7:10]); p["h_e"] = tuple(v[10:13]); p["h_d"] = tuple(v[13:16]); p["h_de"] = tuple(v[16:19])
        p["L"] = v[19]; p["large_message"] = bool(v[20]); p["secretkey_bound"] = v[21]; p["cleartext_bound"] = v[22]
        return p

    def threads(self):
        return self.lib.ref_num_threads()

    def set_threads(self, n):
        self.lib.ref_set_threads(int(n))

    # ---- bulk workloads (see ref_capi.cpp) -------------------------------------------------------
    def bulk_matadd(self, rows, cols, iters, na=0, nc=0, export=False):
        n = rows * cols
        bufs = [np.zeros((2, 2 * na + nc, n), dtype=np.uint32) for _ in range(3)] if export else [None] * 3
        dt = self.lib.ref_bulk_matadd(self.h, rows, cols, iters, _ptr(bufs[0]), _ptr(bufs[1]), _ptr(bufs[2]), na, nc)
        if dt < 0:
            raise RuntimeError("ref_bulk_matadd failed")
        return dt, bufs

    def bulk_wire(self, rows, cols, iters, na, nc):
        """-> (serialise seconds, deserialise seconds, ciphertext SoA limbs, wire bytes) for `iters` passes each"""
        n = rows * cols
        cts = np.zeros((2, 2 * na + nc, n), dtype=np.uint32)
        cap = 64 + n * (48 + 4 * (4 * na + 1) + 2 * (4 * nc + 1))
        wire = np.zeros(cap, dtype=np.uint8)
        wl, dd = ctypes.c_size_t(), ctypes.c_double()
        self.lib.ref_bulk_wire.restype = ctypes.c_double
        self.lib.ref_bulk_wire.argtypes = [ctypes.c_void_p, ctypes.c_uint, ctypes.c_uint, ctypes.c_uint, ctypes.c_void_p, ctypes.c_uint,
                                           ctypes.c_uint, ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t),
                                           ctypes.POINTER(ctypes.c_double)]
        ds = self.lib.ref_bulk_wire(self.h, rows, cols, iters, _ptr(cts), na, nc, _ptr(wire), cap, ctypes.byref(wl), ctypes.byref(dd))
        if ds < 0:
            raise RuntimeError("ref_bulk_wire failed")
        return ds, dd.value, cts, wire[:wl.value].tobytes()

    def bulk_scal_matmul(self, n, m, p, iters=1, scalars=None, na=0, nc=0, export=False):
        bi = np.zeros((2, 2 * na + nc, n * m), dtype=np.uint32) if export else None
        bo = np.zeros((2, 2 * na + nc, n * p), dtype=np.uint32) if export else None
        s = None if scalars is None else " ".join(str(int(x)) for x in scalars).encode()
        dt = self.lib.ref_bulk_scal_matmul(self.h, n, m, p, iters, s, _ptr(bi), _ptr(bo), na, nc)
        if dt < 0:
            raise RuntimeError("ref_bulk_scal_matmul failed")
        return dt, bi, bo

    def bulk_encrypt(self, count, nm, nr, na, nc):
        mo = np.zeros((nm, count), dtype=np.uint32)
        ro = np.zeros((nr, count), dtype=np.uint32)
        co = np.zeros((2, 2 * na + nc, count), dtype=np.uint32)
        dt = self.lib.ref_bulk_encrypt(self.h, count, _ptr(mo), nm, _ptr(ro), nr, _ptr(co), na, nc)
        if dt < 0:
            raise RuntimeError("ref_bulk_encrypt failed")
        return dt, mo, ro, co

    def bulk_part_decrypt(self, count, share, na, nc):
        co = np.zeros((2, 2 * na + nc, count), dtype=np.uint32)
        do = np.zeros((2 * na + nc, count), dtype=np.uint32)
        dt = self.lib.ref_bulk_part_decrypt(self.h, count, str(int(share)).encode(), _ptr(co), _ptr(do), na, nc)
        if dt < 0:
            raise RuntimeError("ref_bulk_part_decrypt failed")
        return dt, co, do

    # ---- the same entry points on caller-supplied SoA limb buffers (full-size parity) -----------------
    def _fn(self, name, argtypes):
        fn = getattr(self.lib, name)
        fn.restype = ctypes.c_double
        fn.argtypes = argtypes
        return fn

    def scal_matmul_given(self, n, m, p, scalars, ct, na, nc, exact_zero=None):
        """reference scal_ciphertext_tensors (exact_zero None) or the exact product built from ClassGroup::nupow /
        nucomp with the given `zero` ciphertext limbs; ct: uint32 [2, fl, n*m] -> (seconds, uint32 [2, fl, n*p])."""
        u, vp, cp = ctypes.c_uint, ctypes.c_void_p, ctypes.c_char_p
        out = np.zeros((2, 2 * na + nc, n * p), dtype=np.uint32)
        s = " ".join(str(int(x)) for x in scalars).encode()
        ct = np.ascontiguousarray(ct, dtype=np.uint32)
        if exact_zero is None:
            dt = self._fn("ref_bulk_scal_matmul_given", [vp, u, u, u, cp, vp, vp, u, u])(self.h, n, m, p, s, _ptr(ct), _ptr(out), na, nc)
        else:
            z = np.ascontiguousarray(exact_zero, dtype=np.uint32)
            dt = self._fn("ref_bulk_matmul_exact", [vp, u, u, u, cp, vp, vp, vp, u, u])(self.h, n, m, p, s, _ptr(ct), _ptr(z), _ptr(out), na, nc)
        if dt < 0:
            raise RuntimeError("reference matmul failed")
        return dt, out

    def encrypt_given(self, m, r, na, nc):
        """m: uint32 [nm, count], r: uint32 [nr, count] -> (seconds, cts [2, fl, count]) by `count` single encryptions"""
        u, vp = ctypes.c_uint, ctypes.c_void_p
        m, r = np.ascontiguousarray(m, dtype=np.uint32), np.ascontiguousarray(r, dtype=np.uint32)
        count = m.shape[1]
        out = np.zeros((2, 2 * na + nc, count), dtype=np.uint32)
        dt = self._fn("ref_bulk_encrypt_given", [vp, u, vp, u, vp, u, vp, u, u])(self.h, count, _ptr(m), m.shape[0], _ptr(r), r.shape[0], _ptr(out), na, nc)
        if dt < 0:
            raise RuntimeError("ref_bulk_encrypt_given failed")
        return dt, out

    def part_decrypt_given(self, share, ct, na, nc):
        u, vp = ctypes.c_uint, ctypes.c_void_p
        ct = np.ascontiguousarray(ct, dtype=np.uint32)
        count = ct.shape[-1]
        out = np.zeros((2 * na + nc, count), dtype=np.uint32)
        dt = self._fn("ref_bulk_part_decrypt_given", [vp, u, ctypes.c_char_p, vp, vp, u, u])(self.h, count, str(int(share)).encode(), _ptr(ct), _ptr(out), na, nc)
        if dt < 0:
            raise RuntimeError("ref_bulk_part_decrypt_given failed")
        return dt, out

    def combine_given(self, ct, ds, nm, na, nc):
        """ct [2, fl, count], ds [t, fl, count] -> (seconds, plaintexts [nm, count])"""
        u, vp = ctypes.c_uint, ctypes.c_void_p
        ct, ds = np.ascontiguousarray(ct, dtype=np.uint32), np.ascontiguousarray(ds, dtype=np.uint32)
        count, t = ct.shape[-1], ds.shape[0]
        out = np.zeros((nm, count), dtype=np.uint32)
        dt = self._fn("ref_bulk_combine_given", [vp, u, u, vp, vp, vp, u, u, u])(self.h, count, t, _ptr(ct), _ptr(ds), _ptr(out), nm, na, nc)
        if dt < 0:
            raise RuntimeError("ref_bulk_combine_given failed")
        return dt, out


def _flatten(a):
    for x in a:
        if isinstance(x, (list, tuple)):
            yield from _flatten(x)
        else:
            yield x
