"""ctypes binding of the C ABI (include/cofhe_b200.h).  Loading fails loudly when the CUDA library has not been
built; there is no CPU fallback anywhere in this package."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libcofhe_b200.so")

OK, EINVAL, ECUDA, EARITH, ENOMEM, ESTATE, ETOOBIG = 0, 1, 2, 3, 4, 5, 6
STATIC_MAX_N = 262144  # COFHE_B200_STATIC_MAX_N
_NAMES = {1: "EINVAL", 2: "ECUDA", 3: "EARITH", 4: "ENOMEM", 5: "ESTATE", 6: "ETOOBIG"}


class CofheError(RuntimeError):
    def __init__(self, code, what):
        super().__init__("cofhe_b200: %s failed with %s" % (what, _NAMES.get(code, code)))
        self.code = code


_lib = None
vp, u32, i32, sz, u64 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int, ctypes.c_size_t, ctypes.c_uint64

SIGNATURES = {
    "cofhe_b200_version": (ctypes.c_char_p, []),
    "cofhe_b200_create": (i32, [ctypes.POINTER(vp), i32, vp, u32]),
    "cofhe_b200_destroy": (None, [vp]),
    "cofhe_b200_shape": (i32, [vp, ctypes.POINTER(u32), ctypes.POINTER(u32), ctypes.POINTER(u32)]),
    "cofhe_b200_status": (i32, [vp, vp]),
    "cofhe_b200_launch_count": (u64, [vp]),
    "cofhe_b200_fast_limbs": (ctypes.c_int, [vp]),
    "cofhe_b200_coop_limbs": (ctypes.c_int, [vp]),
    "cofhe_b200_reg_limbs": (ctypes.c_int, [vp]),
    "cofhe_b200_set_option": (i32, [vp, ctypes.c_char_p, ctypes.c_long]),
    "cofhe_b200_scal_ciphertexts": (i32, [vp, sz, vp, vp, u32, i32, vp, i32, vp, vp]),
    "cofhe_b200_beaver_triplets": (i32, [vp, sz, vp, vp, u32, vp, u32, vp, vp]),
    "cofhe_b200_ct_matmul_expand": (i32, [vp, sz, sz, sz, vp, vp, vp, vp, vp]),
    "cofhe_b200_ct_matmul_reduce": (i32, [vp, sz, sz, sz, vp, vp, vp, vp]),
    "cofhe_b200_make_plaintexts": (i32, [vp, sz, vp, vp, u32, vp, vp]),
    "cofhe_b200_plaintexts_to_floats": (i32, [vp, sz, vp, u32, vp, vp, vp]),
    "cofhe_b200_copy": (i32, [vp, vp, vp, sz, vp]),
    "cofhe_b200_plaintexts_mod2k": (i32, [vp, sz, vp, u32, vp, vp, u32, vp]),
    "cofhe_b200_rerandomize": (i32, [vp, sz, vp, vp, u32, i32, vp, vp]),
    "cofhe_b200_beaver_open": (i32, [vp, sz, vp, vp, vp, vp, vp, vp, vp]),
    "cofhe_b200_beaver_close": (i32, [vp, sz, vp, vp, u32, vp, vp, vp, vp, u32, vp, vp]),
    "cofhe_b200_wire_bound_kind": (sz, [vp, i32, u32, sz, u32]),
    "cofhe_b200_wire_to_batch": (i32, [vp, i32, vp, sz, u32, sz, vp, u32, vp, vp]),
    "cofhe_b200_batch_to_wire": (i32, [vp, i32, u32, vp, sz, vp, u32, vp, vp, sz, ctypes.POINTER(sz), vp]),
    "cofhe_b200_wire_header": (i32, [vp, sz, ctypes.POINTER(u32), vp, u32, ctypes.POINTER(sz)]),
    "cofhe_b200_wire_bound": (sz, [vp, u32, sz]),
    "cofhe_b200_wire_to_cts": (i32, [vp, vp, sz, u32, sz, vp, vp]),
    "cofhe_b200_cts_to_wire": (i32, [vp, u32, vp, sz, vp, vp, sz, ctypes.POINTER(sz), vp]),
    "cofhe_b200_nucomp": (i32, [vp, sz, vp, vp, vp, i32, vp]),
    "cofhe_b200_nudupl": (i32, [vp, sz, vp, vp, vp]),
    "cofhe_b200_reduce": (i32, [vp, sz, vp, u32, u32, vp, vp]),
    "cofhe_b200_nupow": (i32, [vp, sz, vp, vp, u32, i32, vp, i32, vp, vp]),
    "cofhe_b200_add_ciphertexts": (i32, [vp, sz, vp, vp, vp, vp]),
    "cofhe_b200_add_ciphertexts_host": (i32, [vp, sz, vp, vp, vp]),
    "cofhe_b200_add_ciphertexts_static": (i32, [vp, sz, vp, vp, vp, vp]),
    "cofhe_b200_scal_matmul": (i32, [vp, sz, sz, sz, vp, vp, u32, u32, vp, vp, vp]),
    "cofhe_b200_set_scheme": (i32, [vp, u32, vp, u32]),
    "cofhe_b200_set_compact": (i32, [vp, i32]),
    "cofhe_b200_lift": (i32, [vp, sz, vp, vp, vp]),
    "cofhe_b200_set_bases": (i32, [vp, vp, vp, u32]),
    "cofhe_b200_fixed_base_pow": (i32, [vp, sz, vp, u32, vp, vp]),
    "cofhe_b200_power_of_f": (i32, [vp, sz, vp, u32, vp, vp]),
    "cofhe_b200_dlog": (i32, [vp, sz, vp, vp, u32, vp]),
    "cofhe_b200_encrypt": (i32, [vp, sz, vp, u32, vp, u32, i32, vp, vp]),
    "cofhe_b200_part_decrypt": (i32, [vp, sz, vp, vp, u32, i32, vp, vp]),
    "cofhe_b200_combine": (i32, [vp, sz, u32, vp, vp, vp, u32, vp]),
    "cofhe_b200_decrypt": (i32, [vp, sz, vp, vp, u32, vp, u32, vp]),
    "cofhe_b200_device_count": (i32, []),
    "cofhe_b200_dev_alloc": (i32, [vp, sz, ctypes.POINTER(vp)]),
    "cofhe_b200_dev_free": (None, [vp, vp]),
    "cofhe_b200_host_alloc": (i32, [sz, ctypes.POINTER(vp)]),
    "cofhe_b200_host_free": (None, [vp]),
    "cofhe_b200_stream": (vp, [vp]),
    "cofhe_b200_upload": (i32, [vp, vp, vp, sz, vp]),
    "cofhe_b200_download": (i32, [vp, vp, vp, sz, vp]),
    "cofhe_b200_imad_probe": (i32, [i32, i32, ctypes.POINTER(ctypes.c_double)]),
}


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("cofhe_b200: %s is missing -- build it with `python -m cofhe_b200.build` "
                              "(nvcc, sm_100a); there is no CPU fallback" % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc, what):
    if rc != OK:
        raise CofheError(rc, what)
