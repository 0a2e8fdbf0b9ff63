"""Host-side handle on a class group living on one B200: thin Python mirror of the reference's
BICYCL::ClassGroup batched over SoA limb tensors (torch is used only for device memory and streams)."""
import ctypes

import numpy as np
import torch

from . import capi
from . import limbs as L


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def to_device(arr, device):
    """numpy uint32 -> torch int32 on device (bit pattern preserved)."""
    return torch.from_numpy(np.ascontiguousarray(arr).view(np.int32)).to(device, non_blocking=False)


def to_host(t):
    return t.detach().cpu().numpy().view(np.uint32)


class ClassGroupB200:
    """Class group of discriminant D < 0 on CUDA device `device`."""

    def __init__(self, D, device=0):
        if D >= 0:
            raise ValueError("discriminant must be negative")
        if not torch.cuda.is_available():
            raise RuntimeError("cofhe_b200 needs a CUDA device (no CPU fallback)")
        self.lib = capi.load()
        self.D = int(D)
        self.device = torch.device("cuda", device)
        nd = ((-D).bit_length() + 31) // 32
        dl = L.ints_to_limbs([-D], nd)[:, 0].copy()
        h = ctypes.c_void_p()
        capi.check(self.lib.cofhe_b200_create(ctypes.byref(h), device, dl.ctypes.data_as(ctypes.c_void_p), nd), "create")
        self.h = h
        na, nc, nd2 = ctypes.c_uint32(), ctypes.c_uint32(), ctypes.c_uint32()
        capi.check(self.lib.cofhe_b200_shape(self.h, ctypes.byref(na), ctypes.byref(nc), ctypes.byref(nd2)), "shape")
        self.na, self.nc, self.nd = na.value, nc.value, nd2.value
        self.fl = 2 * self.na + self.nc

    def close(self):
        if getattr(self, "h", None):
            self.lib.cofhe_b200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- marshalling -----------------------------------------------------------------------------
    def forms(self, fs):
        return to_device(L.forms_to_limbs(fs, self.na, self.nc), self.device)

    def cts(self, cs):
        return to_device(L.cts_to_limbs(cs, self.na, self.nc), self.device)

    def forms_out(self, t):
        return L.limbs_to_forms(to_host(t), self.na, self.nc)

    def cts_out(self, t):
        return L.limbs_to_cts(to_host(t), self.na, self.nc)

    def status(self):
        capi.check(self.lib.cofhe_b200_status(self.h, _stream()), "status")

    @property
    def fast(self):
        """limbs of the hand-written kernels (register-resident or cooperative) serving this discriminant (0: generic path)"""
        return int(self.lib.cofhe_b200_fast_limbs(self.h))

    @property
    def reg(self):
        """limbs of the register-resident (one thread per form) kernels for this discriminant (0: none)"""
        return int(self.lib.cofhe_b200_reg_limbs(self.h))

    def set_option(self, name, value):
        """tuning knob of the C ABI (cofhe_b200_set_option); -1 = automatic"""
        capi.check(self.lib.cofhe_b200_set_option(self.h, name.encode(), int(value)), "set_option")

    def launches(self):
        return int(self.lib.cofhe_b200_launch_count(self.h))

    # ---- batched form arithmetic (device tensors in, device tensors out) ------------------------------
    def nucomp(self, f1, f2, negf2=False, out=None):
        n = f1.shape[-1]
        out = torch.empty_like(f1) if out is None else out
        capi.check(self.lib.cofhe_b200_nucomp(self.h, n, _ptr(f1), _ptr(f2), _ptr(out), int(negf2), _stream()), "nucomp")
        return out

    def nudupl(self, f, out=None):
        n = f.shape[-1]
        out = torch.empty_like(f) if out is None else out
        capi.check(self.lib.cofhe_b200_nudupl(self.h, n, _ptr(f), _ptr(out), _stream()), "nudupl")
        return out

    def reduce(self, f, na_in, nc_in):
        n = f.shape[-1]
        out = torch.empty((self.fl, n), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_reduce(self.h, n, _ptr(f), na_in, nc_in, _ptr(out), _stream()), "reduce")
        return out

    def nupow(self, f, exps=None, shared=None):
        """exps: list of ints (per element) or shared: one int."""
        n = f.shape[-1]
        out = torch.empty_like(f)
        if shared is not None:
            el = max(1, (abs(shared).bit_length() + 31) // 32)
            ex = to_device(L.ints_to_limbs([abs(shared)], el)[:, 0].copy(), self.device)
            capi.check(self.lib.cofhe_b200_nupow(self.h, n, _ptr(f), _ptr(ex), el, 0, None, int(shared < 0), _ptr(out),
                                                 _stream()), "nupow")
        else:
            el = max(1, max((abs(e).bit_length() + 31) // 32 for e in exps))
            ex = to_device(L.ints_to_limbs([abs(e) for e in exps], el), self.device)
            sg = torch.tensor([1 if e < 0 else 0 for e in exps], dtype=torch.uint8, device=self.device)
            capi.check(self.lib.cofhe_b200_nupow(self.h, n, _ptr(f), _ptr(ex), el, 1, _ptr(sg), 0, _ptr(out), _stream()),
                       "nupow")
        return out

    def add_ciphertexts(self, ct1, ct2, out=None):
        n = ct1.shape[-1]
        out = torch.empty_like(ct1) if out is None else out
        capi.check(self.lib.cofhe_b200_add_ciphertexts(self.h, n, _ptr(ct1), _ptr(ct2), _ptr(out), _stream()), "add")
        return out

    def add_ciphertexts_static(self, ct1, ct2, out):
        """add_ciphertexts through the entry point that allocates nothing (scratch in a fixed array inside the library, at
        most capi.STATIC_MAX_N ciphertexts per call); `out` is the caller's, like the inputs."""
        n = ct1.shape[-1]
        capi.check(self.lib.cofhe_b200_add_ciphertexts_static(self.h, n, _ptr(ct1), _ptr(ct2), _ptr(out), _stream()), "add_static")
        return out

    def scal_matmul(self, ct, scalars, n, m, p, init, ebits=None):
        """out[i,k] = init o prod_j ct[i,j]^{s[j,k]}: ct device tensor [2, fl, n*m]; scalars: m*p non-negative ints
        (row-major) or a device tensor [el, m*p]; init: device tensor [2, fl, 1] (the `zero` ciphertext)."""
        if torch.is_tensor(scalars):
            sc, el = scalars, scalars.shape[0]
            eb = 0 if ebits is None else ebits
        else:
            if any(x < 0 for x in scalars):
                raise ValueError("matmul scalars must be non-negative")
            eb = max([int(x).bit_length() for x in scalars] + [1]) if ebits is None else ebits
            el = max(1, (max([int(x).bit_length() for x in scalars] + [1]) + 31) // 32)
            sc = to_device(L.ints_to_limbs(scalars, el), self.device)
        out = torch.empty((2, self.fl, n * p), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_scal_matmul(self.h, n, m, p, _ptr(ct), _ptr(sc), el, eb, _ptr(init), _ptr(out),
                                                   _stream()), "scal_matmul")
        return out

    # ---- binary tensor wire format (reference serialize/deserialize_ciphertext_tensor) on the device -------------------
    def wire_to_cts(self, data):
        """bytes of CPUCryptoSystem::serialize_ciphertext_tensor -> (shape, device ciphertext batch [2, fl, n])."""
        buf = np.frombuffer(bytes(data), dtype=np.uint8)
        ndim, n = ctypes.c_uint32(), ctypes.c_size_t()
        dims = (ctypes.c_uint32 * 8)()
        capi.check(self.lib.cofhe_b200_wire_header(buf.ctypes.data_as(ctypes.c_void_p), buf.size, ctypes.byref(ndim), dims, 8,
                                                   ctypes.byref(n)), "wire_header")
        dev = torch.from_numpy(buf.copy()).to(self.device)
        out = torch.empty((2, self.fl, n.value), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_wire_to_cts(self.h, _ptr(dev), buf.size, ndim.value, n.value, _ptr(out), _stream()), "wire_to_cts")
        torch.cuda.current_stream().synchronize()  # `dev` must outlive the kernel
        return [int(dims[i]) for i in range(ndim.value)], out

    def cts_to_wire(self, shape, ct):
        """device ciphertext batch -> bytes of the reference's binary tensor format."""
        n = ct.shape[-1]
        cap = int(self.lib.cofhe_b200_wire_bound(self.h, len(shape), n))
        dev = torch.empty(cap, dtype=torch.uint8, device=self.device)
        dims = (ctypes.c_uint32 * max(1, len(shape)))(*shape)
        written = ctypes.c_size_t()
        capi.check(self.lib.cofhe_b200_cts_to_wire(self.h, len(shape), dims, n, _ptr(ct), _ptr(dev), cap, ctypes.byref(written),
                                                   _stream()), "cts_to_wire")
        return bytes(dev[:written.value].cpu().numpy())

    def _wire_in(self, kind, data, ml=0, signs=False):
        buf = np.frombuffer(bytes(data), dtype=np.uint8)
        ndim, n = ctypes.c_uint32(), ctypes.c_size_t()
        dims = (ctypes.c_uint32 * 8)()
        capi.check(self.lib.cofhe_b200_wire_header(buf.ctypes.data_as(ctypes.c_void_p), buf.size, ctypes.byref(ndim), dims, 8,
                                                   ctypes.byref(n)), "wire_header")
        dev = torch.from_numpy(buf.copy()).to(self.device)
        rows = {0: (2, self.fl), 1: (self.fl,), 2: (ml,)}[kind]
        out = torch.empty(rows + (n.value,), dtype=torch.int32, device=self.device)
        sg = torch.zeros(max(1, n.value), dtype=torch.uint8, device=self.device) if signs else None
        capi.check(self.lib.cofhe_b200_wire_to_batch(self.h, kind, _ptr(dev), buf.size, ndim.value, n.value, _ptr(out), ml,
                                                     _ptr(sg) if sg is not None else None, _stream()), "wire_to_batch")
        torch.cuda.current_stream().synchronize()  # `dev` must outlive the kernel
        return [int(dims[i]) for i in range(ndim.value)], out, sg

    def _wire_out(self, kind, shape, batch, ml=0, signs=None):
        n = batch.shape[-1]
        cap = int(self.lib.cofhe_b200_wire_bound_kind(self.h, kind, len(shape), n, ml))
        dev = torch.empty(cap, dtype=torch.uint8, device=self.device)
        dims = (ctypes.c_uint32 * max(1, len(shape)))(*shape)
        written = ctypes.c_size_t()
        capi.check(self.lib.cofhe_b200_batch_to_wire(self.h, kind, len(shape), dims, n, _ptr(batch), ml,
                                                     _ptr(signs) if signs is not None else None, _ptr(dev), cap, ctypes.byref(written),
                                                     _stream()), "batch_to_wire")
        return bytes(dev[:written.value].cpu().numpy())

    def wire_to_forms(self, data):
        """bytes of serialize_part_decryption_result_tensor -> (shape, device form batch [fl, n])"""
        shape, out, _ = self._wire_in(1, data)
        return shape, out

    def forms_to_wire(self, shape, forms):
        return self._wire_out(1, shape, forms)

    def wire_to_plaintexts(self, data, ml):
        """bytes of serialize_plaintext_tensor -> (shape, magnitudes [ml, n], sign bytes [n]) on the device"""
        return self._wire_in(2, data, ml, signs=True)

    def plaintexts_to_wire(self, shape, mags, signs=None):
        return self._wire_out(2, shape, mags, mags.shape[0], signs)

    # ---- CL_HSM2k scheme level (mirrors CPUCryptoSystem's batched entry points on SoA device tensors) --------------
    def set_scheme(self, k, N):
        nN = (int(N).bit_length() + 31) // 32
        nl = L.ints_to_limbs([N], nN)[:, 0].copy()
        capi.check(self.lib.cofhe_b200_set_scheme(self.h, k, nl.ctypes.data_as(ctypes.c_void_p), nN), "set_scheme")
        self.k, self.ml = k, (k + 31) // 32

    def set_compact(self, on=True):
        """compact variant (CPUCryptoSystem(level, k, compact=true)): h, pk and the c1 halves live in Cl(DeltaK); call
        between set_scheme and set_bases"""
        capi.check(self.lib.cofhe_b200_set_compact(self.h, int(bool(on))), "set_compact")

    def lift(self, f):
        """forms of Cl(DeltaK) -> Cl(Delta) (CL_HSM2k::from_Cl_DeltaK_to_Cl_Delta), compact variant"""
        out = torch.empty_like(f)
        capi.check(self.lib.cofhe_b200_lift(self.h, f.shape[-1], _ptr(f), _ptr(out), _stream()), "lift")
        return out

    def set_bases(self, h, pk, ebits):
        hl = L.forms_to_limbs([h], self.na, self.nc)[:, 0].copy()
        pl = None if pk is None else L.forms_to_limbs([pk], self.na, self.nc)[:, 0].copy()
        vp = ctypes.c_void_p
        capi.check(self.lib.cofhe_b200_set_bases(self.h, hl.ctypes.data_as(vp), None if pl is None else pl.ctypes.data_as(vp),
                                                 ebits), "set_bases")
        self.el_r = (ebits + 31) // 32

    def ints(self, vals, nl):
        return to_device(L.ints_to_limbs(vals, nl), self.device)

    def fixed_base_pow(self, exps):
        """exps: list of ints or device tensor [el, n] -> ciphertext-shaped tensor (h^x, pk^x)."""
        ex = exps if torch.is_tensor(exps) else self.ints(exps, self.el_r)
        n = ex.shape[-1]
        out = torch.empty((2, self.fl, n), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_fixed_base_pow(self.h, n, _ptr(ex), ex.shape[0], _ptr(out), _stream()), "fixed_base_pow")
        return out

    def power_of_f(self, ms):
        m = ms if torch.is_tensor(ms) else self.ints([x % (1 << self.k) for x in ms], self.ml)
        n = m.shape[-1]
        out = torch.empty((self.fl, n), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_power_of_f(self.h, n, _ptr(m), m.shape[0], _ptr(out), _stream()), "power_of_f")
        return out

    def dlog(self, f):
        n = f.shape[-1]
        out = torch.empty((self.ml, n), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_dlog(self.h, n, _ptr(f), _ptr(out), self.ml, _stream()), "dlog")
        return out

    def encrypt(self, ms, r):
        """r: one int (shared randomness, reference encrypt_tensor semantics) or a list / device tensor (per element)."""
        m = ms if torch.is_tensor(ms) else self.ints([x % (1 << self.k) for x in ms], self.ml)
        n = m.shape[-1]
        shared = isinstance(r, int)
        rr = r if torch.is_tensor(r) else self.ints([r] if shared else r, self.el_r)
        out = torch.empty((2, self.fl, n), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_encrypt(self.h, n, _ptr(m), m.shape[0], _ptr(rr), rr.shape[0], int(shared), _ptr(out),
                                               _stream()), "encrypt")
        return out

    def scal_ciphertexts(self, ct, exps=None, shared=None):
        """ct[e]^x on both forms: exps = per-element ints / device tensor [el, n], or shared = one int (negate: 2^k - 1)."""
        n = ct.shape[-1]
        out = torch.empty_like(ct)
        if shared is not None:
            el = max(1, (abs(shared).bit_length() + 31) // 32)
            ex = self.ints([abs(shared)], el)
            capi.check(self.lib.cofhe_b200_scal_ciphertexts(self.h, n, _ptr(ct), _ptr(ex), el, 0, None, int(shared < 0), _ptr(out),
                                                            _stream()), "scal_ciphertexts")
        else:
            ex = exps if torch.is_tensor(exps) else self.ints([abs(int(e)) for e in exps], max(1, max((abs(int(e)).bit_length() + 31) // 32 for e in exps)))
            sg = None if torch.is_tensor(exps) else torch.tensor([1 if e < 0 else 0 for e in exps], dtype=torch.uint8, device=self.device)
            capi.check(self.lib.cofhe_b200_scal_ciphertexts(self.h, n, _ptr(ct), _ptr(ex), ex.shape[0], 1, _ptr(sg), 0, _ptr(out),
                                                            _stream()), "scal_ciphertexts")
        return out

    def rerandomize(self, ct, r):
        """(c1 o h^r, c2 o pk^r): r one int (shared) or a list / device tensor [el, n] (per element).  With add_ciphertexts /
        scal_ciphertexts this gives the reference's single-ciphertext add / scal / negate (CL_HSM_utils.inl:188-246), batched."""
        n = ct.shape[-1]
        shared = isinstance(r, int)
        rr = r if torch.is_tensor(r) else self.ints([r] if shared else r, self.el_r)
        out = torch.empty_like(ct)
        capi.check(self.lib.cofhe_b200_rerandomize(self.h, n, _ptr(ct), _ptr(rr), rr.shape[0], int(shared), _ptr(out), _stream()), "rerandomize")
        return out

    def beaver_open(self, ct1, ct2, a, b):
        """(ct1 - a, ct2 - b) as ciphertexts: SMPCCipherTextMultiplier's masking step, on the device."""
        o1, o2 = torch.empty_like(ct1), torch.empty_like(ct2)
        capi.check(self.lib.cofhe_b200_beaver_open(self.h, ct1.shape[-1], _ptr(ct1), _ptr(ct2), _ptr(a), _ptr(b), _ptr(o1), _ptr(o2),
                                                   _stream()), "beaver_open")
        return o1, o2

    def beaver_close(self, pt1, pt2, a, b, c, r):
        """pt1*b + pt2*a + c + Enc_r(pt1*pt2): pt1, pt2 device tensors [ml, n] (the opened values), r one int."""
        n = a.shape[-1]
        rr = self.ints([r], self.el_r)
        out = torch.empty_like(a)
        capi.check(self.lib.cofhe_b200_beaver_close(self.h, n, _ptr(pt1), _ptr(pt2), pt1.shape[0], _ptr(a), _ptr(b), _ptr(c), _ptr(rr),
                                                    rr.shape[0], _ptr(out), _stream()), "beaver_close")
        return out

    def beaver_triplets(self, a, b, r):
        """encrypt_tensor([[a_i, b_i, a_i b_i]]) with the one randomness r: 3n ciphertexts, reference row-major order"""
        ad, bd = (x if torch.is_tensor(x) else self.ints(x, self.ml) for x in (a, b))
        n = ad.shape[-1]
        rr = self.ints([r], self.el_r)
        out = torch.empty((2, self.fl, 3 * n), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_beaver_triplets(self.h, n, _ptr(ad), _ptr(bd), ad.shape[0], _ptr(rr), rr.shape[0], _ptr(out), _stream()),
                   "beaver_triplets")
        return out

    def ct_matmul_expand(self, ct1, ct2, n, m, p):
        o1 = torch.empty((2, self.fl, n * m * p), dtype=torch.int32, device=self.device)
        o2 = torch.empty_like(o1)
        capi.check(self.lib.cofhe_b200_ct_matmul_expand(self.h, n, m, p, _ptr(ct1), _ptr(ct2), _ptr(o1), _ptr(o2), _stream()), "ct_matmul_expand")
        return o1, o2

    def ct_matmul_reduce(self, prod, init, n, m, p):
        out = torch.empty((2, self.fl, n * p), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_ct_matmul_reduce(self.h, n, m, p, _ptr(prod), _ptr(init), _ptr(out), _stream()), "ct_matmul_reduce")
        return out

    def make_plaintexts(self, x, ml=5):
        """float32 device tensor -> (magnitudes [ml, n], signs [n]) = CPUCryptoSystem::make_plaintext of every element"""
        x = x.to(self.device, torch.float32).contiguous()
        n = x.numel()
        mag = torch.empty((ml, n), dtype=torch.int32, device=self.device)
        sg = torch.empty(max(1, n), dtype=torch.uint8, device=self.device)
        capi.check(self.lib.cofhe_b200_make_plaintexts(self.h, n, _ptr(x), _ptr(mag), ml, _ptr(sg), _stream()), "make_plaintexts")
        return mag, sg[:n]

    def plaintexts_to_floats(self, mag, signs=None):
        n = mag.shape[-1]
        out = torch.empty(n, dtype=torch.float32, device=self.device)
        capi.check(self.lib.cofhe_b200_plaintexts_to_floats(self.h, n, _ptr(mag), mag.shape[0], _ptr(signs) if signs is not None else None,
                                                            _ptr(out), _stream()), "plaintexts_to_floats")
        return out

    def _host_exp(self, e):
        el = max(1, (abs(int(e)).bit_length() + 31) // 32)
        return L.ints_to_limbs([abs(int(e))], el)[:, 0].copy(), el

    def part_decrypt(self, ct, share):
        n = ct.shape[-1]
        ex, el = self._host_exp(share)
        out = torch.empty((self.fl, n), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_part_decrypt(self.h, n, _ptr(ct), ex.ctypes.data_as(ctypes.c_void_p), el, int(share < 0),
                                                    _ptr(out), _stream()), "part_decrypt")
        return out

    def combine(self, ct, ds):
        """ds: device tensor [t, fl, n] of partial decryptions (party order as the reference's lambda = (1,-1,-1,...))."""
        n, t = ct.shape[-1], ds.shape[0]
        out = torch.empty((self.ml, n), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_combine(self.h, n, t, _ptr(ct), _ptr(ds), _ptr(out), self.ml, _stream()), "combine")
        return out

    def decrypt(self, ct, sk):
        n = ct.shape[-1]
        ex, el = self._host_exp(sk)
        out = torch.empty((self.ml, n), dtype=torch.int32, device=self.device)
        capi.check(self.lib.cofhe_b200_decrypt(self.h, n, _ptr(ct), ex.ctypes.data_as(ctypes.c_void_p), el, _ptr(out), self.ml,
                                               _stream()), "decrypt")
        return out

    def ints_out(self, t):
        return L.limbs_to_ints(to_host(t))

    def add_ciphertexts_host(self, ct1, ct2, out):
        """numpy uint32 host buffers [2, fl, n]; transfers happen inside the call."""
        n = ct1.shape[-1]
        vp = ctypes.c_void_p
        capi.check(self.lib.cofhe_b200_add_ciphertexts_host(self.h, n, ct1.ctypes.data_as(vp), ct2.ctypes.data_as(vp),
                                                            out.ctypes.data_as(vp)), "add_host")
        return out
