"""Host-side marshalling between Python integers and the C ABI's structure-of-arrays limb buffers.

Layout (include/cofhe_b200.h "Layout"): a batch of `n` integers of `nl` 32-bit limbs is a uint32 array of shape
[nl, n] -- limb-major, element-minor, little-endian limbs -- so that consecutive threads touch consecutive
words.  Signed values (the `b` coefficient) are two's complement over their `nl` limbs.  A batch of forms is
[2*na + nc, n] = a[na] | b[na] | c[nc]; a batch of ciphertexts is [2, 2*na + nc, n] = c1 | c2.
"""
import numpy as np


def ints_to_limbs(vals, nl, signed=False):
    n = len(vals)
    out = np.zeros((nl, n), dtype=np.uint32)
    mod = 1 << (32 * nl)
    for j, v in enumerate(vals):
        v = int(v)
        if signed:
            if not (-(mod >> 1) <= v < (mod >> 1)):
                raise OverflowError("value does not fit %d signed limbs" % nl)
            v %= mod
        elif not (0 <= v < mod):
            raise OverflowError("value does not fit %d limbs" % nl)
        out[:, j] = np.frombuffer(v.to_bytes(4 * nl, "little"), dtype="<u4")
    return out


def limbs_to_ints(arr, signed=False):
    arr = np.ascontiguousarray(arr, dtype=np.uint32)
    nl, n = arr.shape
    cols = np.ascontiguousarray(arr.T)
    mod = 1 << (32 * nl)
    out = []
    for j in range(n):
        v = int.from_bytes(cols[j].astype("<u4").tobytes(), "little")
        if signed and v >= (mod >> 1):
            v -= mod
        out.append(v)
    return out


def forms_to_limbs(forms, na, nc):
    """forms: sequence of (a, b, c) -> uint32 [2*na+nc, n]."""
    a = ints_to_limbs([f[0] for f in forms], na)
    b = ints_to_limbs([f[1] for f in forms], na, signed=True)
    c = ints_to_limbs([f[2] for f in forms], nc)
    return np.concatenate([a, b, c], axis=0)


def limbs_to_forms(arr, na, nc):
    arr = np.asarray(arr)
    a = limbs_to_ints(arr[:na])
    b = limbs_to_ints(arr[na:2 * na], signed=True)
    c = limbs_to_ints(arr[2 * na:2 * na + nc])
    return list(zip(a, b, c))


def cts_to_limbs(cts, na, nc):
    """cts: sequence of ((a,b,c),(a,b,c)) -> uint32 [2, 2*na+nc, n]."""
    return np.stack([forms_to_limbs([c[0] for c in cts], na, nc), forms_to_limbs([c[1] for c in cts], na, nc)])


def limbs_to_cts(arr, na, nc):
    arr = np.asarray(arr)
    return list(zip(limbs_to_forms(arr[0], na, nc), limbs_to_forms(arr[1], na, nc)))


def limb_shape(delta_bits):
    """(na, nc, nd): limbs of a/b (reduced: |b| <= a <= sqrt(|D|/3), plus a sign bit), of c (up to |D|/4) and of |D|."""
    nd = (delta_bits + 31) // 32
    na = (delta_bits // 2 + 1 + 31) // 32
    nc = nd
    return na, nc, nd
