"""ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product path (cofhe_b200/).

CPU restatement, on Python big integers, of the reference's class-group hot path: canonical reduction,
composition, exponentiation, and the CL_HSM2k scheme operations that CoFHE's CPUCryptoSystem drives.
Only tests/, tools/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import it.

Every public reference operation returns the UNIQUE reduced form of its class, so the restatement is free
to use the textbook composition (Cohen, Alg. 5.4.7 style) instead of BICYCL's NUCOMP:
results are identical integer triples.  Each function cites the reference code it restates
(paths relative to /root/reference).

Pinning: tests/test_oracle.py checks this file against
  * the reference's own 150 nupow known-answer vectors (thirdparty/bicycl/tests/test_qfi_nupow.data.hpp,
    committed as tests/golden/bicycl_nupow_kat.json), and
  * fixtures generated from the compiled reference itself (oracle/_ref, tools/gen_golden.py ->
    tests/golden/ref_*.json) for every scheme-level operation,
and, when oracle/_ref is present, against live calls of the reference.
"""
from math import isqrt

# --------------------------------------------------------------------------------------------------
# Forms: tuples (a, b, c) of Python ints, a > 0, disc = b*b - 4*a*c < 0.
# --------------------------------------------------------------------------------------------------


def disc(f):
    a, b, c = f
    return b * b - 4 * a * c


def xgcd(x, y):
    """(g, u, v) with g = u*x + v*y >= 0.  Any valid cofactors do (thirdparty/bicycl/src/bicycl/gmp_extras.inl:1044-1048)."""
    u0, u1, v0, v1 = 1, 0, 0, 1
    while y:
        q, r = divmod(x, y)
        x, y = y, r
        u0, u1 = u1, u0 - q * u1
        v0, v1 = v1, v0 - q * v1
    if x < 0:
        return -x, -u0, -v0
    return x, u0, v0


def normalize(f):
    """-a < b <= a  (thirdparty/bicycl/src/bicycl/qfi.inl:473-485)."""
    a, b, c = f
    # b = 2*a*q + r with -a < r <= a
    q = -((a - b) // (2 * a))
    r = b - 2 * a * q
    c = c - q * (b + r) // 2
    return (a, r, c)


def reduce_form(f):
    """Canonical reduced form: normal, a <= c, and b >= 0 when a == c (qfi.inl:523-548)."""
    a, b, c = normalize(f)
    while a > c:
        a, b, c = normalize((c, -b, a))
    if a == c and b < 0:
        b = -b
    return (a, b, c)


def is_reduced(f):
    a, b, c = f
    return -a < b <= a and a <= c and not (a == c and b < 0)


def identity(D):
    """Neutral form of discriminant D (qfi.inl:1333-1345)."""
    b = D % 2
    return (1, b, (b - D) // 4)


def inverse(f):
    """QFI::neg (qfi.inl:171-176): (a,-b,c), unless that is not reduced (a == c or a == b)."""
    a, b, c = f
    if a == c or a == b:
        return f
    return (a, -b, c)


def compose(f1, f2):
    """Reduced composition f1 o f2 -- same result as ClassGroup::nucomp (qfi.inl:683-818, 1441-1445)."""
    a1, b1, c1 = f1
    a2, b2, c2 = f2
    D = b2 * b2 - 4 * a2 * c2
    s = (b1 + b2) // 2
    g, u, v = xgcd(a1, a2)
    d, x, y = xgcd(g, s)
    v1, v2 = a1 // d, a2 // d
    r = ((x * v) * ((b1 - b2) // 2) - y * c2) % v1
    a3 = v1 * v2
    b3 = b2 + 2 * v2 * r
    c3 = (b3 * b3 - D) // (4 * a3)
    assert (b3 * b3 - D) % (4 * a3) == 0
    return reduce_form((a3, b3, c3))


def compose_inv(f1, f2):
    """f1 o f2^-1 (ClassGroup::nucompinv, qfi.inl:1447-1451).  NB: uses the raw conjugate (a,-b,c) like the
    reference's negf2 flag, which is the true group inverse."""
    a, b, c = f2
    return compose(f1, (a, -b, c))


def square(f):
    """ClassGroup::nudupl (qfi.inl:861-955, 1453-1471)."""
    return compose(f, f)


def power(f, n):
    """ClassGroup::nupow (qfi.inl:977-1064, 1473-1477): f^n, n any sign; f^0 = identity; n<0 -> QFI::neg of f^|n|."""
    D = disc(f)
    if n == 0:
        return identity(D)
    e = abs(n)
    r = f
    for bit in bin(e)[3:]:
        r = square(r)
        if bit == "1":
            r = compose(r, f)
    if n < 0:
        r = inverse(r)
    return r


def nucomp_bound(D):
    """ClassGroup default_nucomp_bound = floor(|D|^(1/4)) (qfi.inl:1303-1308)."""
    return isqrt(isqrt(abs(D)))


# --------------------------------------------------------------------------------------------------
# CL_HSM2k scheme (thirdparty/bicycl/src/bicycl/CL_HSM2k.inl, CL_HSM_utils.inl)
# --------------------------------------------------------------------------------------------------


def val2(x):
    return (x & -x).bit_length() - 1


def mod2k_centered(x, k):
    """Mpz::mod2k_centered (gmp_extras.inl:863-876): representative in [-2^(k-1), 2^(k-1))."""
    m = 1 << k
    r = x % m
    if r >= m // 2:
        r -= m
    return r


class Params:
    """What CL_HSM2k holds after setup (CL_HSM2k.inl:31-103).  Setup itself (random N, class-number bound)
    is out of scope: values come from the reference (tests/golden/*.json) or from explicit arguments."""

    def __init__(self, N, k, exponent_bound, h, d=None, e=None):
        self.N, self.k = N, k
        self.DeltaK = -8 * N                                   # CL_HSM2k.inl:620-627
        self.Delta = self.DeltaK << (2 * (k + 1))              # CL_HSM2k.inl:631-637
        self.M = 1 << k
        self.large_message = (self.M * self.M - 1 + self.DeltaK) > 0   # CL_HSM2k.inl:51-56
        self.exponent_bound = exponent_bound
        self.h = tuple(h)
        self.d = d if d is not None else (exponent_bound.bit_length() + 1) // 2  # CL_HSM2k.inl:92-93
        self.e = e if e is not None else self.d // 2 + 1

    @classmethod
    def from_dict(cls, p):
        return cls(p["N"], p["k"], p["exponent_bound"], p["h"], p.get("d"), p.get("e"))


def power_of_h(P, r):
    """CL_HSM2k::power_of_h (CL_HSM2k.inl:318-322): h^r (the fixed-base algorithm is an implementation detail)."""
    return power(P.h, r)


def power_of_f(P, m):
    """CL_HSM2k::power_of_f (CL_HSM2k.inl:326-405).  Returns the same UNREDUCED triple as the reference."""
    k = P.k
    mk = (1 << k) - 1
    v = val2(m) if m else k
    if m == 0 or v >= k:
        return identity(P.Delta)
    m00, m01, m10, m11 = 1, 0, 0, 1
    minusQ = P.DeltaK - 1
    for i in range(k, 0, -1):
        t0 = (m00 * m00 + m01 * m10) & mk
        t1 = (m00 * m01 + m01 * m11) & mk
        t2 = (m10 * m00 + m11 * m10) & mk
        t3 = (m10 * m01 + m11 * m11) & mk
        m00, m01, m10, m11 = t0, t1, t2, t3
        if (m >> (i - 1)) & 1:
            n00 = (m00 * minusQ + 2 * m10) & mk
            n01 = (m01 * minusQ + 2 * m11) & mk
            m00, m10 = m10, n00
            m01, m11 = m11, n01
    t0 = m00 + m01
    t1 = m01 >> v
    kv = k - v
    t2 = pow(t1, -1, 1 << kv) if kv > 0 else 0
    u = mod2k_centered(t0 * t2, kv)
    a = 1 << (2 * kv)
    b = u << (kv + 1)
    c = (P.N << (2 * v + 3)) + u * u
    return (a, b, c)


def _kernel_representative_2exp(P, f):
    """QFI::kernel_representative_2exp(k+1, DeltaK) (qfi.inl:396-435) incl. to_maximal_order_2exp (:311-328)
    and prime_to_2exp (:590-610)."""
    kk = P.k + 1
    a, b, c = f
    if a % 2 == 0:
        if c % 2 == 0:
            c = c + a + b
            b = -(b + 2 * a)
            a, c = c, a
        else:
            a, c = c, a
            b = -b
    v = pow(a, -1, 1 << kk)
    u = -(v * a - 1)
    assert u % (1 << kk) == 0
    u >>= kk
    b = b * u
    c = (b * b - P.DeltaK) // (4 * a)
    assert (b * b - P.DeltaK) % (4 * a) == 0
    g0, g1 = 1, 0
    a, b, c = normalize((a, b, c))
    while a > c:
        g0, g1 = g0 * b + g1 * P.DeltaK, g1 * b + g0
        a, b, c = normalize((c, -b, a))
    if a != 1 or b != 0:
        raise ValueError("the form is not in the kernel")
    w = val2(g0)
    g0 >>= w
    g1 >>= w
    t = pow(g0, -1, 1 << kk) % (1 << kk)
    return (t * (-g1)) % (1 << kk)


def _F_kerphi_square(P, t):
    k, mk = P.k, (1 << P.k) - 1
    a0 = (((t * t) & mk) * P.DeltaK + 1) & mk
    return (2 * t * pow(a0, -1, 1 << k)) & mk


def _F_kerphi_div(P, t, s):
    k, mk = P.k, (1 << P.k) - 1
    a0 = (-(((t * s) & mk) * P.DeltaK - 1)) & mk
    return ((t - s) * pow(a0, -1, 1 << k)) & mk


def dlog_in_F(P, fm):
    """CL_HSM2k::dlog_in_F (CL_HSM2k.inl:408-455)."""
    if fm == identity(P.Delta):
        return 0
    k = P.k
    if P.large_message:
        tm = _kernel_representative_2exp(P, fm)
    else:
        b = fm[1]
        j = val2(b) - 1
        u = b >> (j + 1)
        tm = (pow(u, -1, 1 << j) if j > 0 else 0) << (k - j)
    tm &= (1 << k) - 1
    t = 1
    m = 0
    for i in range(k):
        if tm != 0 and val2(tm) == i:
            m |= 1 << i
            tm = _F_kerphi_div(P, tm, t)
        t = _F_kerphi_square(P, t)
    return m


def prime_to_2exp(f):
    """QFI::prime_to_2exp (qfi.inl:590-610): an equivalent form whose a is odd."""
    a, b, c = f
    if a % 2 == 0:
        if c % 2 == 0:
            return (a + b + c, -b - 2 * a, a)
        return (c, -b, a)
    return f


def lift_to_delta(P, f):
    """CL_HSM2k::from_Cl_DeltaK_to_Cl_Delta (CL_HSM2k.inl:458-474), compact variant: QFI::lift_2exp(k+1)
    (qfi.inl:263-271), chi from a and c modulo 4 of the lifted reduced form, k squarings in Cl(Delta)
    (raise_to_power_M, :641-648), and for chi = -1 one composition with (4, 4, 1 - DeltaK 2^(2(k-1)))."""
    a, b, c = prime_to_2exp(f)
    k = P.k
    g = reduce_form((a, b << (k + 1), c << (2 * (k + 1))))
    chi_neg = g[0] % 4 == 3 or g[2] % 4 == 3
    g = power(g, 1 << k)
    if chi_neg:
        g = compose(g, (4, 4, 1 - (P.DeltaK << (2 * (k - 1)))))
    return g


def encrypt_compact(P, pk, m, r):
    """CL_HSM_CipherText(C, pk, m, r) with C.compact_variant() (CL_HSM_utils.inl:161-185): c1 = h^r in Cl(DeltaK),
    c2 = lift(pk^r) o f^m in Cl(Delta); P.h and pk are forms of discriminant DeltaK."""
    return (power(P.h, r), compose(lift_to_delta(P, power(pk, r)), power_of_f(P, m)))


def decrypt_compact(P, sk, ct):
    """CL_HSM_ClearText(C, sk, c) with C.compact_variant() (CL_HSM_utils.inl:124-138)."""
    return dlog_in_F(P, compose_inv(ct[1], lift_to_delta(P, power(ct[0], sk))))


def part_decrypt_compact(P, share, ct):
    """partDecrypt with compact_variant (cpu_cryptosystem_distributed.inl:244-254)."""
    return lift_to_delta(P, power(ct[0], share))


def encrypt(P, pk, m, r):
    """CL_HSM_CipherText(C, pk, m, r) (CL_HSM_utils.inl:161-185): (h^r, pk^r o f^m)."""
    c1 = power_of_h(P, r)
    c2 = compose(power(pk, r), power_of_f(P, m))
    return (c1, c2)


def encrypt_shared(P, c1, pkr, m):
    """CL_HSM_CipherText(C, m, c1, pkr) (CL_HSM_utils.inl:248-255): batched encryption with shared randomness
    (CPUCryptoSystem::encrypt_tensor, include/x86_64/cpu_cryptosystem_tensor_ops.inl:1-19)."""
    return (c1, compose(pkr, power_of_f(P, m)))


def decrypt(P, sk, ct):
    """CL_HSM_ClearText(C, sk, c) (CL_HSM_utils.inl:124-138)."""
    c1, c2 = ct
    return dlog_in_F(P, compose_inv(c2, power(c1, sk)))


def add_ciphertexts(ct1, ct2):
    """Batched add, no re-randomisation (cpu_cryptosystem_tensor_ops.inl:242-264)."""
    return (compose(ct1[0], ct2[0]), compose(ct1[1], ct2[1]))


def scal_ciphertext(ct, s):
    """Batched scal, no re-randomisation (cpu_cryptosystem_tensor_ops.inl:280-340)."""
    return (power(ct[0], s), power(ct[1], s))


def scal_matmul(zero, ct, s, n, m, p):
    """CPUCryptoSystem::scal_ciphertext_tensors, matmul branch (cpu_cryptosystem_tensor_ops.inl:346-461):
    res[i,k] = zero o prod_j ct[i,j]^{s[j,k]} accumulated in j order; ct is n x m, s is m x p, row-major lists.
    Uses the mathematically correct power (ClassGroup::nupow), not the defective qfi_nupow."""
    res = []
    for i in range(n):
        for kk in range(p):
            c1, c2 = zero
            for j in range(m):
                c1 = compose(c1, power(ct[i * m + j][0], s[j * p + kk]))
                c2 = compose(c2, power(ct[i * m + j][1], s[j * p + kk]))
            res.append((c1, c2))
    return res


def part_decrypt(share, ct):
    """partDecrypt (include/x86_64/cpu_cryptosystem_distributed.inl:244-254): c1^share."""
    return power(ct[0], share)


def combine(P, ct, ds):
    """finalDecrypt (cpu_cryptosystem_distributed.inl:256-270) with lambda = (1,-1,-1,...) (:214-229):
    m = dlog_F(c2 o (d_0 o d_1^-1 o ...)^-1).  compute_d starts from QFI()=(1,1,1), which composes as the neutral
    element because a == 1; the restatement starts from the true identity."""
    d = identity(P.Delta)
    for i, di in enumerate(ds):
        d = compose(d, power(di, 1 if i == 0 else -1))
    return dlog_in_F(P, compose_inv(ct[1], d))


# ---- binary tensor wire format -------------------------------------------------------------
def _wire_size(x):
    """bytes reserved for one integer: mpz_sizeinbase(x, 2) / 8 + 1 (cpu_cryptosystem.inl:340-356; sizeinbase(0) = 1)."""
    return max(abs(x).bit_length(), 1) // 8 + 1


def serialize_ciphertext_tensor(shape, cts):
    """CPUCryptoSystem::serialize_ciphertext_tensor (include/x86_64/cpu_cryptosystem.inl:322-391): u32 ndim, u32 dims,
    then 6 u64 offsets per ciphertext (c1.a c1.b c1.c c2.a c2.b c2.c; bit 63 set when sgn != 1, i.e. negative OR zero),
    then the magnitudes, little endian (mpz_export order -1, size 1), each in its reserved `_wire_size` bytes."""
    import struct
    ints = [x for ct in cts for f in ct for x in f]
    head = struct.pack("<I", len(shape)) + b"".join(struct.pack("<I", d) for d in shape)
    table, blobs, off = [], [], 0
    for x in ints:
        table.append(off | ((1 << 63) if x <= 0 else 0))
        sz = _wire_size(x)
        blobs.append(abs(x).to_bytes(sz, "little"))
        off += sz
    return head + b"".join(struct.pack("<Q", t) for t in table) + b"".join(blobs)


def deserialize_ciphertext_tensor(data):
    """CPUCryptoSystem::deserialize_ciphertext_tensor (cpu_cryptosystem.inl:393-520): the size of an integer is the
    distance to the next offset (the last one runs to the end of the buffer).  -> (shape, cts)"""
    import struct
    ndim = struct.unpack_from("<I", data, 0)[0]
    shape = list(struct.unpack_from("<%dI" % ndim, data, 4))
    n = 1
    for d in shape:
        n *= d
    base = 4 + 4 * ndim
    table = struct.unpack_from("<%dQ" % (6 * n), data, base)
    body = base + 48 * n
    mask = (1 << 63) - 1
    ints = []
    for i, t in enumerate(table):
        lo = t & mask
        hi = (table[i + 1] & mask) if i + 1 < len(table) else len(data) - body
        v = int.from_bytes(data[body + lo:body + hi], "little")
        ints.append(-v if t >> 63 else v)
    cts = [(tuple(ints[6 * i:6 * i + 3]), tuple(ints[6 * i + 3:6 * i + 6])) for i in range(n)]
    return shape, cts


def serialize_int_tensor(shape, ints):
    """The shared layout of the reference's binary tensor formats for a flat list of integers: serialize_plaintext_tensor
    (include/x86_64/cpu_cryptosystem.inl:229-271; one integer per element) and, with (a, b, c) per element,
    serialize_part_decryption_result_tensor (:510-560)."""
    import struct
    head = struct.pack("<I", len(shape)) + b"".join(struct.pack("<I", d) for d in shape)
    table, blobs, off = [], [], 0
    for x in ints:
        table.append(off | ((1 << 63) if x <= 0 else 0))
        sz = _wire_size(x)
        blobs.append(abs(x).to_bytes(sz, "little"))
        off += sz
    return head + b"".join(struct.pack("<Q", t) for t in table) + b"".join(blobs)


def deserialize_int_tensor(data, per_element):
    """deserialize_plaintext_tensor (cpu_cryptosystem.inl:273-318, per_element = 1) / deserialize_part_decryption_result_tensor
    (:562-635, per_element = 3).  -> (shape, flat integers)"""
    import struct
    ndim = struct.unpack_from("<I", data, 0)[0]
    shape = list(struct.unpack_from("<%dI" % ndim, data, 4))
    n = 1
    for d in shape:
        n *= d
    base = 4 + 4 * ndim
    table = struct.unpack_from("<%dQ" % (per_element * n), data, base)
    body = base + 8 * per_element * n
    mask = (1 << 63) - 1
    ints = []
    for i, t in enumerate(table):
        lo = t & mask
        hi = (table[i + 1] & mask) if i + 1 < len(table) else len(data) - body
        v = int.from_bytes(data[body + lo:body + hi], "little")
        ints.append(-v if t >> 63 else v)
    return shape, ints
