"""Logic tests of the cooperative path (cofhe_b200/csrc/coop*.cuh compiled with g++, the lanes of a group run as fibers by
tests/hostsim/coop_sim.inc): every element it accepts must equal the reference fixtures / the oracle; what it declines
is counted.  Shapes: (T lanes, L limbs per lane) per coefficient width N."""
import random

import numpy as np
import pytest

import hostsim as HS
from conftest import I, load_golden
from cofhe_b200 import limbs as L
from oracle import qf_oracle as O

SHAPES = {7: [(2, 5), (4, 3)], 11: [(2, 7), (4, 4)], 33: [(4, 9), (8, 5)]}


class CoopSim:
    def __init__(self, D, T, LL):
        self.D, self.T, self.L = D, T, LL
        self.dbits = (-D).bit_length()
        self.na, self.nc, self.nd = L.limb_shape(self.dbits)
        self.lib = HS.load()

    def run(self, f1, f2=None, inv=False, dupl=False):
        n = len(f1)
        a = L.forms_to_limbs(f1, self.na, self.nc)
        b = L.forms_to_limbs(f2 if f2 is not None else f1, self.na, self.nc)
        o = np.zeros_like(a)
        bail = np.zeros(n, dtype=np.uint8)
        nb = self.lib.hs_coop_compose(self.T, self.L, self.na, self.dbits, self.nc, n, HS.P(a), HS.P(b), HS.P(o), int(inv), int(dupl), HS.P(bail))
        assert nb >= 0
        return L.limbs_to_forms(o, self.na, self.nc), bail


def check(got, bail, want):
    bad = [i for i, (g, w) in enumerate(zip(got, want)) if not bail[i] and g != w]
    assert not bad, "cooperative path wrong at %s" % bad[:5]


def sims(tier):
    D = int(tier["params"]["Delta"])
    na = L.limb_shape((-D).bit_length())[0]
    return [CoopSim(D, t, l) for t, l in SHAPES.get(na, [])]


def test_coop_fixtures(tier):
    ss = sims(tier)
    if not ss:
        pytest.skip("no cooperative instantiation for this width")
    vs = tier["compose"] + tier["compose_edge"]
    nr = len(tier["compose"])
    f1, f2 = [I(v["f1"]) for v in vs], [I(v["f2"]) for v in vs]
    for S in ss:
        g, b = S.run(f1, f2); check(g, b, [I(v["r"]) for v in vs])
        assert int(b[:nr].sum()) == 0, "random reduced forms must stay on the cooperative path"
        g, b = S.run(f2, f1); check(g, b, [I(v["r"]) for v in vs])
        g, b = S.run(f1, f2, inv=True); check(g, b, [I(v["rinv"]) for v in vs])
        g, b = S.run(f1, dupl=True); check(g, b, [I(v["sq"]) for v in vs])
        assert int(b[:nr].sum()) == 0


@pytest.mark.parametrize("name", ["t0", "node", "t1"])
def test_coop_random_chains(name):
    """products of random group elements: compose, inverse-compose, square; gcd(a1, a2) != 1 pairs included"""
    t = load_golden("ref_%s.json" % name)
    rng = random.Random(77)
    h, pk = I(t["params"]["h"]), I(t["pk"])
    n = 24 if name != "t1" else 8
    base = [O.power(h, rng.getrandbits(40) + 1) for _ in range(n)] + [O.power(pk, rng.getrandbits(40) + 1) for _ in range(n)]
    f1 = base
    f2 = base[::-1]
    for S in sims(t):
        g, b = S.run(f1, f2); check(g, b, [O.compose(x, y) for x, y in zip(f1, f2)])
        assert int(b.sum()) == 0
        g, b = S.run(f1, f2, inv=True); check(g, b, [O.compose_inv(x, y) for x, y in zip(f1, f2)])
        g, b = S.run(f1, dupl=True); check(g, b, [O.square(x) for x in f1])
        assert int(b.sum()) == 0


def test_coop_full_warp_case():
    """one full warp (8 groups of 4 lanes in lock step, each on its own 2053-bit pair) of tests/golden/coop_case_t1.bin"""
    import os
    import struct
    from conftest import GOLDEN
    raw = open(os.path.join(GOLDEN, "coop_case_t1.bin"), "rb").read()
    dbits, nc, n = struct.unpack("III", raw[:12])
    na = 33
    arr = np.frombuffer(raw[12:], dtype=np.uint32).reshape(3, 2 * na + nc, n).copy()
    lib = HS.load()
    lib.hs_coop_groups(0)  # a full warp
    o = np.zeros_like(arr[0])
    bail = np.zeros(n, dtype=np.uint8)
    assert lib.hs_coop_compose(4, 9, na, dbits, nc, n, HS.P(arr[0]), HS.P(arr[1]), HS.P(o), 0, 0, HS.P(bail)) == 0
    lib.hs_coop_groups(2)
    assert (o == arr[2]).all()


@pytest.mark.parametrize("name,shape", [("t0", (2, 5)), ("t0", (4, 3)), ("node", (2, 7)), ("t1", (4, 9))])
def test_coop_pow(name, shape):
    """the exponentiation state machine on the cooperative path: shared and per-element exponents, negative exponents,
    leading zero windows, exponent 0 (declined: the neutral form does not fit the working shape)"""
    t = load_golden("ref_%s.json" % name)
    D = int(t["params"]["Delta"])
    S = CoopSim(D, *shape)
    rng = random.Random(5)
    h, pk = I(t["params"]["h"]), I(t["pk"])
    n = 5 if name != "t1" else 2
    bits = 40 if name != "t1" else 24
    f = [O.power(h, rng.getrandbits(30) + 2) for _ in range(n - 1)] + [pk]
    a = L.forms_to_limbs(f, S.na, S.nc)
    lib = S.lib
    # shared exponent, negative
    e = rng.getrandbits(bits) | (1 << (bits - 1))
    el = 2
    ex = L.ints_to_limbs([e], el)[:, 0].copy()
    o = np.zeros_like(a); bail = np.zeros(n, dtype=np.uint8)
    nb = lib.hs_coop_pow(S.T, S.L, S.na, S.dbits, S.nc, n, HS.P(a), HS.P(o), HS.P(ex), el, 0, None, 1, HS.P(bail))
    assert nb == 0
    assert L.limbs_to_forms(o, S.na, S.nc) == [O.power(x, -e) for x in f]
    # per-element exponents of different lengths (leading zero windows), one of them 0, mixed signs
    es = [rng.getrandbits(rng.choice([3, 9, 17, bits])) + 1 for _ in range(n)]
    es[1] = 0
    sg = np.array([i & 1 for i in range(n)], dtype=np.uint8)
    exs = L.ints_to_limbs(es, el)
    o = np.zeros_like(a); bail = np.zeros(n, dtype=np.uint8)
    nb = lib.hs_coop_pow(S.T, S.L, S.na, S.dbits, S.nc, n, HS.P(a), HS.P(o), HS.P(exs), el, 1, HS.P(sg), 0, HS.P(bail))
    assert bail.tolist() == [1 if x == 0 else 0 for x in es]
    got = L.limbs_to_forms(o, S.na, S.nc)
    for i in range(n):
        if es[i]:
            assert got[i] == O.power(f[i], -es[i] if sg[i] else es[i]), i


@pytest.mark.parametrize("name,shape", [("t0", (2, 5)), ("t1", (4, 9))])
def test_coop_pow_shared_sliding_windows(name, shape):
    """cq_pow with an exponent the launch shares: every small exponent, bit patterns across limb boundaries, exponent 0
    (declined).  (Written for a sliding-window schedule of cq_pow that was taken out again -- DESIGN.md section 9 -- and kept
    for the fixed-window one.)"""
    t = load_golden("ref_%s.json" % name)
    S = CoopSim(int(t["params"]["Delta"]), *shape)
    rng = random.Random(9)
    h, pk = I(t["params"]["h"]), I(t["pk"])
    f = [O.power(h, rng.getrandbits(20) + 2), pk]
    n = len(f)
    a = L.forms_to_limbs(f, S.na, S.nc)
    small = list(range(1, 36)) if name == "t0" else [1, 2, 3, 5, 6, 7, 8, 11, 12, 15, 16, 17, 23, 24, 31, 32, 33]
    big = [(1 << 31) + 1, (1 << 32) - 1, 1 << 32, (1 << 33) | 5, 0xF0F0F0F0F0F, rng.getrandbits(62) | (1 << 61)] if name == "t0" else [(1 << 32) | 9, 0xF0F0F0F0F]
    for e in small + big:
        for el in (2, 3):
            ex = L.ints_to_limbs([e], el)[:, 0].copy()
            o = np.zeros_like(a); bail = np.zeros(n, dtype=np.uint8)
            nb = S.lib.hs_coop_pow(S.T, S.L, S.na, S.dbits, S.nc, n, HS.P(a), HS.P(o), HS.P(ex), el, 0, None, 0, HS.P(bail))
            assert nb == 0, e
            assert L.limbs_to_forms(o, S.na, S.nc) == [O.power(x, e) for x in f], hex(e)
    ex = L.ints_to_limbs([0], 2)[:, 0].copy()
    o = np.zeros_like(a); bail = np.zeros(n, dtype=np.uint8)
    S.lib.hs_coop_pow(S.T, S.L, S.na, S.dbits, S.nc, n, HS.P(a), HS.P(o), HS.P(ex), 2, 0, None, 0, HS.P(bail))
    assert bail.tolist() == [1] * n


def test_coop_slow_compose_fallback(tier):
    """cq_slow_compose (coop_ops.cuh): the robust composition one lane runs for its group when the cooperative composition
    declines inside an exponentiation / table loop -- forced for every group, and for every other group of a warp"""
    ss = sims(tier)
    if not ss:
        pytest.skip("no cooperative instantiation for this width")
    vs = tier["compose"]
    f1, f2 = [I(v["f1"]) for v in vs], [I(v["f2"]) for v in vs]
    lib = HS.load()
    try:
        for S in ss[:1]:
            lib.hs_coop_force_slow(2)
            g, b = S.run(f1, f2)
            assert int(b.sum()) == 0 and g == [I(v["r"]) for v in vs]
            g, b = S.run(f1, dupl=True)
            assert int(b.sum()) == 0 and g == [I(v["sq"]) for v in vs]
            lib.hs_coop_force_slow(1)
            g, b = S.run(f1, f2)
            groups = 32 // S.T
            for i, v in enumerate(vs):
                if (i % groups) % 2 == 0:
                    assert not b[i] and g[i] == I(v["r"])
                else:
                    assert b[i]
    finally:
        lib.hs_coop_force_slow(0)


def test_coop_t2_shape():
    """the 8-lane x 7-limb shape of the 53-limb coefficients (N of 3072 bits: |Delta| = 3333 bits) against reference vectors"""
    t = load_golden("ref_t2.json")
    S = CoopSim(int(t["params"]["Delta"]), 8, 7)
    assert S.na == 53
    vs = t["compose"] + t["compose_edge"]
    nr = len(t["compose"])
    f1, f2 = [I(v["f1"]) for v in vs], [I(v["f2"]) for v in vs]
    g, b = S.run(f1, f2); check(g, b, [I(v["r"]) for v in vs])
    assert int(b[:nr].sum()) == 0
    g, b = S.run(f1, f2, inv=True); check(g, b, [I(v["rinv"]) for v in vs])
    g, b = S.run(f1, dupl=True); check(g, b, [I(v["sq"]) for v in vs])
    assert int(b[:nr].sum()) == 0


@pytest.mark.parametrize("name,T,LL", [("t0", 2, 5), ("t1", 4, 9)])
def test_coop_tabprod_body(name, T, LL):
    """the window-table multi-exponentiation body (cq_tabprod: scal_matmul and fixed base on the cooperative kernels) in the
    simulator: tables T[d][w][base] = base^(d 2^(W w)) from the oracle, a 2 x 3 . 3 x 2 product with an `init` ciphertext half,
    two chunks of the inner index, zero digits and a zero scalar; partial products against the oracle"""
    t = load_golden("ref_%s.json" % name)
    D = int(t["params"]["Delta"])
    S = CoopSim(D, T, LL)
    lib = S.lib
    n, m, p, W = 2, 3, 2, 2
    rng = random.Random(12)
    bases = [I(v["f1"]) for v in t["compose"][:n * m]]
    sc = [rng.getrandbits(7) for _ in range(m * p)]
    sc[1] = 0
    sc[2] = 0b0100011
    el, nwin, nt = 1, 4, (1 << W) - 1
    nb = n * m
    ne = nt * nwin * nb
    ent = [None] * ne
    for b in range(nb):
        for w in range(nwin):
            for d in range(1, nt + 1):
                ent[(d - 1) * nwin * nb + w * nb + b] = O.power(bases[b], d << (W * w))
    tab = L.forms_to_limbs(ent, S.na, S.nc)
    init = I(t["compose"][7]["f1"])
    initl = L.forms_to_limbs([init], S.na, S.nc)
    scal = L.ints_to_limbs(sc, el)
    chunks, jper = 2, 2
    fl = 2 * S.na + S.nc
    out = np.zeros((chunks, fl, n * p), dtype=np.uint32)
    flagged = np.zeros(chunks * n * p, dtype=np.uint8)
    nbad = lib.hs_coop_tabprod(T, LL, S.na, S.dbits, S.nc, n, m, p, nwin, W, el, chunks, jper, HS.P(scal), HS.P(tab), ne, HS.P(initl), HS.P(out), HS.P(flagged))
    assert nbad >= 0
    for c in range(chunks):
        got = L.limbs_to_forms(out[c], S.na, S.nc)
        for i in range(n):
            for k in range(p):
                o = i * p + k
                acc = init if c == 0 else None
                for j in range(c * jper, min((c + 1) * jper, m)):
                    if sc[j * p + k]:
                        f = O.power(bases[i * m + j], sc[j * p + k])
                        acc = f if acc is None else O.compose(acc, f)
                if acc is None:
                    assert flagged[c * n * p + o]      # empty product: robust path
                else:
                    assert not flagged[c * n * p + o] and got[o] == acc, (c, i, k)
