"""Logic tests of the fixed-width fast path (cofhe_b200/csrc/qf_fast.cuh compiled with g++) in isolation: every
element it accepts must equal the oracle / the reference fixtures; what it declines (bail) is counted."""
import random

import numpy as np
import pytest

import hostsim as HS
from conftest import I, load_golden
from cofhe_b200 import limbs as L
from oracle import qf_oracle as O

FAST_N = (2, 4, 7, 11)


class FastSim:
    def __init__(self, D):
        self.D = D
        self.dbits = (-D).bit_length()
        self.na, self.nc, self.nd = L.limb_shape(self.dbits)
        self.lib = HS.load()
        self.absd = L.ints_to_limbs([-D], self.nd)[:, 0].copy()

    def ok(self):
        return self.na in FAST_N

    def run(self, f1, f2=None, inv=False, dupl=False):
        n = len(f1)
        a = L.forms_to_limbs(f1, self.na, self.nc)
        b = L.forms_to_limbs(f2 if f2 is not None else f1, self.na, self.nc)
        o = np.zeros_like(a)
        bail = np.zeros(n, dtype=np.uint8)
        nb = self.lib.hs_fast_compose(self.na, self.dbits, self.nc, n, HS.P(a), HS.P(b), HS.P(o), int(inv), int(dupl), HS.P(bail))
        assert nb >= 0
        return L.limbs_to_forms(o, self.na, self.nc), bail


def check(got, bail, want):
    bad = [i for i, (g, w) in enumerate(zip(got, want)) if not bail[i] and g != w]
    assert not bad, "fast path wrong at %s" % bad[:5]


def test_fast_fixtures(tier):
    S = FastSim(int(tier["params"]["Delta"]))
    if not S.ok():
        pytest.skip("no fast instantiation for %d limbs" % S.na)
    vs = tier["compose"] + tier["compose_edge"]
    f1, f2 = [I(v["f1"]) for v in vs], [I(v["f2"]) for v in vs]
    g, b = S.run(f1, f2); check(g, b, [I(v["r"]) for v in vs])
    nb_rand = int(b[:len(tier["compose"])].sum())
    assert nb_rand == 0, "random reduced forms must stay on the fast path"
    g, b = S.run(f2, f1); check(g, b, [I(v["r"]) for v in vs])
    g, b = S.run(f1, f2, inv=True); check(g, b, [I(v["rinv"]) for v in vs])
    g, b = S.run(f1, dupl=True); check(g, b, [I(v["sq"]) for v in vs])
    assert int(b[:len(tier["compose"])].sum()) == 0


def test_fast_kat_chains():
    """Drive the fast path through the reference's 150 nupow vectors with a Python-side square-and-multiply."""
    kat = load_golden("bicycl_nupow_kat.json")["vectors"]
    by = {}
    for v in kat:
        f, n, r = I(v["f"]), int(v["n"]), I(v["r"])
        by.setdefault(O.disc(f), []).append((f, n, r))
    ran = 0
    for D, vs in by.items():
        S = FastSim(D)
        if not S.ok():
            continue
        for f, n, r in vs[:10]:
            acc = f
            for bit in bin(n)[3:]:
                g, b = S.run([acc], dupl=True)
                acc = g[0] if not b[0] else O.square(acc)
                if bit == "1":
                    g, b = S.run([acc], [f])
                    acc = g[0] if not b[0] else O.compose(acc, f)
            assert acc == r
            ran += 1
    assert ran >= 20


@pytest.mark.parametrize("name", ["t0", "node"])
def test_fast_random_bulk(name):
    t = load_golden("ref_%s.json" % name)
    S = FastSim(int(t["params"]["Delta"]))
    rng = random.Random(77)
    h, pk = I(t["params"]["h"]), I(t["pk"])
    pool = [O.power(h, rng.getrandbits(80) + 1) for _ in range(30)] + [O.power(pk, rng.getrandbits(80) + 1) for _ in range(30)]
    n = 3000
    f1 = [pool[rng.randrange(60)] for _ in range(n)]
    f2 = [pool[rng.randrange(60)] for _ in range(n)]
    memo = {}

    def comp(x, y, inv=False):
        k = (x, y, inv)
        if k not in memo:
            memo[k] = O.compose_inv(x, y) if inv else O.compose(x, y)
        return memo[k]
    g, b = S.run(f1, f2); check(g, b, [comp(x, y) for x, y in zip(f1, f2)])
    distinct = [i for i in range(n) if f1[i] != f2[i]]
    assert sum(int(b[i]) for i in distinct) <= 2, "bail rate on random inputs too high: %d" % int(b.sum())
    g, b = S.run(f1, f2, inv=True); check(g, b, [comp(x, y, True) for x, y in zip(f1, f2)])
    g, b = S.run(pool, dupl=True); check(g, b, [O.square(x) for x in pool])
    assert int(b.sum()) == 0


@pytest.mark.parametrize("name,n,rounds", [("t0", 20000, 6), ("node", 4000, 3)])
def test_fast_vs_generic_chained(name, n, rounds):
    """Fast path against the robust path on chained random compositions (both compiled from the device headers)."""
    from test_hostsim_forms import Sim
    t = load_golden("ref_%s.json" % name)
    D = int(t["params"]["Delta"])
    S, Gn = FastSim(D), Sim(D)
    rng = random.Random(5)
    h, pk = I(t["params"]["h"]), I(t["pk"])
    ns = 1500
    seeds = Gn.pow([h, pk] * (ns // 2), [rng.getrandbits(48) + 1 for _ in range(ns)])
    cur = [seeds[rng.randrange(ns)] for i in range(n)]
    oth = [seeds[rng.randrange(ns)] for i in range(n)]
    total_bail = 0
    for r in range(rounds):
        perm = list(range(n))
        rng.shuffle(perm)
        oth = [cur[p] for p in perm] if r else oth
        for kind in ("comp", "inv", "dupl"):
            if kind == "dupl":
                want = Gn.square(cur)
                got, bail = S.run(cur, dupl=True)
            else:
                want = Gn.compose(cur, oth, inv=(kind == "inv"))
                got, bail = S.run(cur, oth, inv=(kind == "inv"))
            bad = [i for i in range(n) if not bail[i] and got[i] != want[i]]
            assert not bad, (kind, r, bad[:3], cur[bad[0]], oth[bad[0]])
            total_bail += int(bail.sum())
            if kind == "comp":
                nxt = want
        cur = nxt
    # self-compositions (same `a` in both operands) are the only expected bails
    assert total_bail < n * rounds * 3 * 0.002, total_bail
    print("bails", total_bail, "of", n * rounds * 3)


def test_fast_pow_kat_and_fixtures():
    """fr_pow (fixed 4-bit windows on the register path) against the reference's nupow KATs and fixtures."""
    kat = load_golden("bicycl_nupow_kat.json")["vectors"]
    by = {}
    for v in kat:
        f, n, r = I(v["f"]), int(v["n"]), I(v["r"])
        by.setdefault(O.disc(f), []).append((f, n, r))
    for name in ("tiny", "t0", "node"):
        t = load_golden("ref_%s.json" % name)
        by.setdefault(int(t["params"]["Delta"]), []).extend((I(v["f"]), abs(int(v["n"])), None) for v in t["nupow"])
    ran = 0
    for D, vs in by.items():
        S = FastSim(D)
        if not S.ok():
            continue
        n = len(vs)
        el = max(1, max((e.bit_length() + 31) // 32 for _, e, _ in vs))
        inp = L.forms_to_limbs([v[0] for v in vs], S.na, S.nc)
        ex = L.ints_to_limbs([v[1] for v in vs], el)
        out = np.zeros_like(inp)
        bail = np.zeros(n, dtype=np.uint8)
        S.lib.hs_fast_pow(S.na, HS.P(S.absd), S.nd, S.dbits, S.nc, n, HS.P(inp), HS.P(out), HS.P(ex), el, 1, HS.P(bail))
        got = L.limbs_to_forms(out, S.na, S.nc)
        for i, (f, e, r) in enumerate(vs):
            if e == 0:
                assert bail[i]
                continue
            if not bail[i]:
                assert got[i] == (r if r is not None else O.power(f, e)), (D, i)
                ran += 1
        # shared exponent
        e = vs[0][1] or 5
        exs = L.ints_to_limbs([e], max(1, (e.bit_length() + 31) // 32))[:, 0].copy()
        S.lib.hs_fast_pow(S.na, HS.P(S.absd), S.nd, S.dbits, S.nc, n, HS.P(inp), HS.P(out), HS.P(exs), len(exs), 0, HS.P(bail))
        got = L.limbs_to_forms(out, S.na, S.nc)
        for i in range(min(n, 6)):
            if not bail[i]:
                assert got[i] == O.power(vs[i][0], e)
    assert ran >= 100


def test_fast_pow_shared_sliding_windows():
    """fr_pow_shared (what part_decrypt / decrypt / tensor x scalar run on the register path: sliding windows over the odd
    powers) against the oracle: every exponent below 70, powers of two and their neighbours, runs of ones, random shares."""
    t = load_golden("ref_t0.json")
    S = FastSim(int(t["params"]["Delta"]))
    assert S.ok()
    fs = [I(v["f"]) for v in t["nupow"]][:4] + [I(t["params"]["h"]), I(t["pk"])]
    n = len(fs)
    inp = L.forms_to_limbs(fs, S.na, S.nc)
    rng = random.Random(11)
    exps = list(range(1, 70)) + [1 << k for k in (7, 31, 32, 33, 63, 64, 65, 95, 96, 104)]
    exps += [(1 << k) - 1 for k in (8, 31, 32, 33, 64, 97)] + [(1 << k) + 1 for k in (31, 32, 63, 64)]
    exps += [0x11111111111111111111, 0x8000000080000000800000008, 0xF0F0F0F0F0F0F0F0F0F0F0F0F]
    exps += [rng.getrandbits(b) | (1 << (b - 1)) for b in (40, 64, 65, 103, 104, 105, 127, 128, 129) for _ in range(2)]
    for e in exps:
        el = max(1, (e.bit_length() + 31) // 32)
        for pad in (0, 1):  # leading zero limbs as the callers pass them
            exs = L.ints_to_limbs([e], el + pad)[:, 0].copy()
            out = np.zeros_like(inp)
            bail = np.zeros(n, dtype=np.uint8)
            S.lib.hs_fast_pow(S.na, HS.P(S.absd), S.nd, S.dbits, S.nc, n, HS.P(inp), HS.P(out), HS.P(exs), len(exs), 0, HS.P(bail))
            assert int(bail.sum()) == 0, e
            got = L.limbs_to_forms(out, S.na, S.nc)
            assert got == [O.power(f, e) for f in fs], hex(e)
    # e == 0: the neutral form does not fit the register shape -> declined, as on the fixed-window path
    exs = L.ints_to_limbs([0], 2)[:, 0].copy()
    bail = np.zeros(n, dtype=np.uint8)
    S.lib.hs_fast_pow(S.na, HS.P(S.absd), S.nd, S.dbits, S.nc, n, HS.P(inp), HS.P(np.zeros_like(inp)), HS.P(exs), 2, 0, HS.P(bail))
    assert int(bail.sum()) == n


def test_fast_flat_variant(tier):
    """the branch-free scalar Euclid rounds (hgcd32_nb; what the composition kernel runs on small batches): same accepted
    steps, same results as the reference fixtures"""
    S = FastSim(int(tier["params"]["Delta"]))
    if not S.ok():
        pytest.skip("no fast instantiation for %d limbs" % S.na)
    vs = tier["compose"] + tier["compose_edge"]
    f1, f2 = [I(v["f1"]) for v in vs], [I(v["f2"]) for v in vs]
    S.lib.hs_fast_flat(1)
    try:
        g, b = S.run(f1, f2); check(g, b, [I(v["r"]) for v in vs])
        assert int(b[:len(tier["compose"])].sum()) == 0
        g, b = S.run(f1, f2, inv=True); check(g, b, [I(v["rinv"]) for v in vs])
        g, b = S.run(f1, dupl=True); check(g, b, [I(v["sq"]) for v in vs])
        # a longer chain against the oracle
        rng = random.Random(4)
        fs = [f1[rng.randrange(len(tier["compose"]))] for _ in range(64)]
        gs = [f2[rng.randrange(len(tier["compose"]))] for _ in range(64)]
        for _ in range(6):
            out, bail = S.run(fs, gs)
            S.lib.hs_fast_flat(0)
            out0, bail0 = S.run(fs, gs)
            S.lib.hs_fast_flat(1)
            assert list(bail) == list(bail0)   # the same steps are accepted, so the same elements are declined
            want = [O.compose(x, y) for x, y in zip(fs, gs)]
            check(out, bail, want)
            fs = want
    finally:
        S.lib.hs_fast_flat(0)
