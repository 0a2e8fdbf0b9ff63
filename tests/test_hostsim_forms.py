"""Logic tests of the device arithmetic (cofhe_b200/csrc/*.cuh compiled with g++, tests/hostsim) against the oracle
and the reference-generated fixtures.  CPU only; the same checks run on the GPU through the C ABI in test_gpu_*.py."""
import random

import numpy as np
import pytest

from conftest import I, load_golden
from cofhe_b200 import limbs as L
from oracle import qf_oracle as O
import hostsim as HS


class Sim:
    def __init__(self, D):
        self.D = D
        self.absD = -D
        self.na, self.nc, self.nd = L.limb_shape(self.absD.bit_length())
        self.cap = HS.cap_for(self.nd)
        self.dl = L.ints_to_limbs([self.absD], self.nd)[:, 0].copy()
        self.lib = HS.load()

    def _out(self, n):
        return np.zeros((2 * self.na + self.nc, n), dtype=np.uint32)

    def compose(self, f1, f2, inv=False):
        n = len(f1)
        a, b, o = L.forms_to_limbs(f1, self.na, self.nc), L.forms_to_limbs(f2, self.na, self.nc), self._out(n)
        rc = self.lib.hs_compose(self.cap, HS.P(self.dl), self.nd, self.na, self.nc, n, HS.P(a), HS.P(b), HS.P(o), int(inv))
        assert rc == 0, rc
        return L.limbs_to_forms(o, self.na, self.nc)

    def square(self, f):
        n = len(f)
        a, o = L.forms_to_limbs(f, self.na, self.nc), self._out(n)
        rc = self.lib.hs_square(self.cap, HS.P(self.dl), self.nd, self.na, self.nc, n, HS.P(a), HS.P(o))
        assert rc == 0, rc
        return L.limbs_to_forms(o, self.na, self.nc)

    def reduce(self, f, na=None, nc=None):
        # unreduced inputs may need wider limbs: use nd for everything
        n = len(f)
        w = self.nd + 2
        a, o = L.forms_to_limbs(f, w, w), self._out(n)
        rc = self.lib.hs_reduce(self.cap, HS.P(self.dl), self.nd, w, w, self.na, self.nc, n, HS.P(a), HS.P(o))
        assert rc == 0, rc
        return L.limbs_to_forms(o, self.na, self.nc)

    def pow(self, f, exps):
        n = len(f)
        el = max(1, max((abs(e).bit_length() + 31) // 32 for e in exps))
        ex = L.ints_to_limbs([abs(e) for e in exps], el)
        sg = np.array([1 if e < 0 else 0 for e in exps], dtype=np.uint8)
        a, o = L.forms_to_limbs(f, self.na, self.nc), self._out(n)
        rc = self.lib.hs_pow(self.cap, HS.P(self.dl), self.nd, self.na, self.nc, n, HS.P(a), HS.P(o), HS.P(ex), el, 1, HS.P(sg), 0)
        assert rc == 0, rc
        return L.limbs_to_forms(o, self.na, self.nc)

    def pow_shared(self, f, e):
        n = len(f)
        el = max(1, (abs(e).bit_length() + 31) // 32)
        ex = L.ints_to_limbs([abs(e)], el)[:, 0].copy()
        a, o = L.forms_to_limbs(f, self.na, self.nc), self._out(n)
        rc = self.lib.hs_pow(self.cap, HS.P(self.dl), self.nd, self.na, self.nc, n, HS.P(a), HS.P(o), HS.P(ex), el, 0, None, int(e < 0))
        assert rc == 0, rc
        return L.limbs_to_forms(o, self.na, self.nc)


def test_bicycl_kat_hostsim():
    kat = load_golden("bicycl_nupow_kat.json")["vectors"]
    groups = {}
    for v in kat:
        f, n, r = I(v["f"]), int(v["n"]), I(v["r"])
        groups.setdefault(O.disc(f), []).append((f, n, r))
    assert sum(len(g) for g in groups.values()) == 150
    for D, vs in groups.items():
        S = Sim(D)
        got = S.pow([v[0] for v in vs], [v[1] for v in vs])
        assert got == [v[2] for v in vs]


def test_compose_fixtures(tier):
    S = Sim(int(tier["params"]["Delta"]))
    vs = tier["compose"] + tier["compose_edge"]
    f1, f2 = [I(v["f1"]) for v in vs], [I(v["f2"]) for v in vs]
    assert S.compose(f1, f2) == [I(v["r"]) for v in vs]
    assert S.compose(f1, f2, inv=True) == [I(v["rinv"]) for v in vs]
    assert S.square(f1) == [I(v["sq"]) for v in vs]
    # composition is commutative and must not care about operand order
    assert S.compose(f2, f1) == [I(v["r"]) for v in vs]


def test_nupow_fixtures(tier):
    S = Sim(int(tier["params"]["Delta"]))
    vs = tier["nupow"]
    assert S.pow([I(v["f"]) for v in vs], [int(v["n"]) for v in vs]) == [I(v["r"]) for v in vs]
    sk = int(tier["sk"])
    fs = [I(v["f"]) for v in vs[:3]]
    assert S.pow_shared(fs, sk) == [O.power(f, sk) for f in fs]
    assert S.pow_shared(fs, -5) == [O.power(f, -5) for f in fs]
    assert S.pow_shared(fs, 0) == [O.identity(S.D)] * 3


def test_reduce_unreduced(tier):
    S = Sim(int(tier["params"]["Delta"]))
    vs = tier["power_of_f"]
    assert S.reduce([I(v["f"]) for v in vs]) == [I(v["reduced"]) for v in vs]


def test_random_vs_oracle():
    rng = random.Random(11)
    t = load_golden("ref_t0.json")
    S = Sim(int(t["params"]["Delta"]))
    h = I(t["params"]["h"])
    pk = I(t["pk"])
    fs = [O.power(h, rng.getrandbits(100)) for _ in range(40)] + [O.power(pk, rng.getrandbits(90)) for _ in range(40)]
    # small/structured forms: powers of 2 in a, tiny a, identity
    fs += [O.identity(S.D), O.power(h, 1), O.power(h, 2)]
    g = fs[::-1]
    assert S.compose(fs, g) == [O.compose(x, y) for x, y in zip(fs, g)]
    assert S.compose(fs, g, inv=True) == [O.compose_inv(x, y) for x, y in zip(fs, g)]
    assert S.square(fs) == [O.square(x) for x in fs]
