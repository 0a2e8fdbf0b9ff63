"""GPU parity of the cooperative (lane-group) kernels through the C ABI: the reference-generated fixtures with the
"coop" option forced, and equality with the register / generic path on large batches."""
import random

import pytest
import torch

from conftest import I, load_golden
from oracle import qf_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def groups():
    from cofhe_b200.group import ClassGroupB200
    cache = {}

    def get(D):
        if D not in cache:
            cache[D] = ClassGroupB200(D)
        return cache[D]
    yield get
    for g in cache.values():
        g.close()


def coop_group(groups, tier):
    G = groups(int(tier["params"]["Delta"]))
    if not int(G.lib.cofhe_b200_coop_limbs(G.h)):
        pytest.skip("no cooperative instantiation for this width")
    return G


def test_coop_compose_fixtures(groups, tier):
    G = coop_group(groups, tier)
    G.set_option("coop", 1)
    try:
        vs = tier["compose"] + tier["compose_edge"]
        f1, f2 = G.forms([I(v["f1"]) for v in vs]), G.forms([I(v["f2"]) for v in vs])
        assert G.forms_out(G.nucomp(f1, f2)) == [I(v["r"]) for v in vs]
        assert G.forms_out(G.nucomp(f2, f1)) == [I(v["r"]) for v in vs]
        assert G.forms_out(G.nucomp(f1, f2, negf2=True)) == [I(v["rinv"]) for v in vs]
        assert G.forms_out(G.nudupl(f1)) == [I(v["sq"]) for v in vs]
        a = tier["add_tensors"]
        assert G.cts_out(G.add_ciphertexts(G.cts(I(a["ct1"])), G.cts(I(a["ct2"])))) == list(I(a["res"]))
        G.status()
    finally:
        G.set_option("coop", -1)


@pytest.mark.parametrize("name", ["t0", "node", "t1"])
def test_coop_equals_other_path_at_scale(groups, name):
    t = load_golden("ref_%s.json" % name)
    G = coop_group(groups, t)
    rng = random.Random(11)
    h, pk = I(t["params"]["h"]), I(t["pk"])
    base = [O.power(h, rng.getrandbits(64)) for _ in range(8)] + [O.power(pk, rng.getrandbits(64)) for _ in range(8)]
    n = 1 << (15 if name != "t1" else 12)
    f = G.forms([base[i % 16] for i in range(n)])
    g = G.forms([base[(5 * i + 3) % 16] for i in range(n)])
    G.set_option("coop", 0)
    want = (G.nucomp(f, g), G.nucomp(f, g, negf2=True), G.nudupl(f))
    G.status()
    G.set_option("coop", 1)
    try:
        got = (G.nucomp(f, g), G.nucomp(f, g, negf2=True), G.nudupl(f))
        G.status()
        for a, b in zip(got, want):
            assert torch.equal(a, b)
        assert torch.equal(G.nucomp(got[0], g, negf2=True), f)
        G.status()
    finally:
        G.set_option("coop", -1)


def test_coop_lockstep_regression():
    """Eight 2053-bit pairs (one warp of the 4-lane kernel) on which the first warp-converged version fell out of step
    after the lane-divergent scalar Euclid (illegal address on the device; the pair list was found by bisection on a
    B200): expected results from the robust path of the same run (tests/golden/coop_case_t1.bin)."""
    import os
    import struct
    import numpy as np
    from conftest import GOLDEN
    from cofhe_b200.group import ClassGroupB200
    raw = open(os.path.join(GOLDEN, "coop_case_t1.bin"), "rb").read()
    dbits, nc, n = struct.unpack("III", raw[:12])
    t = load_golden("ref_t1.json")
    G = ClassGroupB200(int(t["params"]["Delta"]))
    assert G.nc == nc and (-int(t["params"]["Delta"])).bit_length() == dbits
    arr = np.frombuffer(raw[12:], dtype=np.uint32).reshape(3, G.fl, n).copy()
    dev = lambda a: torch.from_numpy(a.view(np.int32)).to(G.device)
    G.set_option("coop", 1)
    for reps in (1, 64):   # one warp, and the same warp many times over (several waves)
        f, g, want = (dev(np.tile(arr[k], (1, reps))) for k in range(3))
        got = G.nucomp(f, g)
        G.status()
        assert torch.equal(got, want)
    G.close()


@pytest.mark.parametrize("name", ["t0", "node", "t1"])
def test_coop_pow_equals_other_path(groups, name):
    """exponentiation on the cooperative kernels (shared and per-element exponents, negative, zero) against the register /
    generic path and the oracle"""
    t = load_golden("ref_%s.json" % name)
    G = coop_group(groups, t)
    rng = random.Random(21)
    h, pk = I(t["params"]["h"]), I(t["pk"])
    n = 1000 if name != "t1" else 300
    base = [O.power(h, rng.getrandbits(40) + 1) for _ in range(6)] + [pk, O.identity(int(t["params"]["Delta"]))]
    fs = [base[i % len(base)] for i in range(n)]
    f = G.forms(fs)
    sh = int(t["sk"])
    es = [rng.getrandbits(rng.choice([1, 5, 33, 64, 128])) * rng.choice([1, -1]) for _ in range(n)]
    es[3] = 0
    G.set_option("coop", 0)
    want = (G.nupow(f, shared=sh), G.nupow(f, shared=-sh), G.nupow(f, exps=es))
    G.status()
    G.set_option("coop", 1)
    try:
        got = (G.nupow(f, shared=sh), G.nupow(f, shared=-sh), G.nupow(f, exps=es))
        G.status()
        for a, b in zip(got, want):
            assert torch.equal(a, b)
        out = G.forms_out(got[2])
        for i in range(0, n, 97):
            assert out[i] == O.power(fs[i], es[i])
    finally:
        G.set_option("coop", -1)


@pytest.mark.parametrize("name", ["t0", "node", "t1"])
def test_coop_tables_matmul_and_fixed_base(name):
    """the multi-exponentiation from window tables on the cooperative kernels (k_coop_tabprod; the path of T1, forced on
    the narrower widths with the "coop" option set BEFORE set_bases): fixed base and encryption against the reference's
    fixtures, the scal_matmul fixtures, and a larger matmul (several chunks of the inner index, every window width)
    against the oracle"""
    from cofhe_b200.group import ClassGroupB200
    t = load_golden("ref_%s.json" % name)
    p = t["params"]
    G = ClassGroupB200(int(p["Delta"]))
    if not int(G.lib.cofhe_b200_coop_limbs(G.h)):
        G.close()
        pytest.skip("no cooperative instantiation for this width")
    try:
        G.set_option("coop", 1)
        G.set_scheme(int(p["k"]), int(p["N"]))
        G.set_bases(I(p["h"]), I(t["pk"]), int(p["exponent_bound"]).bit_length())
        vs = t["fixed_base"]
        assert G.cts_out(G.fixed_base_pow([int(v["r"]) for v in vs])) == [(I(v["h"]), I(v["pk"])) for v in vs]
        G.status()
        es = t["encrypt"]
        ct = G.encrypt([int(v["m"]) for v in es], [int(v["r"]) for v in es])
        assert G.cts_out(ct) == [I(v["ct"]) for v in es]
        assert G.ints_out(G.decrypt(ct, int(t["sk"]))) == [int(v["dec"]) for v in es]
        G.status()
        for c in t["scal_matmul"]:
            res = G.scal_matmul(G.cts(I(c["ct"])), [int(x) for x in c["s"]], c["n"], c["m"], c["p"], G.cts([I(c["zero"])]))
            assert G.cts_out(res) == list(I(c["res"])), c["label"]
        G.status()
        # a larger product: 2 x 6 . 6 x 5, scalars with zero digits and zero scalars, chunked inner index
        rng = random.Random(77)
        n, m, pp = 2, 6, 5
        cts = [t["add_tensors"]["ct1"][i % 4] if i % 2 else t["add_tensors"]["ct2"][i % 4] for i in range(n * m)]
        cts = [I(c) for c in cts]
        sc = [rng.getrandbits(rng.choice([3, 13, 17, 40])) for _ in range(m * pp)]
        sc[4] = 0
        sc[7] = 1 << 12
        zero = I(t["scal_matmul"][0]["zero"])
        want = O.scal_matmul(zero, cts, sc, n, m, pp)
        for w, ch in ((-1, -1), (2, 3), (3, 1), (5, 6)):
            G.set_option("mm_w", w)
            G.set_option("mm_chunks", ch)
            res = G.scal_matmul(G.cts(cts), sc, n, m, pp, G.cts([zero]))
            G.status()
            assert G.cts_out(res) == want, (w, ch)
    finally:
        G.close()
