"""GPU parity: batched class-group arithmetic through the C ABI (libcofhe_b200.so) against the reference's own
known-answer vectors, the reference-generated fixtures and the oracle.  Bit-exact (integer work)."""
import random

import numpy as np
import pytest

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


def test_bicycl_kat(groups):
    kat = load_golden("bicycl_nupow_kat.json")["vectors"]
    by = {}
    for v in kat:
        f, n, r = I(v["f"]), int(v["n"]), I(v["r"])
        by.setdefault(O.disc(f), []).append((f, n, r))
    total = 0
    for D, vs in by.items():
        G = groups(D)
        got = G.forms_out(G.nupow(G.forms([v[0] for v in vs]), exps=[v[1] for v in vs]))
        G.status()
        assert got == [v[2] for v in vs]
        total += len(vs)
    assert total == 150


def test_compose_fixtures(groups, tier):
    G = groups(int(tier["params"]["Delta"]))
    vs = tier["compose"] + tier["compose_edge"]
    f1, f2 = G.forms([I(v["f1"]) for v in vs]), G.forms([I(v["f2"]) for v in vs])
    assert G.forms_out(G.nucomp(f1, f2)) == [I(v["r"]) for v in vs]
    assert G.forms_out(G.nucomp(f2, f1)) == [I(v["r"]) for v in vs]
    assert G.forms_out(G.nucomp(f1, f2, negf2=True)) == [I(v["rinv"]) for v in vs]
    assert G.forms_out(G.nudupl(f1)) == [I(v["sq"]) for v in vs]
    G.status()


def test_nupow_fixtures(groups, tier):
    G = groups(int(tier["params"]["Delta"]))
    vs = tier["nupow"]
    f = G.forms([I(v["f"]) for v in vs])
    assert G.forms_out(G.nupow(f, exps=[int(v["n"]) for v in vs])) == [I(v["r"]) for v in vs]
    sk = int(tier["sk"])
    fs = [I(v["f"]) for v in vs[:3]]
    f3 = G.forms(fs)
    assert G.forms_out(G.nupow(f3, shared=sk)) == [O.power(x, sk) for x in fs]
    assert G.forms_out(G.nupow(f3, shared=-5)) == [O.power(x, -5) for x in fs]
    assert G.forms_out(G.nupow(f3, shared=0)) == [O.identity(G.D)] * 3
    G.status()


def test_reduce_unreduced(groups, tier):
    from cofhe_b200 import limbs as L
    from cofhe_b200.group import to_device
    G = groups(int(tier["params"]["Delta"]))
    vs = tier["power_of_f"]
    w = G.nd + 2
    inp = to_device(L.forms_to_limbs([I(v["f"]) for v in vs], w, w), G.device)
    assert G.forms_out(G.reduce(inp, w, w)) == [I(v["reduced"]) for v in vs]
    G.status()


def test_add_ciphertexts(groups, tier):
    G = groups(int(tier["params"]["Delta"]))
    a = tier["add_tensors"]
    c1, c2 = G.cts(I(a["ct1"])), G.cts(I(a["ct2"]))
    assert G.cts_out(G.add_ciphertexts(c1, c2)) == list(I(a["res"]))
    G.status()
    # host-pointer entry point (transfers inside the call)
    from cofhe_b200 import limbs as L
    h1, h2 = L.cts_to_limbs(I(a["ct1"]), G.na, G.nc), L.cts_to_limbs(I(a["ct2"]), G.na, G.nc)
    out = np.zeros_like(h1)
    G.add_ciphertexts_host(h1, h2, out)
    assert L.limbs_to_cts(out, G.na, G.nc) == list(I(a["res"]))
    # page-locked buffers: the kernels read / write host memory directly and do not read the c rows at all
    # (wiped here); the neutral form and a == c forms are among the operands
    import torch
    D = int(tier["params"]["Delta"])
    one = O.identity(D)
    cts1 = list(I(a["ct1"])) + [(one, I(a["ct1"])[0][1]), (I(a["ct1"])[0][0], one), (one, one)]
    cts2 = list(I(a["ct2"])) + [(I(a["ct2"])[0][0], one), (one, I(a["ct2"])[0][1]), (one, one)]
    want = [O.add_ciphertexts(x, y) for x, y in zip(cts1, cts2)]
    assert want[:len(a["res"])] == list(I(a["res"]))
    p1 = torch.from_numpy(L.cts_to_limbs(cts1, G.na, G.nc).view(np.int32)).pin_memory()
    p2 = torch.from_numpy(L.cts_to_limbs(cts2, G.na, G.nc).view(np.int32)).pin_memory()
    po = torch.zeros_like(p1).pin_memory()
    for wipe in (False, True):
        if wipe and G.reg:
            # also for the elements the register path declines (neutral operands): the robust redo recomputes c as well
            p1[:, 2 * G.na:, :] = 0x5a5a5a5a
            p2[:, 2 * G.na:, :] = 0x5a5a5a5a
        po.zero_()
        G.add_ciphertexts_host(p1.numpy().view(np.uint32), p2.numpy().view(np.uint32), po.numpy().view(np.uint32))
        assert L.limbs_to_cts(po.numpy().view(np.uint32), G.na, G.nc) == want


def test_ragged_and_empty(groups):
    t = load_golden("ref_t0.json")
    G = groups(int(t["params"]["Delta"]))
    rng = random.Random(3)
    h = I(t["params"]["h"])
    for n in (1, 31, 33, 129, 1000):
        fs = [O.power(h, rng.getrandbits(20) + 1) for _ in range(min(n, 40))]
        fs = (fs * (n // len(fs) + 1))[:n]
        g = fs[::-1]
        got = G.forms_out(G.nucomp(G.forms(fs), G.forms(g)))
        assert got[:40] == [O.compose(x, y) for x, y in zip(fs[:40], g[:40])]
        assert got[-1] == O.compose(fs[-1], g[-1])
    import torch
    e = torch.empty((G.fl, 0), dtype=torch.int32, device=G.device)
    assert G.nucomp(e, e).shape[-1] == 0
    G.status()


def test_bad_form_is_reported(groups):
    from cofhe_b200 import capi
    t = load_golden("ref_t0.json")
    G = groups(int(t["params"]["Delta"]))
    h = I(t["params"]["h"])
    bad = (h[0], h[1] + 2, h[2])  # wrong discriminant: an exact division fails
    G.nucomp(G.forms([bad]), G.forms([h]))
    with pytest.raises(capi.CofheError):
        G.status()
    G.status()  # cleared


def test_group_law_at_scale(groups):
    """Size-independent properties on a large batch: (f o g) o g^-1 == f, f^2 == f o f, commutativity."""
    t = load_golden("ref_t0.json")
    G = groups(int(t["params"]["Delta"]))
    rng = random.Random(9)
    h, pk = I(t["params"]["h"]), I(t["pk"])
    base = [O.power(h, rng.getrandbits(64)) for _ in range(16)] + [O.power(pk, rng.getrandbits(64)) for _ in range(16)]
    n = 1 << 15
    f = G.forms([base[i % 32] for i in range(n)])
    g = G.forms([base[(7 * i + 3) % 32] for i in range(n)])
    fg = G.nucomp(f, g)
    import torch
    assert torch.equal(G.nucomp(fg, g, negf2=True), f)
    assert torch.equal(G.nucomp(g, f), fg)
    assert torch.equal(G.nudupl(f), G.nucomp(f, f))
    G.status()


def test_scal_matmul_fixtures(groups, tier):
    """Matmul branch of scal_ciphertext_tensors against fixtures generated by the compiled reference."""
    G = groups(int(tier["params"]["Delta"]))
    for c in tier["scal_matmul"]:
        n, m, p = c["n"], c["m"], c["p"]
        ct, init = G.cts(I(c["ct"])), G.cts([I(c["zero"])])
        got = G.cts_out(G.scal_matmul(ct, [int(x) for x in c["s"]], n, m, p, init))
        G.status()
        assert got == list(I(c["res"])), c["label"]


@pytest.mark.parametrize("shape,bits", [((3, 5, 4), 13), ((2, 9, 33), 17), ((1, 4, 2), 128), ((4, 1, 3), 40)])
def test_scal_matmul_vs_oracle(groups, shape, bits):
    """Ragged shapes, zero scalars, chunked j ranges, wide scalars: GPU multi-exponentiation == oracle matmul."""
    import os
    t = load_golden("ref_t0.json")
    G = groups(int(t["params"]["Delta"]))
    n, m, p = shape
    rng = random.Random(n * 100 + m * 10 + p)
    h, pk = I(t["params"]["h"]), I(t["pk"])
    cts = [(O.power(h, rng.getrandbits(50) + 1), O.power(pk, rng.getrandbits(50) + 1)) for _ in range(n * m)]
    s = [rng.getrandbits(bits) if rng.random() > 0.15 else 0 for _ in range(m * p)]
    zero = I(t["scal_matmul"][0]["zero"])
    want = O.scal_matmul(zero, cts, s, n, m, p)
    for chunks in (1, 3, -1):
        G.set_option("mm_chunks", chunks)
        got = G.cts_out(G.scal_matmul(G.cts(cts), s, n, m, p, G.cts([zero])))
        G.status()
        assert got == want, chunks


def test_scal_matmul_degenerate(groups):
    """Identical ciphertexts in a row (the fast path declines self-compositions) and all-zero scalars."""
    t = load_golden("ref_t0.json")
    G = groups(int(t["params"]["Delta"]))
    h, pk = I(t["params"]["h"]), I(t["pk"])
    c = (O.power(h, 12345), O.power(pk, 12345))
    zero = I(t["scal_matmul"][0]["zero"])
    cts, s = [c] * 4, [3, 3, 3, 3, 0, 0, 0, 0]
    want = O.scal_matmul(zero, cts, s[:4], 1, 4, 1)
    assert G.cts_out(G.scal_matmul(G.cts(cts), s[:4], 1, 4, 1, G.cts([zero]))) == want
    assert G.cts_out(G.scal_matmul(G.cts(cts), s[4:], 1, 4, 1, G.cts([zero]))) == [zero]
    G.status()


def test_add_host_staged_ranges(groups, monkeypatch):
    """Pageable host buffers take the staged pipelines: one lane per form block (default) and element ranges over four
    lanes (COFHE_B200_HOST_CHUNKS); ragged sizes, results identical to the device-resident call."""
    from cofhe_b200 import limbs as L
    t = load_golden("ref_t0.json")
    G = groups(int(t["params"]["Delta"]))
    rng = random.Random(21)
    h = I(t["params"]["h"])
    base = [O.power(h, rng.getrandbits(50) + 1) for _ in range(20)]
    n = 2500
    c1 = [(base[rng.randrange(20)], base[rng.randrange(20)]) for _ in range(n)]
    c2 = [(base[rng.randrange(20)], base[rng.randrange(20)]) for _ in range(n)]
    want = G.cts_out(G.add_ciphertexts(G.cts(c1), G.cts(c2)))
    G.status()
    assert want[:50] == [O.add_ciphertexts(x, y) for x, y in zip(c1[:50], c2[:50])]
    h1, h2 = L.cts_to_limbs(c1, G.na, G.nc), L.cts_to_limbs(c2, G.na, G.nc)
    for chunks in (-1, 1, 3, 7):
        G.set_option("host_chunks", chunks)
        out = np.zeros_like(h1)
        G.add_ciphertexts_host(h1, h2, out)
        assert L.limbs_to_cts(out, G.na, G.nc) == want, chunks
    G.set_option("host_chunks", -1)
