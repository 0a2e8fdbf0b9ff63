"""GPU parity of the binary tensor wire format kernels against bytes written by the compiled
reference (tests/golden/ref_wire.json) and, at size, against the oracle restatement and the round trip."""
import random

import numpy as np
import pytest

from conftest import I, load_golden
from oracle import qf_oracle as O

pytestmark = pytest.mark.gpu
TIERS = ["tiny", "t0", "node", "t1"]


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


@pytest.mark.parametrize("name", TIERS)
def test_wire_golden(groups, name):
    t = load_golden("ref_wire.json")[name]
    G = groups(int(t["Delta"]))
    cts = [tuple(tuple(int(x) for x in f) for f in ct) for ct in t["cts"]]
    data = bytes.fromhex(t["hex"])
    # reference bytes -> device batch
    shape, dev = G.wire_to_cts(data)
    G.status()
    assert shape == t["shape"]
    assert G.cts_out(dev) == cts
    # device batch -> bytes identical to the reference's
    assert G.cts_to_wire(t["shape"], G.cts(cts)) == data


def test_wire_roundtrip_at_size(groups):
    t = load_golden("ref_t0.json")
    G = groups(int(t["params"]["Delta"]))
    rng = random.Random(11)
    h = I(t["params"]["h"])
    base = [O.power(h, rng.getrandbits(60) + 1) for _ in range(24)] + [O.identity(int(t["params"]["Delta"]))]
    n = 3 * 1111  # not a multiple of the block sizes; several scan blocks (6n > 1024)
    cts = [(base[rng.randrange(len(base))], base[rng.randrange(len(base))]) for _ in range(n)]
    want = O.serialize_ciphertext_tensor([3, 1111], cts)
    dev = G.cts(cts)
    got = G.cts_to_wire([3, 1111], dev)
    assert got == want
    shape, back = G.wire_to_cts(got)
    G.status()
    assert shape == [3, 1111]
    assert np.array_equal(back.cpu().numpy(), dev.cpu().numpy())
    # feeding the result of a device operation straight into the serialiser
    s = G.add_ciphertexts(dev, dev)
    G.status()
    assert O.deserialize_ciphertext_tensor(G.cts_to_wire([n], s))[1] == [O.add_ciphertexts(c, c) for c in cts[:n]]


def test_wire_malformed(groups):
    from cofhe_b200 import capi
    t = load_golden("ref_wire.json")["t0"]
    G = groups(int(t["Delta"]))
    data = bytearray(bytes.fromhex(t["hex"]))
    # truncated buffer: header says 6 ciphertexts but the table is cut
    with pytest.raises(capi.CofheError):
        G.wire_to_cts(bytes(data[:40]))
    # an offset pointing past the end
    bad = bytearray(data)
    bad[4 + 8 + 8 * 3:4 + 8 + 8 * 3 + 8] = (1 << 40).to_bytes(8, "little")
    G.wire_to_cts(bytes(bad))
    with pytest.raises(capi.CofheError):
        G.status()
    # empty tensor
    shape, dev = G.wire_to_cts(O.serialize_ciphertext_tensor([0], []))
    assert shape == [0] and dev.shape[-1] == 0
    assert G.cts_to_wire([0], dev) == O.serialize_ciphertext_tensor([0], [])


@pytest.mark.parametrize("name", TIERS)
def test_wire_forms_and_plaintexts_golden(groups, name):
    """the other two tensor formats (part-decryption results = forms; plaintexts of any sign and size) against bytes
    written by the compiled reference"""
    import torch
    from cofhe_b200 import limbs as L
    t = load_golden("ref_wire.json")[name]
    G = groups(int(t["Delta"]))
    forms = [tuple(int(x) for x in f) for f in t["forms"]]
    data = bytes.fromhex(t["pdr_hex"])
    shape, dev = G.wire_to_forms(data)
    G.status()
    assert shape == t["shape"] and G.forms_out(dev) == forms
    assert G.forms_to_wire(t["shape"], G.forms(forms)) == data
    pts = [int(x) for x in t["pts"]]
    pdata = bytes.fromhex(t["pt_hex"])
    ml = (max(abs(x).bit_length() for x in pts) + 31) // 32
    shape, mags, sg = G.wire_to_plaintexts(pdata, ml)
    G.status()
    assert shape == t["pt_shape"]
    got = [(-m if s else m) for m, s in zip(L.limbs_to_ints(mags.cpu().numpy().view(np.uint32)), sg.cpu().tolist())]
    assert got == pts
    md = G.ints([abs(x) for x in pts], ml)
    sd = torch.tensor([1 if x < 0 else 0 for x in pts], dtype=torch.uint8, device=G.device)
    assert G.plaintexts_to_wire(t["pt_shape"], md, sd) == pdata
    # a value wider than the caller's limbs raises the element status
    if ml > 1:
        from cofhe_b200 import capi
        G.wire_to_plaintexts(pdata, 1)
        with pytest.raises(capi.CofheError):
            G.status()


def test_wire_forms_plaintexts_at_size(groups):
    import torch
    t = load_golden("ref_t0.json")
    G = groups(int(t["params"]["Delta"]))
    rng = random.Random(5)
    h = I(t["params"]["h"])
    base = [O.power(h, rng.getrandbits(60) + 1) for _ in range(16)] + [O.identity(int(t["params"]["Delta"]))]
    n = 2 * 1777
    forms = [base[rng.randrange(len(base))] for _ in range(n)]
    want = O.serialize_int_tensor([2, 1777], [x for f in forms for x in f])
    assert G.forms_to_wire([2, 1777], G.forms(forms)) == want
    shape, back = G.wire_to_forms(want)
    G.status()
    assert shape == [2, 1777] and G.forms_out(back) == forms
    pts = [rng.getrandbits(rng.choice([1, 8, 31, 32, 33, 64, 127, 128])) * rng.choice([1, 1, -1]) for _ in range(n)]
    wantp = O.serialize_int_tensor([n], pts)
    md = G.ints([abs(x) for x in pts], 4)
    sd = torch.tensor([1 if x < 0 else 0 for x in pts], dtype=torch.uint8, device=G.device)
    assert G.plaintexts_to_wire([n], md, sd) == wantp
    shape, mags, sg = G.wire_to_plaintexts(wantp, 4)
    G.status()
    assert torch.equal(mags, md) and torch.equal(sg.cpu(), torch.tensor([1 if x < 0 else 0 for x in pts], dtype=torch.uint8))
