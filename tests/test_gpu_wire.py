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
