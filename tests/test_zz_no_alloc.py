"""GPU test of cofhe_b200_add_ciphertexts_static, the entry point that allocates nothing (its scratch -- the list of
elements handed to the robust path and the counters -- is a fixed `static __device__` array inside the library; no
counterpart in the reference).  Checked through the C ABI against cofhe_b200_add_ciphertexts, bit for bit (integer
work), on two calls in a row: the array is shared and keeps the previous call's contents, so the second call runs on a
dirty list.  Collected last on purpose; it builds its own contexts and passes alone in
a fresh process as well as after the other tests."""
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


def operands(G, t):
    """the fixture's ciphertext pairs plus pairs with neutral forms (the hand-written kernels decline those: they go
    through the list in the fixed scratch and are redone by the robust kernel)"""
    a = t["add_tensors"]
    D = int(t["params"]["Delta"])
    one = O.identity(D)
    x, y = list(I(a["ct1"])), list(I(a["ct2"]))
    cts1 = x + [(one, x[0][1]), (x[0][0], one), (one, one)]
    cts2 = y + [(y[0][0], one), (one, y[0][1]), (one, one)]
    return cts1, cts2, list(I(a["res"]))


def test_static_add_fixtures(groups, tier):
    import torch
    G = groups(int(tier["params"]["Delta"]))
    cts1, cts2, res = operands(G, tier)
    c1, c2 = G.cts(cts1), G.cts(cts2)
    want = G.add_ciphertexts(c1, c2)
    G.status()
    assert G.cts_out(want)[:len(res)] == res
    for call in range(2):
        out = torch.full_like(c1, 0x5a5a5a5a)
        G.add_ciphertexts_static(c1, c2, out)
        G.status()
        assert torch.equal(out, want), call
    # a shorter batch straight after a longer one: what the longer call left in the list is not picked up
    k = 2
    out = torch.full_like(c1[:, :, :k], 0x5a5a5a5a).contiguous()
    G.add_ciphertexts_static(c2[:, :, -k:].contiguous(), c1[:, :, -k:].contiguous(), out)
    G.status()
    assert torch.equal(out, want[:, :, -k:])  # commutative, and the reduced form is unique
    # empty batch
    e = torch.empty((2, G.fl, 0), dtype=torch.int32, device=G.device)
    assert G.add_ciphertexts_static(e, e, e.clone()).shape[-1] == 0


@pytest.mark.parametrize("name", ["t0", "t1"])
def test_static_add_at_size_twice(groups, name):
    """register-resident (t0) and cooperative (t1) kernels at 2^15 ciphertexts with thousands of declined elements, two
    calls in a row with different operands, then the first again; device memory in use does not change across the calls"""
    import torch
    t = load_golden("ref_%s.json" % name)
    G = groups(int(t["params"]["Delta"]))
    cts1, cts2, _ = operands(G, t)
    m = len(cts1)
    c1, c2 = G.cts(cts1), G.cts(cts2)
    n = 1 << 15
    ar = torch.arange(n, device=G.device)
    a = c1[:, :, ar % m].contiguous()
    b = c2[:, :, (7 * ar + 3) % m].contiguous()
    c = c2[:, :, (5 * ar + 1) % m].contiguous()
    want_ab, want_ac = G.add_ciphertexts(a, b), G.add_ciphertexts(a, c)
    G.status()
    out = torch.empty_like(a)
    G.add_ciphertexts_static(a, b, out)  # first use on this device: kernels loaded, address of the array looked up
    G.status()
    assert torch.equal(out, want_ab)
    launches = G.launches()
    for x, y, want in ((a, c, want_ac), (a, b, want_ab), (a, b, want_ab)):
        out.fill_(0x5a5a5a5a)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(G.device)[0]  # read around the call only: torch's own ops may allocate
        G.add_ciphertexts_static(x, y, out)
        G.status()
        assert torch.cuda.mem_get_info(G.device)[0] >= free0, "device memory was allocated inside the call"
        assert torch.equal(out, want)
    assert G.launches() - launches == 3 * 2  # per call: the composition kernel and the robust redo (the memset is not ours)


def test_static_add_refuses_beyond_the_maximum(groups):
    import torch
    from cofhe_b200 import capi
    t = load_golden("ref_tiny.json")
    G = groups(int(t["params"]["Delta"]))
    n = capi.STATIC_MAX_N + 1
    z = torch.zeros((2, G.fl, n), dtype=torch.int32, device=G.device)
    launches = G.launches()
    with pytest.raises(capi.CofheError) as ei:
        G.add_ciphertexts_static(z, z, z)
    assert ei.value.code == capi.ETOOBIG
    assert G.launches() == launches  # nothing was launched
    G.status()
    # the maximum itself is accepted (operands: the neutral ciphertext, every element takes the list)
    one = O.identity(int(t["params"]["Delta"]))
    u = G.cts([(one, one)])
    full = u[:, :, torch.zeros(capi.STATIC_MAX_N, dtype=torch.long, device=G.device)].contiguous()
    out = torch.empty_like(full)
    G.add_ciphertexts_static(full, full, out)
    G.status()
    assert torch.equal(out, full)
