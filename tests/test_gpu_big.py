"""GPU parity at the configurations bench.py times (VERDICT r1 "Pin parity at the benched configurations"):
scal_matmul at p = 256 with the window / chunk / block plan of the benched shard, a sweep over every window width,
chunk count and block size, and BASELINE.json configs 3 / 4 at 2^16 elements.  Expected values come from the compiled
reference (tools/gen_golden_big.py -> tests/golden/big_t0.npz); inputs are regenerated from fixed seeds
(tests/big_inputs.py).  Bit-exact."""
import numpy as np
import pytest
import torch

import big_inputs as B
from conftest import I, load_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    from cofhe_b200.group import ClassGroupB200
    t = load_golden("ref_t0.json")
    G = ClassGroupB200(int(t["params"]["Delta"]))
    G.set_scheme(int(t["params"]["k"]), int(t["params"]["N"]))
    G.set_bases(I(t["params"]["h"]), I(t["pk"]), int(t["params"]["exponent_bound"]).bit_length())
    yield G, t, B.load()
    G.close()


def dev(G, a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int32)).to(G.device)


def host(t):
    return t.cpu().numpy().view(np.uint32)


# (window bits, chunks, block): automatic for this shape; the plan the benched 32-row shard gets (profiles/r1n_launches.csv:
# W = 6, 512-thread blocks, 9 chunks); other corners
PLANS = [(-1, -1, -1), (6, 9, 512), (5, 9, 512), (4, 1, 128), (8, 64, 256)]


@pytest.mark.parametrize("scal", ["pattern", "u128"])
def test_matmul_bench_shape(env, scal):
    """2 x 256 . 256 x 256 (two ct rows of config 5) against CPUCryptoSystem::scal_ciphertext_tensors' output"""
    G, t, z = env
    n, m, p = B.MM_SHAPE
    sc = B.mm_scalars_pattern() if scal == "pattern" else B.mm_scalars_u128()
    assert int(z["mm_ref_agrees_" + scal][0]) == 1  # the reference's tensor op equals the exact product on this input
    ct, zero, want = dev(G, z["mm_ct"]), dev(G, z["mm_zero"]), z["mm_res_" + scal]
    scd = G.ints(sc, (max(x.bit_length() for x in sc) + 31) // 32)
    for W, ch, blk in PLANS:
        G.set_option("mm_w", W); G.set_option("mm_chunks", ch); G.set_option("mm_block", blk)
        got = host(G.scal_matmul(ct, scd, n, m, p, zero, ebits=max(x.bit_length() for x in sc)))
        G.status()
        assert (got == want).all(), (scal, W, ch, blk)
    for o in ("mm_w", "mm_chunks", "mm_block"):
        G.set_option(o, -1)


def test_matmul_plan_sweep(env):
    """every window width x chunk count x block size on 2 x 16 . 16 x 256 with 0 / 40..128-bit scalars: digits cross limb
    boundaries for every W that does not divide 32"""
    G, t, z = env
    n, m, p = B.SWEEP_SHAPE
    sc = B.sweep_scalars()
    ct, zero, want = dev(G, z["sweep_ct"]), dev(G, z["mm_zero"]), z["sweep_res"]
    scd = G.ints(sc, 4)
    bad = []
    for W in range(2, 9):
        for ch in (1, 3, -1):
            for blk in (128, 256, 512):
                G.set_option("mm_w", W); G.set_option("mm_chunks", ch); G.set_option("mm_block", blk)
                got = host(G.scal_matmul(ct, scd, n, m, p, zero, ebits=128))
                G.status()
                if not (got == want).all():
                    bad.append((W, ch, blk))
    for o in ("mm_w", "mm_chunks", "mm_block"):
        G.set_option(o, -1)
    assert not bad, bad


def test_config3_config4_full_size(env):
    """2^16 per-element-r encryptions, three part_decrypt_tensor passes and combine: SHA-256 of every output buffer
    equals the compiled reference's, and the plaintexts come back"""
    G, t, z = env
    k, bound = int(t["params"]["k"]), int(t["params"]["exponent_bound"])
    ms, rs = B.enc_inputs(k, bound)
    md, rd = G.ints(ms, G.ml), G.ints(rs, G.el_r)
    ct = G.encrypt(md, rd)
    G.status()
    h = host(ct)
    assert (h[:, :, :8] == z["enc_head"]).all()
    assert B.sha(h) == B.stored_sha(z, "enc_sha")
    ds = []
    for q in range(3):
        d = G.part_decrypt(ct, int(t["shares"][q][0]))
        G.status()
        hd = host(d)
        assert (hd[:, :8] == z["pd_head_%d" % q]).all(), q
        assert B.sha(hd) == B.stored_sha(z, "pd_sha_%d" % q), q
        ds.append(d)
    m = G.combine(ct, torch.stack(ds))
    G.status()
    assert torch.equal(m, md)
    assert B.sha(host(m)) == B.stored_sha(z, "comb_sha")
    # decrypt_tensor at the same size (c1^sk, nucompinv, dlog)
    assert torch.equal(G.decrypt(ct, int(t["sk"])), md)
    G.status()
