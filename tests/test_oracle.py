"""Pins oracle/qf_oracle.py (the CPU restatement) against
 (1) the reference's own 150 nupow known-answer vectors, and
 (2) fixtures generated from the compiled reference (tools/gen_golden.py), for every hot-path operation,
 (3) live calls into oracle/_ref when it is present (build container only)."""
import random

import pytest

from conftest import I, load_golden
from oracle import qf_oracle as O


def params_of(t):
    p = t["params"]
    return O.Params(int(p["N"]), int(p["k"]), int(p["exponent_bound"]), I(p["h"]), int(p["d"]), int(p["e"]))


def test_bicycl_nupow_kat():
    kat = load_golden("bicycl_nupow_kat.json")["vectors"]
    assert len(kat) == 150
    for v in kat:
        f, n, r = I(v["f"]), int(v["n"]), I(v["r"])
        assert O.is_reduced(f) and O.is_reduced(r)
        assert O.power(f, n) == r


def test_params(tier):
    P = params_of(tier)
    p = tier["params"]
    assert P.Delta == int(p["Delta"]) and P.DeltaK == int(p["DeltaK"])
    assert P.large_message == p["large_message"]
    assert O.nucomp_bound(P.Delta) == int(p["L"])
    assert O.identity(P.Delta) == I(tier["one"])
    assert O.power(P.h, 1 << P.e) == I(p["h_e"])
    assert O.power(P.h, 1 << P.d) == I(p["h_d"])
    assert O.power(P.h, 1 << (P.d + P.e)) == I(p["h_de"])
    assert O.power_of_h(P, int(tier["sk"])) == I(tier["pk"])


def test_compose(tier):
    for v in tier["compose"] + tier["compose_edge"]:
        f1, f2 = I(v["f1"]), I(v["f2"])
        assert O.compose(f1, f2) == I(v["r"])
        assert O.compose_inv(f1, f2) == I(v["rinv"])
        assert O.square(f1) == I(v["sq"])


def test_nupow(tier):
    for v in tier["nupow"]:
        assert O.power(I(v["f"]), int(v["n"])) == I(v["r"])


def test_fixed_base(tier):
    P = params_of(tier)
    pk = I(tier["pk"])
    for v in tier["fixed_base"]:
        assert O.power_of_h(P, int(v["r"])) == I(v["h"])
        assert O.power(pk, int(v["r"])) == I(v["pk"])


def test_power_of_f_and_dlog(tier):
    P = params_of(tier)
    for v in tier["power_of_f"]:
        m = int(v["m"])
        f = O.power_of_f(P, m)
        assert f == I(v["f"])
        red = O.reduce_form(f)
        assert red == I(v["reduced"])
        assert O.dlog_in_F(P, red) == int(v["dlog"]) == m % (1 << P.k)


def test_encrypt_decrypt(tier):
    P = params_of(tier)
    pk, sk = I(tier["pk"]), int(tier["sk"])
    for v in tier["encrypt"]:
        ct = O.encrypt(P, pk, int(v["m"]), int(v["r"]))
        assert ct == I(v["ct"])
        assert O.decrypt(P, sk, ct) == int(v["dec"])


def test_tensor_ops(tier):
    a = tier["add_tensors"]
    for c1, c2, r in zip(I(a["ct1"]), I(a["ct2"]), I(a["res"])):
        assert O.add_ciphertexts(c1, c2) == r
    P = params_of(tier)
    ng = tier["negate_tensor"]
    for c, r in zip(I(ng["ct"]), I(ng["res"])):
        # the exponent is make_plaintext(-1) from the reference's float codec (2^k - 1 when mpf precision allows)
        assert O.scal_ciphertext(c, int(ng["s"])) == r
    if P.k <= 128:
        assert int(ng["s"]) == (1 << P.k) - 1


def test_scal_matmul(tier):
    for v in tier["scal_matmul"]:
        res = O.scal_matmul(I(v["zero"]), I(v["ct"]), I(v["s"]), v["n"], v["m"], v["p"])
        assert tuple(res) == I(v["res"])


def test_threshold(tier):
    P = params_of(tier)
    shares = I(tier["shares"])
    for v in tier["threshold"]:
        ct = I(v["ct"])
        ds = [O.part_decrypt(shares[p][0], ct) for p in (0, 1, 2)]
        assert tuple(ds) == I(v["ds"])
        assert O.combine(P, ct, ds) == int(v["m"])


def test_live_reference_random():
    """Random cross-check against the compiled reference (skipped where oracle/_ref is absent)."""
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    R = ref.Reference(128, 128)
    p = R.params()
    P = O.Params.from_dict(p)
    rng = random.Random(5)
    fs = [tuple(R.call("power_of_h", rng.randrange(p["exponent_bound"]))) for _ in range(6)]
    for i in range(6):
        f1, f2 = fs[i], fs[(i + 1) % 6]
        assert O.compose(f1, f2) == tuple(R.call("nucomp", f1, f2))
        e = rng.getrandbits(100) - (1 << 99)
        assert O.power(f1, e) == tuple(R.call("nupow", f1, e))
        m = rng.getrandbits(128)
        assert O.power_of_f(P, m) == tuple(R.call("power_of_f", m))


def test_wire_format_matches_reference_bytes():
    """oracle restatement of the binary tensor format vs bytes written by the compiled reference (tests/golden/ref_wire.json,
    made by tools/gen_golden_wire.py)."""
    w = load_golden("ref_wire.json")
    for tier, t in w.items():
        cts = [tuple(tuple(int(x) for x in f) for f in ct) for ct in t["cts"]]
        data = bytes.fromhex(t["hex"])
        assert O.serialize_ciphertext_tensor(t["shape"], cts) == data, tier
        assert O.deserialize_ciphertext_tensor(data) == (t["shape"], cts), tier


def test_big_fixture_heads():
    """tests/golden/big_t0.npz (outputs of the compiled reference at the benched sizes) against the restatement on
    the leading elements: one matmul output column of the sweep shape, the first encryptions / partial decryptions"""
    import big_inputs as B
    from cofhe_b200 import limbs as L
    t = load_golden("ref_t0.json")
    P = params_of(t)
    z = B.load()
    na, nc, _ = L.limb_shape((-P.Delta).bit_length())
    n, m, p = B.SWEEP_SHAPE
    cts = L.limbs_to_cts(z["sweep_ct"], na, nc)
    zero = L.limbs_to_cts(z["mm_zero"], na, nc)[0]
    sc = B.sweep_scalars()
    want = L.limbs_to_cts(z["sweep_res"], na, nc)
    got = O.scal_matmul(zero, cts[:m], [sc[j * p + 1] for j in range(m)], 1, m, 1)
    assert got[0] == want[1]
    ms, rs = B.enc_inputs(P.k, P.exponent_bound)
    pk = I(t["pk"])
    heads = L.limbs_to_cts(z["enc_head"], na, nc)
    for i in range(2):
        ct = O.encrypt(P, pk, ms[i], rs[i])
        assert ct == heads[i]
        for q in range(2):
            assert O.part_decrypt(int(t["shares"][q][0]), ct) == L.limbs_to_forms(z["pd_head_%d" % q], na, nc)[i]


def test_wire_forms_plaintexts_match_reference_bytes():
    """the shared layout for part-decryption-result and plaintext tensors vs bytes written by the compiled reference"""
    w = load_golden("ref_wire.json")
    for name, t in w.items():
        forms = [int(x) for f in t["forms"] for x in f]
        assert O.serialize_int_tensor(t["shape"], forms) == bytes.fromhex(t["pdr_hex"]), name
        assert O.deserialize_int_tensor(bytes.fromhex(t["pdr_hex"]), 3) == (t["shape"], forms)
        pts = [int(x) for x in t["pts"]]
        assert O.serialize_int_tensor(t["pt_shape"], pts) == bytes.fromhex(t["pt_hex"]), name
        assert O.deserialize_int_tensor(bytes.fromhex(t["pt_hex"]), 1) == (t["pt_shape"], pts)
