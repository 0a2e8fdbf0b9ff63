#!/usr/bin/env python
"""Golden fixtures of the COMPACT variant of CL_HSM2k (CPUCryptoSystem(level, k, compact=true): h, pk and every c1 in
Cl(DeltaK)) from the compiled reference -> tests/golden/ref_compact.json.

TEST INFRASTRUCTURE.  Runs only in the build container (needs /root/reference and oracle/_ref/libcofhe_ref.so).
Per tier: parameters, keys, forms of Cl(DeltaK) with their images in Cl(Delta) (from_Cl_DeltaK_to_Cl_Delta), compositions
and powers in Cl(DeltaK), encryptions with explicit randomness, the batched tensor operations, threshold decryption.
All integers are decimal strings."""
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.ref import Reference  # noqa: E402
from tools.gen_golden import S, cts6  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def gen_tier(name, nbits, k, n_forms, n_enc, mat):
    rng = random.Random(4321 + nbits + k)
    R = Reference(nbits, k, compact=True)
    P = R.params()
    out = {"tier": name, "nbits": nbits, "k": k, "compact": True,
           "params": {kk: S(v) if not isinstance(v, bool) else v for kk, v in P.items()}}
    sk, *pk = R.call("keygen")
    out["sk"], out["pk"] = S(sk), S(pk)
    bound = P["exponent_bound"]
    one = tuple(R.call("one"))
    # forms of Cl(DeltaK): h^r, pk^r, small powers of h (a even / c even cases of prime_to_2exp show up among them), the identity
    elems = [tuple(R.call("power_of_h" if i % 2 == 0 else "pk_exp", rng.randrange(bound))) for i in range(n_forms)]
    elems += [tuple(R.call("nupow", P["h"], e)) for e in (1, 2, 3, 5, -7)] + [one]
    out["lift"] = [{"f": S(f), "lifted": S(R.call("lift", f))} for f in elems]
    comp = []
    for i in range(n_forms):
        f1, f2 = elems[i], elems[(i * 5 + 2) % len(elems)]
        e = rng.getrandbits([13, 64, k, bound.bit_length() - 1][i % 4]) | 1
        comp.append({"f1": S(f1), "f2": S(f2), "r": S(R.call("nucomp", f1, f2)), "rinv": S(R.call("nucompinv", f1, f2)),
                     "sq": S(R.call("nudupl", f1)), "n": S(e), "pow": S(R.call("nupow", f1, e))})
    out["compose"] = comp
    out["fixed_base"] = [{"r": S(r), "h": S(R.call("power_of_h", r)), "pk": S(R.call("pk_exp", r))}
                         for r in [0, 1, bound - 1] + [rng.randrange(bound) for _ in range(4)]]
    enc = []
    for i in range(n_enc):
        m = rng.getrandbits(k) if i else 0
        r = rng.randrange(bound)
        ct = R.call("encrypt_r", m, r)
        enc.append({"m": S(m), "r": S(r), "ct": S(cts6(ct)[0]), "dec": S(R.call("decrypt", ct)[0])})
        assert int(enc[-1]["dec"]) == m
    out["encrypt"] = enc
    # RandGen-driven calls in this order after keygen on a fresh object (draw schedule)
    R2 = Reference(nbits, k, compact=True)
    sk2, *_ = R2.call("keygen")
    assert sk2 == sk
    pts = [i + 1 for i in range(6)]
    out["schedule"] = {"encrypt_tensor": {"rows": 2, "cols": 3, "pt": S(pts), "ct": S(cts6(R2.call("encrypt_tensor", 2, 3, pts)))},
                       "encrypt_single": {"pt": "42", "ct": S(cts6(R2.call("encrypt", 42))[0])}}
    cts = [cts6(R.call("encrypt_r", (i + 1), rng.randrange(bound)))[0] for i in range(8)]
    a1, a2 = cts[:4], cts[4:]
    out["add_tensors"] = {"ct1": S(a1), "ct2": S(a2), "res": S(cts6(R.call("add_tensors", 2, 2, a1, a2)))}
    out["negate_tensor"] = {"ct": S(a1[:2]), "s": S(R.call("make_plaintext", -1)[0]), "res": S(cts6(R.call("negate_tensor", 2, a1[:2])))}
    n, m, p = mat
    ctm = [cts6(R.call("encrypt_r", (i + 1), rng.randrange(bound)))[0] for i in range(n * m)]
    for label, sc in (("small", [rng.randrange(1, 1 << 13) for _ in range(m * p)]),
                      ("u64", [rng.randrange(1, 1 << 64) for _ in range(m * p)])):
        R.call("set_seed", 777)
        zr = R.call("rand_mpz", bound)[0]
        R.call("set_seed", 777)
        res = cts6(R.call("scal_matmul", n, m, p, ctm, sc))
        zero = cts6(R.call("encrypt_r", 0, zr))[0]
        out.setdefault("scal_matmul", []).append({"label": label, "n": n, "m": m, "p": p, "ct": S(ctm), "s": S(sc),
                                                  "zero_r": S(zr), "zero": S(zero), "res": S(res)})
    sh = R.call("keygen_shares", 3, 5)
    npart, per = sh[0], sh[1]
    shares = [sh[2 + i * per: 2 + (i + 1) * per] for i in range(npart)]
    out["shares"] = S(shares)
    th = []
    for i in range(n_enc):
        ct = cts6(R.call("encrypt_r", enc[i]["m"], enc[i]["r"]))[0]
        ds = [tuple(R.call("part_decrypt", shares[q][0], ct)) for q in (0, 1, 2)]
        mrec = R.call("combine", ct, ds)[0]
        assert mrec == int(enc[i]["m"]) % (1 << k)
        th.append({"ct": S(ct), "ds": S(ds), "m": S(mrec)})
    out["threshold"] = th
    print(name, "DeltaK bits", abs(P["DeltaK"]).bit_length(), "Delta bits", abs(P["Delta"]).bit_length())
    return out


if __name__ == "__main__":
    tiers = {"t0": gen_tier("t0", 128, 128, n_forms=12, n_enc=4, mat=(2, 3, 2)),
             "tiny": gen_tier("tiny", 64, 16, n_forms=10, n_enc=3, mat=(2, 2, 2)),
             "n256k64": gen_tier("n256k64", 256, 64, n_forms=6, n_enc=2, mat=(1, 2, 2))}
    path = os.path.join(OUT, "ref_compact.json")
    json.dump(tiers, open(path, "w"), indent=0)
    print("->", path, os.path.getsize(path) // 1024, "KiB")
