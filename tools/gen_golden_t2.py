#!/usr/bin/env python
"""Golden fixture of the largest parameter set exercised here, "t2": N of 3072 bits, k = 128 -> |Delta| = 3333 bits, 53-limb
coefficients (the width served by the 8-lane cooperative kernels) -> tests/golden/ref_t2.json.

TEST INFRASTRUCTURE.  Runs only in the build container (needs /root/reference and oracle/_ref/libcofhe_ref.so).  A reduced set
of what tools/gen_golden.py writes per tier (the file would be > 1 MB otherwise).  All integers are decimal strings."""
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.ref import Reference  # noqa: E402
from tools.gen_golden import S, cts6  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "ref_t2.json")

if __name__ == "__main__":
    nbits, k = 3072, 128
    rng = random.Random(9000 + nbits + k)
    R = Reference(nbits, k)
    P = R.params()
    out = {"tier": "t2", "nbits": nbits, "k": k, "params": {kk: S(v) if not isinstance(v, bool) else v for kk, v in P.items()}}
    sk, *pk = R.call("keygen")
    out["sk"], out["pk"] = S(sk), S(pk)
    bound = P["exponent_bound"]
    one = tuple(R.call("one"))
    elems = [tuple(R.call("power_of_h" if i % 2 == 0 else "pk_exp", rng.randrange(bound))) for i in range(6)]
    comp = []
    for i in range(6):
        f1, f2 = elems[i], elems[(i * 5 + 1) % 6]
        comp.append({"f1": S(f1), "f2": S(f2), "r": S(R.call("nucomp", f1, f2)), "rinv": S(R.call("nucompinv", f1, f2)),
                     "sq": S(R.call("nudupl", f1))})
    out["compose"] = comp
    f = elems[0]
    edge = [(one, f), (f, (f[0], -f[1], f[2])), (f, f)]
    out["compose_edge"] = [{"f1": S(a), "f2": S(b), "r": S(R.call("nucomp", a, b)), "rinv": S(R.call("nucompinv", a, b)),
                            "sq": S(R.call("nudupl", a))} for a, b in edge]
    pw = []
    for i, e in enumerate([0, 1, -1, (1 << k) - 1, rng.getrandbits(64), -rng.getrandbits(200), rng.getrandbits(bound.bit_length() - 1)]):
        pw.append({"f": S(elems[i % 6]), "n": S(e), "r": S(R.call("nupow", elems[i % 6], e))})
    out["nupow"] = pw
    out["fixed_base"] = [{"r": S(r), "h": S(R.call("power_of_h", r)), "pk": S(R.call("pk_exp", r))}
                         for r in [0, bound - 1, rng.randrange(bound)]]
    enc = []
    for i in range(3):
        m = rng.getrandbits(k) if i else 0
        r = rng.randrange(bound)
        ct = R.call("encrypt_r", m, r)
        enc.append({"m": S(m), "r": S(r), "ct": S(cts6(ct)[0]), "dec": S(R.call("decrypt", ct)[0])})
    out["encrypt"] = enc
    cts = [cts6(R.call("encrypt_r", (i + 1), rng.randrange(bound)))[0] for i in range(4)]
    out["add_tensors"] = {"ct1": S(cts[:2]), "ct2": S(cts[2:]), "res": S(cts6(R.call("add_tensors", 1, 2, cts[:2], cts[2:])))}
    out["negate_tensor"] = {"ct": S(cts[:1]), "s": S(R.call("make_plaintext", -1)[0]), "res": S(cts6(R.call("negate_tensor", 1, cts[:1])))}
    n, m, p = 1, 2, 2
    sc = [rng.randrange(1, 1 << 17) for _ in range(m * p)]
    R.call("set_seed", 777)
    zr = R.call("rand_mpz", bound)[0]
    R.call("set_seed", 777)
    res = cts6(R.call("scal_matmul", n, m, p, cts[:2], sc))
    zero = cts6(R.call("encrypt_r", 0, zr))[0]
    out["scal_matmul"] = [{"label": "17bit", "n": n, "m": m, "p": p, "ct": S(cts[:2]), "s": S(sc), "zero_r": S(zr), "zero": S(zero), "res": S(res)}]
    sh = R.call("keygen_shares", 3, 5)
    npart, per = sh[0], sh[1]
    shares = [sh[2 + i * per: 2 + (i + 1) * per] for i in range(npart)]
    out["shares"] = S(shares)
    th = []
    for i in range(2):
        ct = cts6(R.call("encrypt_r", enc[i]["m"], enc[i]["r"]))[0]
        ds = [tuple(R.call("part_decrypt", shares[q][0], ct)) for q in (0, 1, 2)]
        mrec = R.call("combine", ct, ds)[0]
        assert mrec == int(enc[i]["m"]) % (1 << k)
        th.append({"ct": S(ct), "ds": S(ds), "m": S(mrec)})
    out["threshold"] = th
    json.dump(out, open(OUT, "w"), indent=0)
    print("t2 Delta bits", abs(P["Delta"]).bit_length(), "->", OUT, os.path.getsize(OUT) // 1024, "KiB")
