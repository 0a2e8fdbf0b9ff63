#!/usr/bin/env python
"""Full-size fixtures for the configurations bench.py times (tests/golden/big_t0.npz), generated from the compiled
reference (oracle/_ref).  TEST INFRASTRUCTURE; runs only in the build container.

  mm_*     scalar x ciphertext matmul 2x256 . 256x256 (two ct rows of BASELINE.json config 5): inputs
           encrypt_tensor(pt[i] = i+1), `zero` = encrypt(0) with the recorded draw, scalars = benchmark pattern i+1 and
           uniform 128-bit; outputs of CPUCryptoSystem::scal_ciphertext_tensors, cross-checked against the product of
           ClassGroup::nupow / nucomp (the exact value)
  sweep_*  2x16 . 16x256 with mixed 0 / 40..128-bit scalars (window-width / chunk / block sweep of the multi-exponentiation)
  enc_*, pd_*, comb_*   config 3 / 4 at 2^16: SHA-256 of the reference's output buffers for inputs that
           tests/big_inputs.py regenerates from fixed seeds (only hashes and a few leading columns are stored)
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle.ref import Reference  # noqa: E402
from cofhe_b200 import limbs as L  # noqa: E402
import big_inputs as B  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint32).tobytes()).hexdigest()


def cts6(v):
    return [(tuple(v[i:i + 3]), tuple(v[i + 3:i + 6])) for i in range(0, len(v), 6)]


def main():
    t = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_t0.json")))
    nbits, k = t["nbits"], t["k"]
    D = int(t["params"]["Delta"])
    na, nc, _ = L.limb_shape((-D).bit_length())
    bound = int(t["params"]["exponent_bound"])
    R = Reference(nbits, k)
    sk, *pk = R.call("keygen")
    assert sk == int(t["sk"])
    out = {}
    # ---- matmul at the benched shape
    n, m, p = B.MM_SHAPE
    ct = L.cts_to_limbs(cts6(R.call("encrypt_tensor", n, m, [i + 1 for i in range(n * m)])), na, nc)
    R.call("set_seed", 777)
    zr = R.call("rand_mpz", bound)[0]
    zero = L.cts_to_limbs(cts6(R.call("encrypt_r", 0, zr)), na, nc)
    out["mm_ct"], out["mm_zero"] = ct, zero
    for name, sc in (("pattern", B.mm_scalars_pattern()), ("u128", B.mm_scalars_u128())):
        R.call("set_seed", 777)
        dt, res = R.scal_matmul_given(n, m, p, sc, ct, na, nc)
        dte, exact = R.scal_matmul_given(n, m, p, sc, ct, na, nc, exact_zero=zero)
        same = bool((res == exact).all())
        print("matmul %s: reference %.1f s, exact %.1f s, reference == exact: %s" % (name, dt, dte, same))
        out["mm_res_" + name] = exact
        out["mm_ref_agrees_" + name] = np.array([int(same)])
    # ---- window / chunk / block sweep shape
    n, m, p = B.SWEEP_SHAPE
    sct = L.cts_to_limbs(cts6(R.call("encrypt_tensor", n, m, [3 * i + 7 for i in range(n * m)])), na, nc)
    dte, exact = R.scal_matmul_given(n, m, p, B.sweep_scalars(), sct, na, nc, exact_zero=zero)
    out["sweep_ct"], out["sweep_res"] = sct, exact
    print("sweep shape exact: %.1f s" % dte)
    # ---- config 3 / 4 at 2^16
    cnt = B.BIG_N
    ms, rs = B.enc_inputs(k, bound, cnt)
    ml, rl = L.ints_to_limbs(ms, (k + 31) // 32), L.ints_to_limbs(rs, (bound.bit_length() + 31) // 32)
    dt, cts = R.encrypt_given(ml, rl, na, nc)
    print("encrypt 2^16: %.1f s" % dt)
    out["enc_sha"] = np.frombuffer(sha(cts).encode(), dtype=np.uint8)
    out["enc_head"] = cts[:, :, :8].copy()
    shares = [int(t["shares"][q][0]) for q in range(3)]
    ds = []
    for q, s in enumerate(shares):
        dt, d = R.part_decrypt_given(s, cts, na, nc)
        print("part_decrypt share %d: %.1f s" % (q, dt))
        ds.append(d)
        out["pd_sha_%d" % q] = np.frombuffer(sha(d).encode(), dtype=np.uint8)
        out["pd_head_%d" % q] = d[:, :8].copy()
    dt, mm = R.combine_given(cts, np.stack(ds), (k + 31) // 32, na, nc)
    print("combine: %.1f s" % dt)
    assert (mm == ml).all(), "threshold decryption of the reference does not return the plaintexts"
    out["comb_sha"] = np.frombuffer(sha(mm).encode(), dtype=np.uint8)
    path = os.path.join(ROOT, "tests", "golden", "big_t0.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
