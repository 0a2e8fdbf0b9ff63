"""Device timings of the shared-exponent exponentiation (part_decrypt) and encryption per tier: python tools/pow_time.py [tier ...]"""
import os, random, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from conftest import I, load_golden
from cofhe_b200.group import ClassGroupB200


def timeit(fn, reps=2):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    return best


for tier in (sys.argv[1:] or ["t1", "t0"]):
    t = load_golden("ref_%s.json" % tier)
    p = t["params"]
    G = ClassGroupB200(int(p["Delta"]))
    G.set_scheme(int(p["k"]), int(p["N"]))
    G.set_bases(I(p["h"]), I(t["pk"]), int(p["exponent_bound"]).bit_length())
    rng = random.Random(3)
    sk = int(t["sk"])
    for n in ([2048, 8192, 65536] if tier == "t0" else [2048, 8192]):
        ms = [rng.getrandbits(int(p["k"])) for _ in range(n)]
        ct = G.encrypt(ms, rng.randrange(1, int(p["exponent_bound"])))
        G.status()
        for mode in ((0, 1, 256, 4) if int(G.lib.cofhe_b200_coop_limbs(G.h)) else (0,)):   # 256: cooperative kernels in 256-thread blocks; 4: 128 threads, 4 blocks per SM
            G.set_option("coop", 1 if mode else 0)
            G.set_option("coop_pb", mode if mode in (256, 4) else 128)
            dt = timeit(lambda: G.part_decrypt(ct, sk))
            G.status()
            dd = timeit(lambda: G.decrypt(ct, sk), reps=1)
            ok = G.ints_out(G.decrypt(ct, sk)) == ms
            G.status()
            print("%s n=%6d coop=%d part_decrypt(%d-bit) %9.3f ms %9.1f k/s | decrypt %9.3f ms %9.1f k/s ok=%s" % (tier, n, mode, sk.bit_length(), dt * 1e3, n / dt / 1e3, dd * 1e3, n / dd / 1e3, ok))
        G.set_option("coop", -1)
        G.set_option("coop_pb", -1)
        for pb in (128, 256):
            G.set_option("coop_pb", pb)
            # encryption with own randomness per ciphertext (2 fixed-base exponentiations + f^m + 1 composition each)
            rs = [rng.randrange(1, int(p["exponent_bound"])) for _ in range(n)]
            rd, md = G.ints(rs, G.el_r), G.ints(ms, G.ml)
            de = timeit(lambda: G.encrypt(md, rd))
            G.status()
            ok = G.ints_out(G.decrypt(G.encrypt(md, rd), sk)) == ms
            G.status()
            print("%s n=%6d encrypt pb=%d (own r, %d-bit) %9.3f ms %9.1f k/s roundtrip=%s" % (tier, n, pb, int(p["exponent_bound"]).bit_length(), de * 1e3, n / de / 1e3, ok))
        G.set_option("coop_pb", -1)
    G.close()
