"""Saturated shared-exponent exponentiation at T1 by block configuration of the cooperative kernel: python tools/pow_sat.py [n]"""
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


n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
t = load_golden("ref_t1.json")
p = t["params"]
G = ClassGroupB200(int(p["Delta"]))
G.set_scheme(int(p["k"]), int(p["N"]))
G.set_bases(I(p["h"]), I(t["pk"]), int(p["exponent_bound"]).bit_length())
rng = random.Random(3)
sk = int(t["sk"])
ms = [rng.getrandbits(int(p["k"])) for _ in range(n)]
ct = G.encrypt(ms, rng.randrange(1, int(p["exponent_bound"])))
G.status()
ref = None
for pb in (128, 4, 256, -1):
    G.set_option("coop_pb", pb)
    dt = timeit(lambda: G.part_decrypt(ct, sk), reps=1)
    out = G.part_decrypt(ct, sk)
    G.status()
    if ref is None: ref = out
    print("t1 n=%d coop_pb=%d part_decrypt %.1f ms %.1f k/s equal=%s" % (n, pb, dt * 1e3, n / dt / 1e3, bool(torch.equal(ref, out))))
G.close()
