"""Multi-wave check of the shared-exponent (sliding-window) exponentiation, DESIGN.md section 9 item 4: the same exponent
once as ONE shared exponent (fr_pow_shared / cq_pow's shared schedule) and once as an exponent per element (fixed
windows) on a batch of several waves; the two results must be identical.  python tools/pow_waves.py [tier] [n] [bits]"""
import os, random, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from conftest import I, load_golden
from cofhe_b200.group import ClassGroupB200

tier = sys.argv[1] if len(sys.argv) > 1 else "t0"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 18
t = load_golden("ref_%s.json" % tier)
bits = int(sys.argv[3]) if len(sys.argv) > 3 else int(t["sk"]).bit_length()
G = ClassGroupB200(int(t["params"]["Delta"]))
rng = random.Random(7)
nb = 1 << 12
seeds = G.forms([I(t["params"]["h"]), I(t["pk"])] * (nb // 2))
base = G.nupow(seeds, exps=[rng.getrandbits(40) | 1 for _ in range(nb)])
G.status()
f = base.repeat(1, n // nb).contiguous()
for k in range(3):  # three different exponents back to back, as the threshold decryption issues them
    e = rng.getrandbits(bits) | (1 << (bits - 1))
    a = G.nupow(f, shared=e)
    b = G.nupow(f, exps=[e] * n)
    G.status()
    print("%s n=%d %d-bit exponent %d: shared == per-element: %s" % (tier, n, bits, k, bool(torch.equal(a, b))))
G.close()
