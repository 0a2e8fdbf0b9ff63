"""One launch of the cooperative table kernel (fixed-base encryption) and of the cooperative exponentiation kernel (part_decrypt)
at T1 for profiling (ncu -k regex:k_coop_tabprod / k_coop_pow): python tools/coop_prof2.py [n_enc] [n_pdec]"""
import os, random, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from conftest import I, load_golden
from cofhe_b200.group import ClassGroupB200

ne = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
npd = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
t = load_golden("ref_t1.json")
p = t["params"]
G = ClassGroupB200(int(p["Delta"]))
G.set_scheme(int(p["k"]), int(p["N"]))
G.set_bases(I(p["h"]), I(t["pk"]), int(p["exponent_bound"]).bit_length())
rng = random.Random(3)
ms = [rng.getrandbits(int(p["k"])) for _ in range(ne)]
rs = [rng.randrange(1, int(p["exponent_bound"])) for _ in range(ne)]
ct = G.encrypt(G.ints(ms, G.ml), G.ints(rs, G.el_r))
G.status()
d = G.part_decrypt(ct[:, :, :npd].contiguous(), int(t["sk"]))
G.status()
print("done", ct.shape, d.shape)
