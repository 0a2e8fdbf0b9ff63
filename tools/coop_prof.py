"""One launch of each batched form primitive for profiling (ncu -k regex:...): python tools/coop_prof.py tier n coop"""
import os, random, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from conftest import I, load_golden
from cofhe_b200.group import ClassGroupB200

tier, n, coop = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
t = load_golden("ref_%s.json" % tier)
G = ClassGroupB200(int(t["params"]["Delta"]))
h, pk = I(t["params"]["h"]), I(t["pk"])
rng = random.Random(1)
nb = min(n, 1 << 12)
G.set_option("coop", 0)
seeds = G.forms([h, pk] * (nb // 2))
fb = G.nupow(seeds, exps=[rng.getrandbits(40) | 1 for _ in range(nb)])
gb = G.nupow(seeds, exps=[rng.getrandbits(40) | 1 for _ in range(nb)])
G.status()
reps = n // nb
f = fb.repeat(1, reps).contiguous()
g = torch.cat([torch.roll(gb, k + 1, dims=1) for k in range(reps)], dim=1).contiguous()
G.set_option("coop", coop)
out = torch.empty_like(f)
for _ in range(2):
    G.nucomp(f, g, out=out)
    G.nudupl(f, out=out)
G.status()
print("done")
