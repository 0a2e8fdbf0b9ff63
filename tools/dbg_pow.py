import json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import I, load_golden
from cofhe_b200.group import ClassGroupB200
from oracle import qf_oracle as O
import torch
tier = load_golden("ref_%s.json" % sys.argv[1])
G = ClassGroupB200(int(tier["params"]["Delta"]))
vs = tier["nupow"]
idx = [int(x) for x in sys.argv[2].split(",")]
sub = [vs[i] for i in idx]
f = G.forms([I(v["f"]) for v in sub])
try:
    got = G.forms_out(G.nupow(f, exps=[int(v["n"]) for v in sub]))
    print(sys.argv[2], got == [I(v["r"]) for v in sub])
    for g, v in zip(got, sub):
        print("  exp", v["n"], "ok" if g == I(v["r"]) else ("BAD got %s want %s" % (g, I(v["r"]))))
    try:
        G.status(); print("  status ok")
    except Exception as e:
        print("  status", e)
except Exception as e:
    print(sys.argv[2], "EXC", str(e)[:60])
