"""Ad-hoc device timing of the matmul path: python tools/mm_time.py [tier] [n,m,p ...]"""
import os, random, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from conftest import I, load_golden
from cofhe_b200.group import ClassGroupB200
import bench as B

tier = sys.argv[1] if len(sys.argv) > 1 else "t0"
shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[2:]] or [(8, 64, 64), (64, 64, 64)]
fx = B.fixture(tier)
t = load_golden("ref_%s.json" % tier)
G = ClassGroupB200(fx["Delta"])
G.set_scheme(fx["k"], fx["N"])
G.set_bases(fx["h"], fx["pk"], fx["bound"].bit_length())
zero = G.cts([I(t["scal_matmul"][0]["zero"])])
for (n, m, p) in shapes:
    ct = B.random_cts(G, fx, n * m, 5)
    for label, s in (("pattern i+1", [i + 1 for i in range(m * p)]), ("uniform k-bit", [random.Random(3).getrandbits(fx["k"]) for _ in range(m * p)])):
        G.scal_matmul(ct, s, n, m, p, zero); torch.cuda.synchronize()
        best = 1e9
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); G.scal_matmul(ct, s, n, m, p, zero); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) * 1e-3)
        G.status()
        units = 2 * n * m * p
        print("%s %dx%d . %dx%d %-14s W=%s %.3f ms  %.2f M form-exps/s" % (tier, n, m, m, p, label, os.environ.get("COFHE_B200_MM_W", "4"), best * 1e3, units / best / 1e6))
