"""Cooperative (lane-group) kernels against the register / generic path: equality of results and device timings.
    python tools/coop_time.py [tier ...]"""
import json, os, random, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from conftest import I, load_golden
from cofhe_b200.group import ClassGroupB200


def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    return best


for tier in (sys.argv[1:] or ["t0", "node", "t1"]):
    t = load_golden("ref_%s.json" % tier)
    G = ClassGroupB200(int(t["params"]["Delta"]))
    print("== %s: fast_limbs %d, coop_limbs %d" % (tier, G.fast, int(G.lib.cofhe_b200_coop_limbs(G.h))))
    h, pk = I(t["params"]["h"]), I(t["pk"])
    rng = random.Random(1)
    sizes = [1 << 13, 1 << 15, 1 << 17, 1 << 20] if tier != "t1" else [1 << 13, 1 << 15, 1 << 17]
    nmax = max(sizes)
    G.set_option("coop", 0)
    nb = 1 << 12
    seeds = G.forms([h, pk] * (nb // 2))
    fb = G.nupow(seeds, exps=[rng.getrandbits(40) | 1 for _ in range(nb)])
    gb = G.nupow(seeds, exps=[rng.getrandbits(40) | 1 for _ in range(nb)])
    G.status()
    reps = nmax // nb
    F, Gm = fb.repeat(1, reps), torch.roll(gb, 1, dims=1).repeat(1, reps)
    # decorrelate the pairs a little: shift the second operand by a per-copy offset
    Gm = torch.cat([torch.roll(gb, k + 1, dims=1) for k in range(reps)], dim=1)
    for n in sizes:
        f, g = F[:, :n].contiguous(), Gm[:, :n].contiguous()
        res = {}
        for mode in (0, 4, 14, 24):   # 0: register / generic path; 1, 4, 5: cooperative kernels compiled for 3, 4, 5 blocks per SM; 14: 4 + phase barriers
            G.set_option("coop", 1 if mode else 0)
            G.set_option("coop_mb", mode if mode > 1 else 3)
            out = torch.empty_like(f)
            dt = timeit(lambda: G.nucomp(f, g, out=out))
            G.status()
            res[mode] = out.clone()
            out2 = torch.empty_like(f)
            dt2 = timeit(lambda: G.nudupl(f, out=out2))
            G.status()
            res[(mode, 2)] = out2.clone()
            print("%s n=%7d coop=%d  nucomp %8.3f ms %9.2f M/s | nudupl %8.3f ms %9.2f M/s" % (tier, n, mode, dt * 1e3, n / dt / 1e6, dt2 * 1e3, n / dt2 / 1e6))
        print("   equal:", [bool(torch.equal(res[0], res[m])) and bool(torch.equal(res[(0, 2)], res[(m, 2)])) for m in (4, 14, 24)])
    G.set_option("coop", -1)
    G.close()
