"""Ad-hoc device timings of the batched primitives (not the bench): python tools/quick_time.py [tier]"""
import ctypes, json, os, random, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from conftest import I, load_golden
from cofhe_b200 import capi
from cofhe_b200.group import ClassGroupB200

tier = sys.argv[1] if len(sys.argv) > 1 else "t0"
t = load_golden("ref_%s.json" % tier)
G = ClassGroupB200(int(t["params"]["Delta"]))
lib = capi.load()
for wide in (1, 0):
    r = ctypes.c_double()
    lib.cofhe_b200_imad_probe(0, wide, ctypes.byref(r))
    print("imad probe wide=%d: %.3f T MAC/s" % (wide, r.value / 1e12))
h, pk = I(t["params"]["h"]), I(t["pk"])
rng = random.Random(1)
def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    return best
sizes = [int(x) for x in (sys.argv[2].split(",") if len(sys.argv) > 2 else ["32768", "262144"])]
for n in sizes:
    seeds = G.forms([h, pk] * (n // 2))
    f = G.nupow(seeds, exps=[rng.getrandbits(40) | 1 for _ in range(n)])
    g = G.nupow(seeds, exps=[rng.getrandbits(40) | 1 for _ in range(n)])
    G.status()
    out = torch.empty_like(f)
    ref = None
    for fb in (128, 256, 1128):   # threads per block of the register composition kernel; 1128: 128 + branch-free Euclid rounds
        G.set_option("fast_block", fb % 1000)
        G.set_option("fast_flat", 1 if fb > 1000 else 0)
        dt = timeit(lambda: G.nucomp(f, g, out=out)); print("n=%d block=%d nucomp  %.3f ms  %.1f M/s" % (n, fb, dt * 1e3, n / dt / 1e6))
        if ref is None: ref = out.clone()
        else: assert torch.equal(ref, out)
        dt = timeit(lambda: G.nudupl(f, out=out)); print("n=%d block=%d nudupl  %.3f ms  %.1f M/s" % (n, fb, dt * 1e3, n / dt / 1e6))
    G.set_option("fast_block", -1)
    G.set_option("fast_flat", -1)
    e = int(t["sk"])
    dt = timeit(lambda: G.nupow(f, shared=e), reps=2); print("n=%d nupow(%d bits shared) %.3f ms  %.1f k/s" % (n, e.bit_length(), dt * 1e3, n / dt / 1e3))
    G.status()
