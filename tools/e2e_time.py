"""Where the host-pointer add spends its time: raw pinned copies vs cofhe_b200_add_ciphertexts_host per chunk count."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench as B
from cofhe_b200.group import ClassGroupB200, to_host

fx = B.fixture("t0")
G = ClassGroupB200(fx["Delta"], 0)
for n in (16384, 131072):
    ct1, ct2 = B.synth_cts(G, fx, n, 1), B.synth_cts(G, fx, n, 2)
    h1, h2 = to_host(ct1).copy(), to_host(ct2).copy()
    p1, p2 = torch.from_numpy(h1.view(np.int32)).pin_memory(), torch.from_numpy(h2.view(np.int32)).pin_memory()
    po = torch.empty_like(p1).pin_memory()
    d = torch.empty_like(ct1)
    def t(fn, reps=10):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps
    mb = h1.nbytes / 1e6
    th = t(lambda: d.copy_(p1, non_blocking=True)); td = t(lambda: po.copy_(d, non_blocking=True))
    print("n=%d  %.2f MB per operand: H2D %.1f GB/s (%.3f ms)  D2H %.1f GB/s (%.3f ms)" % (n, mb, mb / th / 1e3, th * 1e3, mb / td / 1e3, td * 1e3))
    out = torch.empty_like(ct1)
    tk = t(lambda: G.add_ciphertexts(ct1, ct2, out=out))
    print("   kernel only (device resident): %.3f ms" % (tk * 1e3))
    n1, n2, no = p1.numpy().view(np.uint32), p2.numpy().view(np.uint32), po.numpy().view(np.uint32)
    for c in (0, 1, 2, 4, 8, 12):
        os.environ["COFHE_B200_HOST_CHUNKS"] = str(c)
        te = t(lambda: G.add_ciphertexts_host(n1, n2, no))
        print("   chunks=%2d: %.3f ms  %.1f M ct-adds/s" % (c, te * 1e3, n / te / 1e6))
