"""Experiment: kernels reading/writing pinned host memory directly (UVA zero-copy) vs staged copies, after idle gaps."""
import ctypes, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench as B
from cofhe_b200 import capi
from cofhe_b200.group import ClassGroupB200, to_host

fx = B.fixture("t0")
G = ClassGroupB200(fx["Delta"], 0)
vp = ctypes.c_void_p
for n in (16384, 131072):
    ct1, ct2 = B.synth_cts(G, fx, n, 1), B.synth_cts(G, fx, n, 2)
    h1, h2 = to_host(ct1).copy(), to_host(ct2).copy()
    p1, p2 = torch.from_numpy(h1.view(np.int32)).pin_memory(), torch.from_numpy(h2.view(np.int32)).pin_memory()
    po = torch.empty_like(p1).pin_memory()
    ref = to_host(G.add_ciphertexts(ct1, ct2)); G.status()
    st = torch.cuda.current_stream().cuda_stream
    def zc_all():
        capi.check(G.lib.cofhe_b200_add_ciphertexts(G.h, n, vp(p1.data_ptr()), vp(p2.data_ptr()), vp(po.data_ptr()), vp(st)), "zc")
        torch.cuda.synchronize()
    dout = torch.empty_like(ct1)
    def zc_in():
        capi.check(G.lib.cofhe_b200_add_ciphertexts(G.h, n, vp(p1.data_ptr()), vp(p2.data_ptr()), vp(dout.data_ptr()), vp(st)), "zc")
        po.copy_(dout, non_blocking=True); torch.cuda.synchronize()
    d1, d2 = torch.empty_like(ct1), torch.empty_like(ct1)
    def staged():
        d1.copy_(p1, non_blocking=True); d2.copy_(p2, non_blocking=True)
        capi.check(G.lib.cofhe_b200_add_ciphertexts(G.h, n, vp(d1.data_ptr()), vp(d2.data_ptr()), vp(dout.data_ptr()), vp(st)), "st")
        po.copy_(dout, non_blocking=True); torch.cuda.synchronize()
    for name, fn in (("zero-copy in+out", zc_all), ("zero-copy in, D2H copy out", zc_in), ("staged copies", staged)):
        po.zero_(); fn(); fn()
        ok = np.array_equal(po.numpy().view(np.uint32).reshape(ref.shape), ref)
        for gap in (0.0, 0.001):
            tt = 0
            for _ in range(10):
                if gap: time.sleep(gap)
                t0 = time.perf_counter(); fn(); tt += time.perf_counter() - t0
            print("n=%d %-28s gap %.0f us: %.3f ms  %.1f M ct-adds/s  parity %s" % (n, name, gap * 1e6, tt / 10 * 1e3, n / (tt / 10) / 1e6, ok))
