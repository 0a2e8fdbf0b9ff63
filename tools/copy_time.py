"""Pinned-copy timings: back-to-back vs one copy per synchronise (what a blocking host call sees)."""
import time, torch
for mb in (1.77, 3.54, 7.08, 28.3):
    n = int(mb * 1e6 / 4)
    p = torch.empty(n, dtype=torch.int32).pin_memory(); d = torch.empty(n, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    for name, src, dst in (("H2D", p, d), ("D2H", d, p)):
        dst.copy_(src, non_blocking=True); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(20): dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize(); tb = (time.perf_counter() - t0) / 20
        ts = 0
        for _ in range(20):
            time.sleep(0.0005)
            t0 = time.perf_counter(); dst.copy_(src, non_blocking=True); torch.cuda.synchronize(); ts += time.perf_counter() - t0
        ts /= 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        time.sleep(0.001); e0.record(); dst.copy_(src, non_blocking=True); e1.record(); torch.cuda.synchronize()
        print("%5.2f MB %s: back-to-back %.3f ms (%.1f GB/s); single+sync %.3f ms (%.1f GB/s); single by events %.3f ms" % (mb, name, tb * 1e3, mb / tb / 1e3, ts * 1e3, mb / ts / 1e3, e0.elapsed_time(e1)))
