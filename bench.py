"""bench.py -- CoFHE ciphertext hot path on B200.  BASELINE.json's metric has two halves; both are measured here:

  * ciphertext adds/s   -- `ciphertext_matadd` 128x128 (BASELINE.json configs[1]; reference benchmarks/local.cpp:65-138 at
                           128x128): the headline `value` / `e2e` / `roofline` of the JSON line;
  * form exponentiations/s -- scalar x ciphertext matmul 256x256 . 256x256 (configs[4]; benchmarks/local.cpp:140-215 shape
                           rules): the structured `form_exps` object of the same line (value, same-host
                           reference value, parity, IMAD roofline, e2e through host buffers), strong-scaled over the ranks.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--tier t0|t1] [--no-cpu] [--no-extra]

One process per GPU.  matadd: every rank owns an independent 128x128 matrix (weak scaling, no data-path collective:
SURVEY.md 8e); a step = one add_ciphertext_tensors pass (res <- res + ct2), timed with CUDA events on the launching stream,
inputs resident in HBM, L2 flushed between steps.  `e2e` goes through the host-pointer C-ABI entry point (page-locked host
buffers; transfers inside the timed region); `e2e_class` through B200CryptoSystem on Tensor<CipherText*> (the API the
reference's callers use).  Inputs are the reference benchmark's: ct = encrypt_tensor(pt[i] = i+1), one shared randomness
per tensor.  The oracle / compiled reference (oracle/) is used here only as the checker and for the `cpu_baseline` /
`--impl reference` legs.
"""
import argparse
import ctypes
import json
import os
import random
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

TIERS = {"t0": ("ref_t0.json", 128, 128), "t1": ("ref_t1.json", 1792, 128)}
ROWS = COLS = 128
MM = (256, 256, 256)            # config 5: ct 256x256, scalars 256x256
MM_CPU_ROWS = 8                 # rows of config 5 the reference runs on the host cores (scaled linearly)
BIG_N = 1 << 16                 # configs 3 / 4

METRIC = "ciphertext adds/sec (ciphertext_matadd 128x128)"
DATA = "synthetic: ct1, ct2 = encrypt_tensor(pt[i] = i+1), one shared randomness per tensor, as benchmarks/local.cpp:65-138"
DTYPE = "u32 limbs (multi-precision integer)"


def config_of(tier):
    _, nbits, k = TIERS[tier]
    return {"workload": "ciphertext_matadd 128x128 per GPU: one add_ciphertext_tensors pass per step, chained res <- res + ct2 "
                        "(benchmarks/local.cpp:65-138 at 128x128); tier %s (N %d bits, k %d)" % (tier, nbits, k),
            "l2": "GPU arm: L2 flushed between steps (256 MiB memset), per-step CUDA events on the launch stream, summed; "
                  "reference arm: host caches as they are",
            "sharding": "one independent matrix per rank, no collective (the reference arm runs one matrix on the host cores)"}


def fixture(tier):
    with open(os.path.join(ROOT, "tests", "golden", TIERS[tier][0])) as fh:
        t = json.load(fh)
    conv = lambda x: tuple(int(y) for y in x)
    return {"tier": tier, "N": int(t["params"]["N"]), "Delta": int(t["params"]["Delta"]), "h": conv(t["params"]["h"]), "pk": conv(t["pk"]),
            "bound": int(t["params"]["exponent_bound"]), "sk": int(t["sk"]), "k": int(t["params"]["k"]),
            "shares": [int(row[0]) for row in t.get("shares", [])[:3]]}


def limb_s(dbits):
    """s of SURVEY.md section 8d: 32-bit limbs of a reduced coefficient."""
    return ((dbits + 1) // 2 + 31) // 32


def w_comp(s):
    return 16 * s * s


def w_dupl(s):
    return 14 * s * s


def w_exp(s, e):
    return (e - 1) * w_dupl(s) + -(-e // 3) * w_comp(s)


def host_threads():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 - 0.05 <= t <= t1 + 0.15 and len(r) >= 6] or [r for _, r in self.rows if len(r) >= 6]
        mhz = sorted(int(r[0]) for r in rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i] == "Active" for r in rows)]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": int(rows[0][1]) if rows and rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(rows)}


def dist_env():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    return rank, world, local


# ---------------------------------------------------------------------------------------------------------------------
# the reference on the host cores (oracle/_ref: the UNMODIFIED CPUCryptoSystem + BICYCL + libgmp)
# ---------------------------------------------------------------------------------------------------------------------
def open_reference(tier):
    from oracle import ref as R
    if not R.available():
        return None
    _, nbits, k = TIERS[tier]
    ref = R.Reference(nbits, k)
    ks = ref.call("keygen")
    ref.set_threads(host_threads())
    ref.sk, ref.pk = int(ks[0]), tuple(int(x) for x in ks[1:4])
    return ref


def cpu_baseline_matadd(tier, G=None, iters=50):
    """config 2 exactly (128x128, `iters` chained adds) on all OpenMP threads, then 1 thread on a shorter chain.  When a
    GPU group is given, the GPU repeats the same chain on the reference's own inputs and the final ciphertexts are compared
    bit for bit."""
    ref = open_reference(tier)
    cores = ref.threads()
    na, nc = (G.na, G.nc) if G is not None else (0, 0)
    dt, bufs = ref.bulk_matadd(ROWS, COLS, iters, na, nc, export=G is not None)
    out = {"value": ROWS * COLS * iters / dt, "unit": "ct_adds/s", "cores": cores, "kind": "reference",
           "sample": "%dx%d ciphertext_matadd, %d chained add_ciphertext_tensors, OMP threads=%d, %.2f s" % (ROWS, COLS, iters, cores, dt)}
    ref.set_threads(1)
    it1 = max(1, iters // 10)
    dt1, _ = ref.bulk_matadd(ROWS, COLS, it1)
    out["value_1core"] = ROWS * COLS * it1 / dt1
    ref.set_threads(cores)
    if G is not None:
        from cofhe_b200.group import to_device, to_host
        a, b = to_device(bufs[0], G.device), to_device(bufs[1], G.device)
        res = G.add_ciphertexts(a, b)
        for _ in range(iters - 1):
            res = G.add_ciphertexts(res, b)
        G.status()
        out["parity"] = bool((to_host(res) == bufs[2]).all())
    ref.close()
    return out


def reference_matmul(ref, G_shape, rows, zero_seed=777):
    """CPUCryptoSystem::scal_ciphertext_tensors on `rows` ct rows of config 5 (pattern scalars i+1), all host threads.
    Returns (seconds, units, ct limbs, zero randomness, result limbs): inputs encrypt_tensor(pt[i] = i+1); the call's `zero`
    draw is pinned by re-seeding the RandGen, so that the GPU can repeat the call bit for bit."""
    from cofhe_b200 import limbs as L
    na, nc = G_shape
    n, m, p = rows, MM[1], MM[2]
    v = ref.call("encrypt_tensor", n, m, [i + 1 for i in range(n * m)])
    cts = [(tuple(v[i:i + 3]), tuple(v[i + 3:i + 6])) for i in range(0, len(v), 6)]
    ct = L.cts_to_limbs(cts, na, nc)
    bound = ref.params()["exponent_bound"]
    ref.call("set_seed", zero_seed)
    zr = ref.call("rand_mpz", bound)[0]
    ref.call("set_seed", zero_seed)
    dt, out = ref.scal_matmul_given(n, m, p, [i + 1 for i in range(m * p)], ct, na, nc)
    return dt, 2 * n * m * p, ct, zr, out


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    from oracle import ref as R
    if not R.available():
        emit_line({"impl": "reference", "unavailable": "oracle/_ref/libcofhe_ref.so (compiled reference) missing or libgmp.so.10 not loadable"})
        return
    ref = open_reference(args.tier)
    cores = ref.threads()
    if args.warmup:
        ref.bulk_matadd(ROWS, COLS, args.warmup)
    dt, _ = ref.bulk_matadd(ROWS, COLS, args.steps)
    n = ROWS * COLS
    v = n * args.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "ct_adds/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": DATA,
            "config": config_of(args.tier),
            "cpu_baseline": {"value": v, "unit": "ct_adds/s", "cores": cores, "kind": "reference",
                             "sample": "%d chained full 128x128 add_ciphertext_tensors passes of CPUCryptoSystem (BICYCL + GMP + OpenMP)" % args.steps},
            "e2e": {"value": v, "unit": "ct_adds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if not args.no_extra:
        try:
            from cofhe_b200 import limbs as L
            dbits = (-int(ref.params()["Delta"])).bit_length()
            na, nc, _ = L.limb_shape(dbits)
            dtm, units, _, _, _ = reference_matmul(ref, (na, nc), MM_CPU_ROWS)
            line["form_exps"] = {"metric": "form exponentiations/sec (scal_matmul %dx%d . %dx%d)" % (MM[0], MM[1], MM[1], MM[2]),
                                 "value": units / dtm, "unit": "form_exps/s", "cores": cores,
                                 "sample": "%d of %d ct rows, pattern scalars i+1, one scal_ciphertext_tensors call, %.2f s; "
                                           "scales linearly in the rows" % (MM_CPU_ROWS, MM[0], dtm)}
        except Exception as e:  # the second metric must not take the headline down
            line["form_exps"] = {"error": str(e)[:160]}
    ref.close()
    emit_line(line)


# ---------------------------------------------------------------------------------------------------------------------
# the product arm
# ---------------------------------------------------------------------------------------------------------------------
def enc_pattern(G, fx, n, seed):
    """encrypt_tensor(pt[i] = i+1) on the device: ONE randomness for the tensor (c1 = h^r for every element,
    c2_i = pk^r o f^(i+1)), reference cpu_cryptosystem_tensor_ops.inl:1-19 as benchmarks/local.cpp calls it."""
    r = random.Random(seed).randrange(1, fx["bound"])
    ct = G.encrypt([i + 1 for i in range(n)], r)
    G.status()
    return ct


def random_cts(G, fx, n, seed):
    """n ciphertexts with own randomness each (c1 = h^r_i, c2 = pk^r_i o f^m_i): no two compositions alike"""
    rng = random.Random(seed)
    ct = G.encrypt([rng.getrandbits(fx["k"]) for _ in range(n)], [rng.randrange(1, fx["bound"]) for _ in range(n)])
    G.status()
    return ct


def bind_near_gpu(index):
    """One process per GPU: run on (and allocate page-locked host buffers from) the CPU socket the GPU hangs off, so
    that the host-pointer path does not cross the socket interconnect.  Best effort (NVML's CPU affinity mask)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass


def _timed(fn, world, reps=2):
    """best-of-reps device time of fn() in seconds: CUDA events on the current stream, max over ranks."""
    import torch
    import torch.distributed as dist
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) * 1e-3], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, float(t.item()))
    return best


def _wall(fn, world, reps=2):
    """best-of-reps wall time of a blocking host call, barrier + synchronize on both sides, max over ranks"""
    import torch
    import torch.distributed as dist
    fn()
    best = 1e30
    for _ in range(reps):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, float(t.item()))
    return best


def run_product(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from cofhe_b200.group import ClassGroupB200, to_host
    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        bind_near_gpu(local)
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout for the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    fx = fixture(args.tier)
    G = ClassGroupB200(fx["Delta"], local)
    G.set_scheme(fx["k"], fx["N"])
    G.set_bases(fx["h"], fx["pk"], fx["bound"].bit_length())
    dbits = (-fx["Delta"]).bit_length()
    s = limb_s(dbits)
    n = ROWS * COLS
    ct1, ct2 = enc_pattern(G, fx, n, 1000 + rank), enc_pattern(G, fx, n, 2000 + rank)
    res = torch.empty_like(ct1)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=G.device)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(first):
        G.add_ciphertexts(ct1 if first else res, ct2, out=res)

    for i in range(args.warmup):
        step(i == 0)
        flush.zero_()
    step(True)
    barrier()
    l0 = G.launches()
    sampler = ClockSampler(local) if rank == 0 else None
    evs = []
    t0 = time.time()
    for i in range(args.steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step(False)
        e1.record()
        evs.append((e0, e1))
    barrier()
    t1 = time.time()
    G.status()
    launches = G.launches() - l0
    kern_ms = sum(a.elapsed_time(b) for a, b in evs)
    # ---- e2e: host buffers through the C-ABI host entry point (H2D + kernel + D2H inside the timed region)
    h1, h2 = to_host(ct1).copy(), to_host(ct2).copy()
    p1, p2 = torch.from_numpy(h1.view(np.int32)).pin_memory(), torch.from_numpy(h2.view(np.int32)).pin_memory()
    po = torch.empty_like(p1).pin_memory()
    n1, n2, no = p1.numpy().view(np.uint32), p2.numpy().view(np.uint32), po.numpy().view(np.uint32)
    e2e_steps = max(3, min(args.steps, 20))
    for _ in range(2):
        G.add_ciphertexts_host(n1, n2, no)
    barrier()
    te = time.perf_counter()
    for _ in range(e2e_steps):
        G.add_ciphertexts_host(n1, n2, no)
    barrier()
    e2e_s = time.perf_counter() - te
    clocks = sampler.stop(t0, time.time()) if sampler else None
    times = torch.tensor([kern_ms, e2e_s * 1e3], dtype=torch.float64, device=G.device)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    kern_ms, e2e_ms = times.tolist()
    imad = ctypes.c_double()
    G.lib.cofhe_b200_imad_probe(G.device.index, 1, ctypes.byref(imad))
    # the same launch on ciphertexts with independent randomness (every composition different: the harder input)
    rnd = None
    if not args.no_extra:
        r1, r2 = random_cts(G, fx, n, 31 + rank), random_cts(G, fx, n, 32 + rank)
        ro = torch.empty_like(r1)
        dt = _timed(lambda: G.add_ciphertexts(r1, r2, out=ro), world, reps=5)
        rnd = {"workload": "128x128 ciphertexts with own randomness each", "ct_adds_per_s": world * n / dt, "ms": dt * 1e3}
        del r1, r2, ro
    form_exps = None if args.no_extra else guarded(lambda: matmul_metric(args, G, fx, s, world, rank, imad.value))
    extra = None if args.no_extra else extras(args, G, fx, s, world, rank, imad.value)
    if rank == 0:
        emit(args, G, fx, s, world, n, kern_ms, e2e_ms, e2e_steps, launches, clocks, h1.nbytes, imad.value, form_exps, extra, rnd)
    if world > 1:
        dist.destroy_process_group()


def guarded(fn):
    try:
        return fn()
    except Exception as e:  # a secondary measurement must not take the headline line down
        return {"error": "%s: %s" % (type(e).__name__, str(e)[:200])}


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the committed `ncu --set full` capture of
    this same command (profiles/ncu_traffic.json); None when there is no capture for this tier's kernel."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
            t = json.load(fh)[kernel]
        return int(t["dram_bytes_read"]) + int(t["dram_bytes_write"])
    except Exception:
        return None


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def matmul_metric(args, G, fx, s, world, rank, imad_peak):
    """BASELINE.json's other half: form exponentiations/s of scal_ciphertext_tensors' matmul on config 5
    (256x256 . 256x256, pattern scalars i+1 <= 17 bits), the ct rows split over the ranks (strong scaling), one NCCL
    all-gather of the result rows inside the timed region.  At N = 1 the compiled reference runs MM_CPU_ROWS rows of the
    same problem on all host cores in this run, and the GPU's rows are compared with its output bit for bit."""
    import numpy as np
    import torch
    from cofhe_b200 import shard as S
    from cofhe_b200.group import to_device, to_host
    n, m, p = MM
    if not G.reg:
        n = 8 * world  # the 2053-bit tier (cooperative kernels, 25x the arithmetic per unit): 8 ct rows per GPU of the same 256-wide problem
    rows = n // world
    units = 2 * n * m * p
    sc = [i + 1 for i in range(m * p)]
    ebits = max(x.bit_length() for x in sc)
    scd = G.ints(sc, (ebits + 31) // 32)
    out = {"metric": "form exponentiations/sec (scal_matmul %dx%d . %dx%d)" % (MM[0], m, m, p), "unit": "form_exps/s", "scaling": "strong",
           "workload": "scal_ciphertext_tensors matmul: ct %dx%d = encrypt_tensor(pt[i]=i+1), scalars %dx%d = i+1 (<= %d bits), "
                       "%d ct rows per GPU, one NCCL all-gather of the result rows; 1 unit = one form raised to one scalar incl. its "
                       "accumulate composition = %d units" % (n, m, m, p, ebits, rows, units)}
    ref = zr = None
    ref_ct = ref_out = None
    if world == 1 and not args.no_cpu:
        ref = open_reference(args.tier)
    if ref is not None:
        dtc, cunits, ref_ct, zr, ref_out = reference_matmul(ref, (G.na, G.nc), MM_CPU_ROWS)
        out["reference"] = {"value": cunits / dtc, "unit": "form_exps/s", "cores": ref.threads(), "kind": "reference",
                            "sample": "%d of %d ct rows on all host cores, one scal_ciphertext_tensors call, %.2f s (scales linearly in the rows)"
                                      % (MM_CPU_ROWS, n, dtc)}
        ref.close()
    else:
        zr = random.Random(777).randrange(1, fx["bound"])
    zero = G.encrypt([0], zr)  # the call's `zero` = encrypt(pk, 0)
    ct_all = enc_pattern(G, fx, n * m, 1)  # same tensor on every rank (same seed); each rank takes its rows
    if ref_ct is not None:
        # the reference's own encrypt_tensor(i+1) for its rows, so that the comparison is on identical inputs
        ct_all[:, :, :MM_CPU_ROWS * m] = to_device(ref_ct, G.device)
    ct = ct_all[:, :, rank * rows * m:(rank + 1) * rows * m].contiguous()
    del ct_all
    hold = {}

    def run():
        hold["local"] = G.scal_matmul(ct, scd, rows, m, p, zero, ebits=ebits)
        hold["all"] = S.gather_last(hold["local"], [rows * p] * world)
    dt = _timed(run, world, reps=3)
    G.status()
    out["value"] = units / dt
    out["seconds"] = dt
    macs = units * (w_exp(s, ebits) + w_comp(s)) / dt
    out["roofline_imad"] = {"bound": "imad", "achieved": macs / world / 1e12, "peak": imad_peak / 1e12, "unit": "T MAC/s (32x32->64) per GPU",
                            "frac": macs / world / imad_peak,
                            "model": "W_exp(%d bit) + W_comp = %d nominal MACs per unit (SURVEY.md 8d: algorithm-independent; the kernel "
                                     "does ~3 table compositions per unit)" % (ebits, w_exp(s, ebits) + w_comp(s))}
    if ref_out is not None:
        got = to_host(hold["local"])[:, :, :MM_CPU_ROWS * p]
        out["parity"] = bool((got == ref_out).all())
        out["ratio_vs_reference"] = out["value"] / out["reference"]["value"]
    # ---- e2e: page-locked host buffers in and out (ct rows + scalars + zero up, result rows down inside the timed region)
    hc, hs = to_host(ct), to_host(scd)
    pc, ps = torch.from_numpy(hc.view(np.int32).copy()).pin_memory(), torch.from_numpy(hs.view(np.int32).copy()).pin_memory()
    po = torch.empty((2, G.fl, rows * p), dtype=torch.int32).pin_memory()
    dc, ds = torch.empty_like(ct), torch.empty_like(scd)

    def e2e():
        dc.copy_(pc, non_blocking=True)
        ds.copy_(ps, non_blocking=True)
        r = G.scal_matmul(dc, ds, rows, m, p, zero, ebits=ebits)
        po.copy_(r, non_blocking=True)
        torch.cuda.current_stream().synchronize()
    dte = _wall(e2e, world)
    G.status()
    out["e2e"] = {"value": units / dte, "unit": "form_exps/s", "h2d_bytes_per_step": pc.numel() * 4 + ps.numel() * 4,
                  "d2h_bytes_per_step": po.numel() * 4,
                  "api": "cofhe_b200_upload + cofhe_b200_scal_matmul + cofhe_b200_download pattern on page-locked SoA buffers, per rank; no gather"}
    # uniform k-bit scalars (the general case of the same call)
    k = fx["k"]
    rng = random.Random(4242)
    sck = G.ints([rng.getrandbits(k) for _ in range(m * p)], (k + 31) // 32)

    def runk():
        S.gather_last(G.scal_matmul(ct, sck, rows, m, p, zero, ebits=k), [rows * p] * world)
    dtk = _timed(runk, world, reps=1)
    G.status()
    out["uniform_k_bit_scalars"] = {"value": units / dtk, "unit": "form_exps/s", "seconds": dtk,
                                    "imad_frac_nominal": units * (w_exp(s, k) + w_comp(s)) / dtk / world / imad_peak}
    return out


def extras(args, G, fx, s, world, rank, imad_peak):
    """The other BASELINE.json configs (3: encryption of 2^16 plaintexts with own randomness; 4: threshold decryption of
    2^16 ciphertexts, 3 shares + combine), element ranges split over the ranks (strong scaling); Beaver pipeline; wire
    format.  At tier t0 the inputs are the ones of tests/big_inputs.py, whose reference outputs are pinned by SHA-256 in
    tests/golden/big_t0.npz (generated from the compiled reference by tools/gen_golden_big.py)."""
    import numpy as np
    import torch
    import big_inputs as B
    from cofhe_b200 import shard as S
    from cofhe_b200.group import to_host
    out = {}
    k, bound = fx["k"], fx["bound"]
    cnt = BIG_N
    lo, hi = rank * cnt // world, (rank + 1) * cnt // world
    ms_all, rs_all = B.enc_inputs(k, bound, cnt)
    ms, rs = G.ints(ms_all[lo:hi], G.ml), G.ints(rs_all[lo:hi], G.el_r)
    z = B.load() if (fx["tier"] == "t0" and os.path.exists(B.GOLDEN)) else None
    sizes = [hi - lo] * world if cnt % world == 0 else None
    hold = {}

    def gather(t):
        return S.gather_last(t, sizes) if sizes else t

    def enc():
        hold["ct"] = G.encrypt(ms, rs)
        hold["ct_all"] = gather(hold["ct"])
    dt = _timed(enc, world)
    G.status()
    eb = bound.bit_length()
    out["config3_encrypt"] = guarded(lambda: {
        "workload": "%d plaintexts (random k-bit), own randomness per ciphertext (2 fixed-base exps + f^m + 1 composition each), "
                    "%d per GPU + NCCL gather" % (cnt, hi - lo),
        "encryptions_per_s": cnt / dt, "fixed_base_form_exps_per_s": 2 * cnt / dt, "seconds": dt,
        "imad_frac_nominal": 2 * cnt * w_exp(s, eb) / dt / world / imad_peak,
        "parity": (B.sha(to_host(hold["ct_all"])) == B.stored_sha(z, "enc_sha")) if (z is not None and sizes) else None})
    shares = fx["shares"] if len(fx["shares"]) == 3 else [fx["sk"]] * 3

    def pdec():
        hold["ds"] = [G.part_decrypt(hold["ct"], sh) for sh in shares]
    dtp = _timed(pdec, world)
    G.status()

    def comb():
        hold["m"] = G.combine(hold["ct"], torch.stack(hold["ds"]))
        hold["m_all"] = gather(hold["m"])
    dtc = _timed(comb, world)
    G.status()
    sb = max(abs(x).bit_length() for x in shares)

    def c4():
        par = None
        if z is not None and sizes:
            par = all(B.sha(to_host(gather(hold["ds"][q]))) == B.stored_sha(z, "pd_sha_%d" % q) for q in range(3)) and \
                B.sha(to_host(hold["m_all"])) == B.stored_sha(z, "comb_sha")
        return {"workload": "%d ciphertexts, 3-of-5 threshold: part_decrypt_tensor for 3 shares (%d-bit shared exponents) + "
                            "combine_part_decryption_results_tensor, %d per GPU + NCCL gather of the plaintexts" % (cnt, sb, hi - lo),
                "threshold_decryptions_per_s": cnt / (dtp + dtc), "part_decrypt_form_exps_per_s": 3 * cnt / dtp,
                "part_decrypt_seconds": dtp, "combine_seconds": dtc,
                "imad_frac_nominal_part_decrypt": 3 * cnt * w_exp(s, sb) / dtp / world / imad_peak,
                "plaintexts_recovered": bool(torch.equal(hold["m"], ms)), "parity": par}
    out["config4_threshold"] = guarded(c4)

    def dec():
        hold["md"] = G.decrypt(hold["ct"], fx["sk"])
    dtd = _timed(dec, world, reps=1)
    G.status()
    out["decrypt"] = {"workload": "%d ciphertexts (c1^sk, nucompinv, dlog_F), %d per GPU" % (cnt, hi - lo), "decryptions_per_s": cnt / dtd,
                      "roundtrip_ok": bool(torch.equal(hold["md"], ms))}
    # ---- section 8f-2: Beaver-triple ciphertext x ciphertext multiplication (device side, without the network round)
    def beaver():
        nb = 1 << 13
        A = hold["ct"][:, :, :nb].contiguous()
        msb = ms[:, :nb].contiguous()
        Bc = G.encrypt(msb, 4242)
        C = G.add_ciphertexts(A, Bc)  # any ciphertext batch serves as c for timing; correctness is in tests/test_gpu_scheme.py
        t_open = _timed(lambda: G.beaver_open(A, Bc, A, Bc), world, reps=1)
        t_close = _timed(lambda: G.beaver_close(msb, msb, A, Bc, C, 31337), world, reps=1)
        G.status()
        return {"workload": "%d ciphertext pairs/GPU: open (2 negate + 2 add) + close (2 scal + encrypt_tensor + 3 add), k = %d bit exponents; "
                            "the threshold-decryption round between them is not included" % (nb, k),
                "open_ms": t_open * 1e3, "close_ms": t_close * 1e3, "ct_mults_per_s": world * nb / (t_open + t_close)}
    out["beaver"] = guarded(beaver)
    del hold
    # ---- the reference's own CPU code on bounded samples of the same workloads, same run (rank 0 of a 1-GPU run)
    if world == 1 and not args.no_cpu:
        out["reference_cpu"] = guarded(lambda: reference_extras(G, fx))
    # ---- section 8f-1: the reference's binary tensor wire format <-> SoA on the device (128x128 ciphertexts)
    out["wire"] = guarded(lambda: wire_extra(args, G, fx, world, rank))
    return out


def reference_extras(G, fx):
    """Bounded samples of configs 1, 3, 4 on the UNMODIFIED reference (oracle/_ref) with all host threads, and the GPU on
    the reference's own inputs where the outputs are comparable."""
    import numpy as np
    import torch
    ref = open_reference(fx["tier"])
    if ref is None:
        return {"unavailable": "oracle/_ref not loadable"}
    sk, pk = ref.sk, ref.pk
    res = {"cores": ref.threads(), "kind": "reference"}
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int32)).to(G.device)
    # config 1 EXACTLY: benchmarks/local.cpp scal_matmul (8x64 . 64x64, pattern i+1, 50 chained calls)
    if fx["tier"] == "t0":
        dt, _, _ = ref.bulk_scal_matmul(8, 64, 64, iters=50)
        res["config1_scal_matmul"] = {"sample": "8x64 . 64x64, pattern i+1, 50 chained scal_ciphertext_tensors calls (benchmarks/local.cpp:140-215)",
                                      "seconds": dt, "form_exps_per_s": 50 * 2 * 8 * 64 * 64 / dt}
    # config 3 mode B: own r per ciphertext; the GPU repeats it with the reference's (m, r) under the reference's pk
    cnt = 4096
    dt, mo, ro, co = ref.bulk_encrypt(cnt, G.ml, G.el_r, G.na, G.nc)
    G.set_bases(fx["h"], pk, fx["bound"].bit_length())
    mine = G.encrypt(dev(mo), dev(ro))
    G.status()
    res["encrypt"] = {"sample": "%d ciphertexts, own r each" % cnt, "seconds": dt, "encryptions_per_s": cnt / dt,
                      "gpu_parity": bool(torch.equal(mine, dev(co)))}
    # config 4: part_decrypt_tensor with one share
    cnt = 2048
    dt, co, do = ref.bulk_part_decrypt(cnt, sk, G.na, G.nc)
    mine = G.part_decrypt(dev(co), sk)
    G.status()
    res["part_decrypt"] = {"sample": "%d ciphertexts, %d-bit share" % (cnt, sk.bit_length()), "seconds": dt,
                           "form_exps_per_s": cnt / dt, "gpu_parity": bool(torch.equal(mine, dev(do)))}
    G.set_bases(fx["h"], fx["pk"], fx["bound"].bit_length())
    ref.close()
    return res


def wire_extra(args, G, fx, world, rank):
    """serialize_ciphertext_tensor / deserialize_ciphertext_tensor of a 128x128 tensor as device kernels (wire bytes and SoA
    batch both resident in HBM), bytes/s of the wire buffer.  On rank 0 of a 1-GPU run the compiled reference does the
    same tensor on the host cores and its bytes are the parity check."""
    import numpy as np
    import torch
    from cofhe_b200 import capi
    n = ROWS * COLS
    ct = random_cts(G, fx, n, 7)
    ref_line = None
    if world == 1 and not args.no_cpu:
        try:
            ref = open_reference(fx["tier"])
            if ref is not None:
                ds, dd, cts, wire = ref.bulk_wire(ROWS, COLS, 3, G.na, G.nc)
                cores = ref.threads()
                ref.close()
                dev = torch.from_numpy(cts.view(np.int32)).to(G.device)
                mine = G.cts_to_wire([ROWS, COLS], dev)
                shape, back = G.wire_to_cts(wire)
                G.status()
                ref_line = {"serialize_s": ds / 3, "deserialize_s": dd / 3, "bytes": len(wire), "cores": cores,
                            "parity": bool(mine == wire and shape == [ROWS, COLS] and torch.equal(back, dev))}
                ct = dev
        except Exception as e:  # the checker is optional on the GPU box
            ref_line = {"error": str(e)[:100]}
    cap = int(G.lib.cofhe_b200_wire_bound(G.h, 2, n))
    wbuf = torch.empty(cap, dtype=torch.uint8, device=G.device)
    dims = (ctypes.c_uint32 * 2)(ROWS, COLS)
    written = ctypes.c_size_t()
    st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ptr = lambda t: ctypes.c_void_p(t.data_ptr())

    def ser():
        capi.check(G.lib.cofhe_b200_cts_to_wire(G.h, 2, dims, n, ptr(ct), ptr(wbuf), cap, ctypes.byref(written), st()), "cts_to_wire")
    t_ser = _timed(ser, world, reps=3)
    nbytes = written.value
    back = torch.empty_like(ct)

    def de():
        capi.check(G.lib.cofhe_b200_wire_to_cts(G.h, ptr(wbuf), nbytes, 2, n, ptr(back), st()), "wire_to_cts")
    t_de = _timed(de, world, reps=3)
    G.status()
    out = {"workload": "%dx%d ciphertext tensor, wire bytes and SoA batch resident in HBM" % (ROWS, COLS), "wire_bytes": nbytes,
           "serialize_ms": t_ser * 1e3, "deserialize_ms": t_de * 1e3, "serialize_GBps": nbytes / t_ser / 1e9,
           "deserialize_GBps": nbytes / t_de / 1e9, "roundtrip_ok": bool(torch.equal(back, ct))}
    if ref_line:
        out["reference_cpu"] = ref_line
    return out


def class_e2e(steps):
    """The same 128x128 chained add through B200CryptoSystem::add_ciphertext_tensors on Tensor<CipherText*> (BICYCL objects in,
    BICYCL objects out, marshalling inside the timed region) -- the API the reference's callers use -- next
    to the reference class on the same host, compared bit for bit (tests/cpp/b200_cs_parity.cpp --bench)."""
    exe = os.path.join(ROOT, "tests", "_cpp", "b200_cs_parity")
    if not os.path.exists(exe):
        return {"unavailable": "tests/_cpp/b200_cs_parity not built (needs the reference headers at build time)"}
    env = dict(os.environ)
    env["OMP_NUM_THREADS"] = str(host_threads())
    r = subprocess.run([exe, "--bench", "matadd", str(ROWS), str(COLS), str(steps), str(min(steps, 5))], capture_output=True, text=True,
                       timeout=900, env=env)
    if r.returncode != 0 and not r.stdout.strip():
        return {"error": (r.stderr or "exit %d" % r.returncode)[-200:]}
    d = json.loads(r.stdout.strip().splitlines()[-1])
    return {"value": d["b200_units_per_s"], "unit": "ct_adds/s", "reference_value": d["cpu_units_per_s"], "reference_threads": d["cpu_threads"],
            "parity": d["mismatches"] == 0, "steps": d["iters"],
            "device_resident_value": d.get("b200_resident_units_per_s"),
            "device_resident_api": "B200CryptoSystem::upload_ciphertexts x2, add_ciphertext_tensors(pk, DeviceBatch, DeviceBatch) x steps, "
                                   "download_ciphertexts: the same chain with the tensors kept on the device between the calls",
            "api": "B200CryptoSystem::add_ciphertext_tensors(pk, Tensor<CipherText*>, Tensor<CipherText*>) -> Tensor<CipherText*>: "
                   "mpz_export of 2 x 16384 ciphertexts, upload, kernel, download, 16384 new CipherText per step; reference_value = "
                   "CPUCryptoSystem::add_ciphertext_tensors on the same host in the same process"}


def emit(args, G, fx, s, world, n, kern_ms, e2e_ms, e2e_steps, launches, clocks, ct_bytes, imad_peak, form_exps, extra, rnd):
    peak_gbs, peak_src = measured_peaks()
    ms_step = kern_ms / args.steps
    value = world * n * args.steps / (kern_ms * 1e-3)
    alg_bytes = 72 * s * n  # (2 in + 1 out) x 6 coefficients x s limbs x 4 B per ct-add
    # bytes that cross PCIe towards the device per step: on the register-resident tiers the host entry point reads only
    # the a and b rows of both operands (c is recomputed on the device); all rows of the result travel back
    if G.reg:
        h2d_bytes = 2 * ct_bytes * (2 * G.na) // G.fl
        h2d_note = "a and b rows only: %d of %d limb rows per form" % (2 * G.na, G.fl)
    else:
        h2d_bytes, h2d_note = 2 * ct_bytes, "all limb rows, staged"
    achieved = alg_bytes / (ms_step * 1e-3) / 1e9
    macs = 2 * n * w_comp(s) / (ms_step * 1e-3)
    kernel = "k_fast_compose" if G.reg else "k_coop_compose"
    line = {"metric": METRIC, "value": value, "unit": "ct_adds/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": DTYPE, "data": DATA, "config": config_of(args.tier),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs,
                         "traffic": ncu_traffic(kernel), "peak_source": peak_src, "kernel": kernel,
                         "alg_bytes_per_launch": alg_bytes, "alg_bytes_model": "72*s = %d B per ct-add (SURVEY.md 8d), s = %d limbs" % (72 * s, s),
                         "layout_bytes_per_launch": 3 * ct_bytes},
            "roofline_imad": {"bound": "imad", "achieved": macs / 1e12, "peak": imad_peak / 1e12, "unit": "T MAC/s (32x32->64)",
                              "frac": macs / imad_peak, "model": "2*W_comp = 32*s^2 MACs per ct-add (SURVEY.md 8d)",
                              "peak_source": "measured in this run (cofhe_b200_imad_probe, IMAD.WIDE.U32 chains on all SMs)"},
            "e2e": {"value": world * n * e2e_steps / (e2e_ms * 1e-3), "unit": "ct_adds/s", "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": ct_bytes,
                    "api": "cofhe_b200_add_ciphertexts_host: blocking call on page-locked host SoA buffers; the kernels read "
                           "the operands from and write the result to host memory over PCIe inside the call (%s)" % h2d_note},
            "gpu_launches": launches, "clocks": clocks}
    if rnd is not None:
        line["matadd_random_inputs"] = rnd
    if world == 1 and not args.no_cpu:
        from oracle import ref as R
        if R.available():
            line["cpu_baseline"] = guarded(lambda: cpu_baseline_matadd(args.tier, G))
            if args.tier == "t0":
                line["e2e_class"] = guarded(lambda: class_e2e(max(5, min(args.steps, 20))))
        else:
            line["cpu_baseline"] = {"value": None, "unit": "ct_adds/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not loadable"}
    if form_exps is not None:
        line["form_exps"] = form_exps
    if extra is not None:
        line["extra"] = extra
    emit_line(line)


_REAL_STDOUT = None


def guard_stdout():
    """Keep stdout for the one JSON line: libraries (NCCL's version banner, torchrun notices) that write to file descriptor 1
    are sent to stderr for the rest of the run; emit_line() writes to the original descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit_line(obj):
    data = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tier", default="t0", choices=sorted(TIERS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the legs that run the reference on the host (profiling runs)")
    ap.add_argument("--no-extra", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    guard_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_product(args)


if __name__ == "__main__":
    main()
