"""bench.py -- CoFHE ciphertext hot path on B200: ciphertext adds/s for `ciphertext_matadd` 128x128 (BASELINE.json
configs[1], reference benchmarks/local.cpp:65-138 at 128x128) plus, in `extra`, form exponentiations/s for the
threshold / scal paths.  One process per GPU; ranks own independent 128x128 ciphertext matrices (weak scaling, no
data-path collective: SURVEY.md section 8e).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--tier t0|t1]

A step = one add_ciphertext_tensors pass (res <- res + ct2) over the rank's matrix.  `value` is timed with CUDA
events per step on the launching stream, inputs resident in HBM, L2 flushed between steps; `e2e` goes through the
host-pointer C-ABI entry point (pinned host buffers, H2D + kernel + D2H inside the timed region).
The oracle / compiled reference (oracle/) is used here only for the `cpu_baseline` leg and `--impl reference`.
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

TIERS = {"t0": ("ref_t0.json", 128, 128), "t1": ("ref_t1.json", 1792, 128)}
ROWS = COLS = 128


def fixture(tier):
    with open(os.path.join(ROOT, "tests", "golden", TIERS[tier][0])) as fh:
        t = json.load(fh)
    conv = lambda x: tuple(int(y) for y in x)
    return {"tier": tier, "N": int(t["params"]["N"]), "Delta": int(t["params"]["Delta"]), "h": conv(t["params"]["h"]), "pk": conv(t["pk"]),
            "bound": int(t["params"]["exponent_bound"]), "sk": int(t["sk"]), "k": int(t["params"]["k"])}


def limb_s(dbits):
    """s of SURVEY.md section 8d: 32-bit limbs of a reduced coefficient."""
    return ((dbits + 1) // 2 + 31) // 32


def w_comp(s):
    return 16 * s * s


def w_dupl(s):
    return 14 * s * s


def w_exp(s, e):
    return (e - 1) * w_dupl(s) + -(-e // 3) * w_comp(s)


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


def cpu_baseline_matadd(tier, G=None, iters=50):
    """The UNMODIFIED reference (oracle/_ref) on the host cores: config 2 exactly (128x128, `iters` chained adds),
    all OpenMP threads, then 1 thread on a shorter chain.  When a GPU group is given, the GPU repeats the same chain
    on the reference's own inputs and the final ciphertexts are compared bit for bit."""
    from oracle import ref as R
    _, nbits, k = TIERS[tier]
    ref = R.Reference(nbits, k)
    ref.call("keygen")
    ref.set_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
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
        import torch
        from cofhe_b200.group import to_device, to_host
        a, b = to_device(bufs[0], G.device), to_device(bufs[1], G.device)
        res = G.add_ciphertexts(a, b)
        for _ in range(iters - 1):
            res = G.add_ciphertexts(res, b)
        G.status()
        out["parity"] = bool((to_host(res) == bufs[2]).all())
    ref.close()
    return out


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    from oracle import ref as R
    if not R.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libcofhe_ref.so (compiled reference) missing or libgmp.so.10 not loadable"}))
        return
    _, nbits, k = TIERS[args.tier]
    ref = R.Reference(nbits, k)
    ref.call("keygen")
    # all the host threads this process may use: torchrun exports OMP_NUM_THREADS=1, which would throttle the reference
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    ref.set_threads(cores)
    cores = ref.threads()
    if args.warmup:
        ref.bulk_matadd(ROWS, COLS, args.warmup)
    dt, _ = ref.bulk_matadd(ROWS, COLS, args.steps)
    n = ROWS * COLS
    v = n * args.steps / dt
    line = {"impl": "reference", "metric": "ciphertext adds/sec (ciphertext_matadd 128x128)", "value": v, "unit": "ct_adds/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32 limbs (multi-precision integer)",
            "data": "synthetic: encrypt_tensor(pt[i]=i+1), default-seeded RandGen, as benchmarks/local.cpp",
            "config": {"workload": "ciphertext_matadd 128x128, tier %s (N %d bits, k %d)" % (args.tier, nbits, k),
                       "note": "reference CPUCryptoSystem::add_ciphertext_tensors (BICYCL+GMP+OpenMP) on host cores; one matrix regardless of --gpus"},
            "cpu_baseline": {"value": v, "unit": "ct_adds/s", "cores": cores, "kind": "reference",
                             "sample": "%d chained full 128x128 passes" % args.steps},
            "e2e": {"value": v, "unit": "ct_adds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def synth_cts(G, fx, n, seed):
    """n synthetic ciphertexts (c1 = h^r, c2 = pk^r'), random r < exponent bound, computed on the GPU itself."""
    import torch
    rng = random.Random(seed)
    fl = G.fl
    out = torch.empty((2, fl, n), dtype=torch.int32, device=G.device)
    for j, base in enumerate((fx["h"], fx["pk"])):
        seeds = G.forms([base] * n)
        out[j] = G.nupow(seeds, exps=[rng.randrange(1, fx["bound"]) for _ in range(n)])
    G.status()
    return out


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
    dbits = (-fx["Delta"]).bit_length()
    s = limb_s(dbits)
    n = ROWS * COLS
    ct1, ct2 = synth_cts(G, fx, n, 1000 + rank), synth_cts(G, fx, n, 2000 + rank)
    res = torch.empty_like(ct1)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=G.device)  # > 126 MB L2
    lib, h = G.lib, G.h

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
    extra = None if args.no_extra else extras(G, fx, s, world, rank, imad.value)
    if rank == 0:
        emit(args, G, fx, s, world, n, kern_ms, e2e_ms, e2e_steps, launches, clocks, h1.nbytes, imad.value, extra)
    if world > 1:
        dist.destroy_process_group()


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


def extras(G, fx, s, world, rank, imad_peak):
    """The other BASELINE.json configs, weak-scaled (fixed work per GPU, results gathered over NCCL where the config
    says so): scal_matmul rows (config 5: 256x256 . 256x256 at 8 GPUs = 32 ct rows per GPU), threshold part_decrypt +
    combine (config 4), per-element-r encryption (config 3).  Every rank runs them; times are max over ranks."""
    import torch
    from conftest_free import golden_zero
    from cofhe_b200 import shard as S
    out = {}
    k = fx["k"]
    G.set_scheme(k, fx["N"])
    G.set_bases(fx["h"], fx["pk"], fx["bound"].bit_length())
    rng = random.Random(4242)
    # ---- config 5: scalar x ciphertext matmul, row-sharded, one all-gather of the result rows
    rows, m, p = 32, 256, 256
    ct = synth_cts(G, fx, rows * m, 31 + rank)
    zero = G.cts([golden_zero(fx["tier"])])
    for label, sc in (("pattern i+1 (<= 17 bit, as benchmarks/local.cpp)", [i + 1 for i in range(m * p)]),
                      ("uniform %d-bit" % k, [rng.getrandbits(k) for _ in range(m * p)])):
        ebits = max(x.bit_length() for x in sc)
        scd = G.ints(sc, (ebits + 31) // 32)
        parts = [(r * rows, (r + 1) * rows) for r in range(world)]

        def run():
            local = G.scal_matmul(ct, scd, rows, m, p, zero, ebits=ebits)
            return S.gather_last(local, [rows * p] * world)
        dt = _timed(run, world)
        G.status()
        units = 2 * world * rows * m * p
        macs = units * (w_exp(s, ebits) + w_comp(s)) / dt
        out.setdefault("scal_matmul", []).append({
            "workload": "%dx%d . %dx%d, %s, row-sharded %d rows/GPU + NCCL gather" % (world * rows, m, m, p, label, rows),
            "form_exps_per_s": units / dt, "seconds": dt, "imad_frac_nominal": macs / imad_peak / world})
    # ---- config 4: threshold decryption, element-sharded; config 3: encryption with per-element randomness
    n = 1 << 13
    ms = G.ints([rng.getrandbits(k) for _ in range(n)], G.ml)
    rs = G.ints([rng.randrange(1, fx["bound"]) for _ in range(n)], G.el_r)
    holder = {}

    def enc():
        holder["ct"] = G.encrypt(ms, rs)
    dt = _timed(enc, world)
    out["encrypt"] = {"workload": "%d ciphertexts/GPU, own r per ciphertext (2 fixed-base exps + f^m + 1 comp each)" % n,
                      "encryptions_per_s": world * n / dt, "fixed_base_form_exps_per_s": 2 * world * n / dt,
                      "imad_frac_nominal": 2 * world * n * w_exp(s, fx["bound"].bit_length()) / dt / imad_peak / world}
    share = fx["sk"]

    def pdec():
        holder["d"] = G.part_decrypt(holder["ct"], share)
    dt = _timed(pdec, world)
    out["part_decrypt"] = {"workload": "%d ciphertexts/GPU, shared %d-bit exponent" % (n, share.bit_length()),
                           "form_exps_per_s": world * n / dt,
                           "imad_frac_nominal": world * n * w_exp(s, share.bit_length()) / dt / imad_peak / world}

    def dec():
        holder["m"] = G.decrypt(holder["ct"], share)
    dt = _timed(dec, world, reps=1)
    out["decrypt"] = {"workload": "%d ciphertexts/GPU (c1^sk, nucompinv, dlog_F)" % n, "decryptions_per_s": world * n / dt,
                      "roundtrip_ok": bool(torch.equal(holder["m"], ms))}
    G.status()
    # ---- section 8f-2: Beaver-triple ciphertext x ciphertext multiplication (device side, without the network round)
    try:
        A, B = holder["ct"], G.encrypt(ms, 4242)
        C = G.add_ciphertexts(A, B)  # any ciphertext batch serves as c for timing; correctness is in tests/test_gpu_scheme.py
        hb = {}
        def bopen():
            hb["o"] = G.beaver_open(holder["ct"], B, A, B)
        t_open = _timed(bopen, world, reps=1)
        def bclose():
            hb["r"] = G.beaver_close(ms, ms, A, B, C, 31337)
        t_close = _timed(bclose, world, reps=1)
        out["beaver"] = {"workload": "%d ciphertext pairs/GPU: open (2 negate + 2 add) + close (2 scal + encrypt_tensor + 3 add), k = %d bit exponents; "
                                     "the threshold-decryption round between them is not included" % (n, k),
                         "open_ms": t_open * 1e3, "close_ms": t_close * 1e3, "ct_mults_per_s": world * n / (t_open + t_close),
                         "form_exps_per_s": world * n * 8 / (t_open + t_close)}
        G.status()
    except Exception as e:
        out["beaver"] = {"error": str(e)[:120]}
    # ---- the reference's own CPU code on bounded samples of the same workloads, same run (rank 0 of a 1-GPU run)
    if world == 1:
        out["reference_cpu"] = reference_extras(G, fx)
    # ---- section 8f-1: the reference's binary tensor wire format <-> SoA on the device (128x128 ciphertexts)
    out["wire"] = wire_extra(G, fx, world, rank)
    return out


def reference_extras(G, fx):
    """Bounded samples of configs 3-5 on the UNMODIFIED reference (oracle/_ref) with all host threads, and the GPU on the
    reference's own inputs where the outputs are comparable (encryption with the reference's (m, r), part-decryption)."""
    import numpy as np
    import torch
    try:
        from oracle import ref as R
        if not R.available():
            return {"unavailable": "oracle/_ref not loadable"}
        _, nbits, k = TIERS[fx["tier"]]
        ref = R.Reference(nbits, k)
        ks = ref.call("keygen")
        sk, pk = int(ks[0]), tuple(int(x) for x in ks[1:4])
        cores = len(os.sched_getaffinity(0))
        ref.set_threads(cores)
        res = {"cores": cores, "kind": "reference"}
        dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int32)).to(G.device)
        # config 5a: scalar x ciphertext matmul, 2 ct rows of the 256-wide problem, benchmark scalars i+1
        rows, m, p = 2, 256, 256
        dt, _, _ = ref.bulk_scal_matmul(rows, m, p)
        res["scal_matmul"] = {"sample": "%dx%d . %dx%d, pattern i+1, one call" % (rows, m, m, p), "seconds": dt,
                              "form_exps_per_s": 2 * rows * m * p / dt}
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
    except Exception as e:
        return {"error": str(e)[:160]}


def wire_extra(G, fx, world, rank):
    """serialize_ciphertext_tensor / deserialize_ciphertext_tensor of a 128x128 tensor as device kernels (wire bytes and SoA
    batch both resident in HBM), bytes/s of the wire buffer.  On rank 0 of a 1-GPU run the compiled reference does the
    same tensor on the host cores and its bytes are the parity check."""
    import ctypes
    import numpy as np
    import torch
    from cofhe_b200 import capi
    n = ROWS * COLS
    ct = synth_cts(G, fx, n, 7)
    ref_line = None
    if world == 1:
        try:
            from oracle import ref as R
            if R.available():
                _, nbits, k = TIERS[fx["tier"]]
                ref = R.Reference(nbits, k)
                ref.call("keygen")
                ref.set_threads(len(os.sched_getaffinity(0)))
                ds, dd, cts, wire = ref.bulk_wire(ROWS, COLS, 3, G.na, G.nc)
                ref.close()
                dev = torch.from_numpy(cts.view(np.int32)).to(G.device)
                mine = G.cts_to_wire([ROWS, COLS], dev)
                shape, back = G.wire_to_cts(wire)
                G.status()
                ref_line = {"serialize_s": ds / 3, "deserialize_s": dd / 3, "bytes": len(wire), "cores": len(os.sched_getaffinity(0)),
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


def emit(args, G, fx, s, world, n, kern_ms, e2e_ms, e2e_steps, launches, clocks, ct_bytes, imad_peak, extra):
    peak_gbs, peak_src = measured_peaks()
    ms_step = kern_ms / args.steps
    value = world * n * args.steps / (kern_ms * 1e-3)
    alg_bytes = 3 * ct_bytes                       # read ct1, ct2; write res -- in this build's limb layout
    # bytes that cross PCIe towards the device per step: on the register-resident tiers the host entry point reads only
    # the a and b rows of both operands (c is recomputed on the device); all rows of the result travel back
    if G.fast:
        h2d_bytes = 2 * ct_bytes * (2 * G.na) // G.fl
        h2d_note = "a and b rows only: %d of %d limb rows per form" % (2 * G.na, G.fl)
    else:
        h2d_bytes, h2d_note = 2 * ct_bytes, "all limb rows, staged"
    achieved = alg_bytes / (ms_step * 1e-3) / 1e9
    imad = ctypes.c_double(imad_peak)
    macs = 2 * n * w_comp(s) / (ms_step * 1e-3)
    line = {"metric": "ciphertext adds/sec (ciphertext_matadd 128x128)", "value": value, "unit": "ct_adds/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32 limbs (multi-precision integer)",
            "data": "synthetic: c1 = h^r, c2 = pk^r' for random r below the encryption exponent bound, generated on the GPU",
            "config": {"workload": "ciphertext_matadd 128x128 per GPU, tier %s (|Delta| %d bits, s=%d limbs)" % (args.tier, (-fx["Delta"]).bit_length(), s),
                       "l2": "flushed between steps (256 MiB memset); per-step CUDA events on the launch stream, summed",
                       "sharding": "one independent matrix per rank, no collective"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs,
                         "traffic": ncu_traffic("k_fast_compose"), "peak_source": peak_src, "kernel": "k_fast_compose",
                         "alg_bytes_per_launch": alg_bytes},
            "roofline_imad": {"bound": "imad", "achieved": macs / 1e12, "peak": imad.value / 1e12, "unit": "T MAC/s (32x32->64)",
                              "frac": macs / imad.value, "model": "2*W_comp = 32*s^2 MACs per ct-add (SURVEY.md 8d)",
                              "peak_source": "measured in this run (cofhe_b200_imad_probe, IMAD.WIDE.U32 chains on all SMs)"},
            "e2e": {"value": world * n * e2e_steps / (e2e_ms * 1e-3), "unit": "ct_adds/s", "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": ct_bytes,
                    "api": "cofhe_b200_add_ciphertexts_host: blocking call on page-locked host SoA buffers; the kernels read "
                           "the operands from and write the result to host memory over PCIe inside the call (%s)" % h2d_note},
            "gpu_launches": launches, "clocks": clocks}
    if world == 1 and not args.no_cpu:
        from oracle import ref as R
        if R.available():
            line["cpu_baseline"] = cpu_baseline_matadd(args.tier, G)
        else:
            line["cpu_baseline"] = {"value": None, "unit": "ct_adds/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not loadable"}
    if extra is not None:
        line["extra"] = extra
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tier", default="t0", choices=sorted(TIERS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-extra", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_product(args)


if __name__ == "__main__":
    main()
