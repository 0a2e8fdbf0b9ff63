#!/bin/bash
# One GPU round trip while iterating: parity tests, the bench line (short form) and the primitive timings.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/quick_bench.json 2>gpurun_out/err.log
python - <<'P'
import json
d = json.load(open("gpurun_out/quick_bench.json"))
x = d.get("extra") or {}
print("add %.1f M/s  e2e %.1f M/s  %.4f ms/step" % (d["value"] / 1e6, d["e2e"]["value"] / 1e6, d["ms_per_step"]))
if x:
    print("matmul", [round(m["form_exps_per_s"] / 1e6, 1) for m in x["scal_matmul"]], "enc %.2f M/s" % (x["encrypt"]["encryptions_per_s"] / 1e6),
          "pdec %.2f M/s" % (x["part_decrypt"]["form_exps_per_s"] / 1e6), "dec %.2f M/s" % (x["decrypt"]["decryptions_per_s"] / 1e6))
P
python tools/quick_time.py t0 2>&1 | tail -6
