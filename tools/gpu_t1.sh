mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_coop.py -m gpu -x -q -k "tables" > gpurun_out/r2m_tab.log 2>&1; tail -25 gpurun_out/r2m_tab.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2m_gputest.log 2>&1; tail -5 gpurun_out/r2m_gputest.log
timeout 600 python tools/pow_time.py t1 > gpurun_out/r2m_pow_time.txt 2>&1; tail -8 gpurun_out/r2m_pow_time.txt
timeout 900 python bench.py --tier t1 --steps 20 > gpurun_out/r2m_bench_1gpu_t1.json 2> gpurun_out/r2m_err.log; echo rc $?; tail -3 gpurun_out/r2m_err.log
python - <<'P'
import json
d=json.loads(open("gpurun_out/r2m_bench_1gpu_t1.json").read().strip().splitlines()[-1])
print("t1 add", d["value"], d["ms_per_step"], "imad", d["roofline_imad"]["frac"])
fe=d.get("form_exps") or {}
print("t1 matmul", fe.get("value"), fe.get("parity"), fe.get("error"), (fe.get("roofline_imad") or {}).get("frac"))
x=d.get("extra") or {}
print({k:(v.get("encryptions_per_s") or v.get("threshold_decryptions_per_s") or v.get("error")) for k,v in x.items() if isinstance(v,dict)})
P
timeout 600 python tools/coop_time.py t1 > gpurun_out/r2m_coop_time.txt 2>&1; tail -22 gpurun_out/r2m_coop_time.txt
