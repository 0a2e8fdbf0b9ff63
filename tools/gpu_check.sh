#!/bin/bash
# Short validation trip for one kernel change: parity suite, the no-alloc test alone, bench lines (T0, then T1), primitive timings.
#   gpurun --timeout 600 -- 'bash tools/gpu_check.sh <tag>'
tag=${1:-rX}
O=gpurun_out
mkdir -p $O
timeout 400 python -m pytest tests -m gpu -x -q > $O/${tag}_gputest.log 2>&1
tail -2 $O/${tag}_gputest.log
timeout 60 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/${tag}_noalloc_alone.log 2>&1
timeout 300 python bench.py > $O/${tag}_bench_1gpu.json 2> $O/${tag}_bench_err.log
echo "bench rc $?"
timeout 200 python tools/coop_time.py t0 t1 > $O/${tag}_coop_time.txt 2>&1
grep "coop=0 \|coop=-1" $O/${tag}_coop_time.txt | head -20
timeout 300 python bench.py --tier t1 --steps 20 > $O/${tag}_bench_1gpu_t1.json 2>> $O/${tag}_bench_err.log
echo "t1 bench rc $?"
