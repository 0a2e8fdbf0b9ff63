#!/bin/bash
# Validation trip for a change to the register exponentiation: parity suite, no-alloc test alone, T0 bench line, exponentiation timings.
tag=${1:-rX}
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -m gpu -x -q > $O/${tag}_gputest.log 2>&1
tail -2 $O/${tag}_gputest.log
timeout 60 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/${tag}_noalloc_alone.log 2>&1
timeout 200 python bench.py > $O/${tag}_bench_1gpu.json 2> $O/${tag}_bench_err.log
echo "bench rc $?"
timeout 100 python tools/pow_time.py t0 > $O/${tag}_pow_time.txt 2>&1
grep "coop=0" $O/${tag}_pow_time.txt
