#!/bin/bash
# Validation trip for a change to the cooperative exponentiation: parity suite, no-alloc test alone, T1 exponentiation timings.
tag=${1:-rX}
O=gpurun_out
mkdir -p $O
timeout 200 python -m pytest tests -m gpu -x -q > $O/${tag}_gputest.log 2>&1
tail -2 $O/${tag}_gputest.log
timeout 40 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/${tag}_noalloc_alone.log 2>&1
timeout 90 python tools/pow_time.py t1 > $O/${tag}_pow_time.txt 2>&1
cat $O/${tag}_pow_time.txt
