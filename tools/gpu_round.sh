#!/bin/bash
# One full GPU round trip for profiles/: parity suite, bench lines, ncu launch list, ncu --set full captures of the two
# composition kernels (register path at T0, cooperative path at T1), primitive timings.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round.sh <tag>'
tag=${1:-rX}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/${tag}_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/${tag}_gputest.log 2>&1
tail -2 $O/${tag}_gputest.log
timeout 120 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/${tag}_noalloc_alone.log 2>&1
timeout 600 python bench.py > $O/${tag}_bench_1gpu.json 2> $O/${tag}_bench_err.log
echo "bench rc $?"
timeout 600 python bench.py --impl reference > $O/${tag}_bench_reference_arm.json 2>> $O/${tag}_bench_err.log
echo "ref rc $?"
timeout 600 python tools/coop_time.py t0 t1 > $O/${tag}_coop_time.txt 2>&1
timeout 600 python tools/pow_time.py t1 t0 >> $O/${tag}_coop_time.txt 2>&1
tail -12 $O/${tag}_coop_time.txt
# launch list of the bench command (times under ncu are cold-cache and serialised: shares only)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${tag}_launches.csv \
  python bench.py --steps 2 --warmup 3 > $O/${tag}_ncu_bench.log 2>&1
echo "launch list rc $?"
# full captures: the headline kernel (128x128 add at T0) and the cooperative composition at T1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_fast_compose -s 2 -c 1 -f -o $O/${tag}_fast \
  python tools/coop_prof.py t0 32768 0 > $O/${tag}_ncu_fast.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_coop_compose -s 2 -c 1 -f -o $O/${tag}_coop \
  python tools/coop_prof.py t1 32768 -1 > $O/${tag}_ncu_coop.log 2>&1
for k in fast coop; do
  ncu -i $O/${tag}_$k.ncu-rep --page details > $O/${tag}_${k}_details.txt 2>/dev/null
  ncu -i $O/${tag}_$k.ncu-rep --page raw --csv > $O/${tag}_${k}_raw.csv 2>/dev/null
  ncu -i $O/${tag}_$k.ncu-rep --page source --csv 2>/dev/null | gzip > $O/${tag}_${k}_sass.csv.gz
done
ls -la $O | tail -30
