#!/bin/bash
# The round-closing GPU trip: tools/gpu_round.sh (parity suite, bench lines, launch list, ncu --set full of the two
# composition kernels, primitive timings) + the T1 bench line + ncu --set full of the two other hot kernels at T0
# (k_mm_accum of the benched matmul shard, k_fast_pow at 8 192 shared-exponent exponentiations).
#   gpurun --timeout 1700 -- 'bash tools/gpu_final.sh <tag>'
tag=${1:-rX}
O=gpurun_out
bash tools/gpu_round.sh $tag
rm -f $O/${tag}_fast.ncu-rep $O/${tag}_coop.ncu-rep   # gpurun_out/ comes back only below 64 MiB: the exported pages stay
timeout 600 python bench.py --tier t1 --steps 20 > $O/${tag}_bench_1gpu_t1.json 2>> $O/${tag}_bench_err.log
echo "t1 bench rc $?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_mm_accum -c 6 -f -o $O/${tag}_mm \
  python tools/mm_time.py t0 32,256,256 > $O/${tag}_ncu_mm.log 2>&1
echo "ncu mm rc $?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_fast_pow -s 7 -c 1 -f -o $O/${tag}_pow \
  python tools/pow_time.py t0 > $O/${tag}_ncu_pow.log 2>&1
echo "ncu pow rc $?"
for k in mm pow; do
  ncu -i $O/${tag}_$k.ncu-rep --page details > $O/${tag}_${k}_details.txt 2>/dev/null
  ncu -i $O/${tag}_$k.ncu-rep --page raw --csv > $O/${tag}_${k}_raw.csv 2>/dev/null
  ncu -i $O/${tag}_$k.ncu-rep --page source --csv 2>/dev/null | gzip > $O/${tag}_${k}_sass.csv.gz
  rm -f $O/${tag}_$k.ncu-rep
done
du -sh $O
ls -la $O | tail -40
