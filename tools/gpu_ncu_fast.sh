#!/bin/bash
# ncu --set full of the register composition at T0 in its two regimes: 32 768 compositions (the 128x128 add: one warp's
# latency) and 2^20 (saturated).  gpurun --timeout 400 -- 'bash tools/gpu_ncu_fast.sh <tag>'
tag=${1:-rX}
O=gpurun_out
mkdir -p $O
for n in 32768 1048576; do
  timeout 180 ncu --set full --clock-control none --import-source on -k regex:k_fast_compose -s 2 -c 1 -f -o $O/${tag}_fast_$n \
    python tools/coop_prof.py t0 $n 0 > $O/${tag}_ncu_fast_$n.log 2>&1
  echo "ncu $n rc $?"
  ncu -i $O/${tag}_fast_$n.ncu-rep --page details > $O/${tag}_fast_${n}_details.txt 2>/dev/null
  ncu -i $O/${tag}_fast_$n.ncu-rep --page raw --csv > $O/${tag}_fast_${n}_raw.csv 2>/dev/null
  ncu -i $O/${tag}_fast_$n.ncu-rep --page source --csv 2>/dev/null | gzip > $O/${tag}_fast_${n}_sass.csv.gz
  rm -f $O/${tag}_fast_$n.ncu-rep
  grep -m3 "k_fast_compose\|Duration\|Executed Ipc Active" $O/${tag}_fast_${n}_details.txt
done
du -sh $O
