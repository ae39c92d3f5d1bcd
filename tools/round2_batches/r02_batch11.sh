#!/bin/bash
TAG=${1:-r03k}
O=gpurun_out
mkdir -p $O
CMD="python bench.py --steps 1 --warmup 3 --no-breakdown --no-cpu-baseline"
for k in ln_bwd_group_sacc attn_bwd_mma ln_fwd_group; do
  timeout 300 ncu --set full --clock-control none --import-source on --graph-profiling node -k "regex:$k" -s 40 -c 1 \
      -f -o $O/${TAG}_$k $CMD > $O/${TAG}_ncu_$k.log 2>&1
  ncu -i $O/${TAG}_$k.ncu-rep --page source --csv > $O/${TAG}_${k}_source.csv 2>/dev/null
  ncu -i $O/${TAG}_$k.ncu-rep --page raw --csv > $O/${TAG}_${k}_raw.csv 2>/dev/null
  python tools/ncu_sass_hot.py $O/${TAG}_${k}_source.csv 45 > $O/${TAG}_${k}_hot.txt 2>&1
  rm -f $O/${TAG}_$k.ncu-rep $O/${TAG}_${k}_source.csv
  head -28 $O/${TAG}_${k}_hot.txt
done
