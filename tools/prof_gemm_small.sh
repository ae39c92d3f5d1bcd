#!/bin/bash
# ncu --set full capture of k3m_gemm_bf16 (plain epilogue) and of the cuBLAS kernel next to it on the latency-bound
# shapes of tools/bench_gemm.py; raw pages exported as CSV.  usage: tools/prof_gemm_small.sh <tag>
set -u
TAG="${1:-r01}"
mkdir -p gpurun_out
IFS='|' read -ra LIST <<< "${SHAPES:-tiny|co small|co ffn2|co wffn1}"
for s in "${LIST[@]}"; do
  f=$(echo "$s" | tr ' ' '_')
  python tools/bench_gemm.py "$s" > gpurun_out/${TAG}_plain_$f.log 2>&1 || { cat gpurun_out/${TAG}_plain_$f.log; exit 1; }
  ncu --set full --clock-control none --import-source on -s 2 -c 1 -k regex:gemm_kernel -f -o gpurun_out/${TAG}_small_$f \
      python tools/bench_gemm.py "$s" > gpurun_out/${TAG}_ncu_$f.log 2>&1
  ncu -i gpurun_out/${TAG}_small_$f.ncu-rep --page raw --csv > gpurun_out/${TAG}_small_$f.csv 2>/dev/null
  [ -n "${KEEP_REP:-}" ] || rm -f gpurun_out/${TAG}_small_$f.ncu-rep      # gpurun_out/ travels back only under 64 MiB
  # the cuBLAS kernel of the same shape: duration only
  ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_launches_$f.csv \
      python tools/bench_gemm.py "$s" > /dev/null 2>&1
  cat gpurun_out/${TAG}_plain_$f.log
done
