#!/bin/bash
# ncu --set full captures of k3m_gemm_bf16 on one K3M shape, one report per epilogue variant.
# usage: tools/prof_gemm.sh "<shape note of tools/bench_gemm.py>" <tag>
set -u
SHAPE="${1:-text ffn1}"
TAG="${2:-r01}"
mkdir -p gpurun_out
for e in plain gelu dgelu drop_add; do
  epi=$e; [ $e = plain ] && epi=""
  K3M_BENCH_EPI=$epi python tools/bench_gemm.py "$SHAPE" > gpurun_out/${TAG}_plain_$e.log 2>&1 || { echo "plain run failed for $e"; cat gpurun_out/${TAG}_plain_$e.log; exit 1; }
done
for e in plain gelu dgelu drop_add; do
  epi=$e; [ $e = plain ] && epi=""
  K3M_BENCH_EPI=$epi ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 1 -c 1 \
      -f -o gpurun_out/${TAG}_gemm_$e python tools/bench_gemm.py "$SHAPE" > gpurun_out/${TAG}_ncu_$e.log 2>&1
  tail -2 gpurun_out/${TAG}_ncu_$e.log
done
cat gpurun_out/${TAG}_plain_*.log
