#!/bin/bash
TAG=${1:-r03i}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "gemm or xent" > $O/${TAG}_gemm_tests.log 2>&1
tail -3 $O/${TAG}_gemm_tests.log
for e in plain gelu dgelu drop_add; do
  epi=$e; [ $e = plain ] && epi=""
  for s in "text ffn1" "text attn-out"; do
    K3M_BENCH_EPI=$epi python tools/bench_gemm.py "$s" 2>&1 | head -1 | sed "s/^/$e: /" >> $O/${TAG}_epi.txt
  done
done
cat $O/${TAG}_epi.txt
python tools/bench_gemm_group.py > $O/${TAG}_group.txt 2>&1
tail -1 $O/${TAG}_group.txt
for v in 1 2; do
  python bench.py --steps 20 --warmup 3 --no-breakdown --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_per_step %.3f  e2e %.3f  launches %d' % (d['ms_per_step'], d['e2e_ms_per_step'], d['launches_per_step']))" >> $O/${TAG}_ab.txt 2>&1
done
cat $O/${TAG}_ab.txt
