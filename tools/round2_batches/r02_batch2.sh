#!/bin/bash
# kernel tests + GEMM epilogue micro + grouped micro + bench (timing only, then full line)
TAG=${1:-r03b}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "gemm or epilogue or group" > $O/${TAG}_gemm_tests.log 2>&1
tail -3 $O/${TAG}_gemm_tests.log
for e in plain gelu dgelu drop_add; do
  epi=$e; [ $e = plain ] && epi=""
  for s in "text ffn1" "text attn-out" "co dffn2"; do
    K3M_BENCH_EPI=$epi python tools/bench_gemm.py "$s" 2>&1 | head -1 | sed "s/^/$e: /" >> $O/${TAG}_epi.txt
  done
done
cat $O/${TAG}_epi.txt
python tools/bench_gemm_group.py > $O/${TAG}_group.txt 2>&1
tail -1 $O/${TAG}_group.txt
python bench.py --steps 20 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
python -c "
import json
d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['e2e']['ms_per_step'], d['launches_per_step'], d['roofline']['achieved'], d['roofline']['frac'], d['roofline']['ms_per_step_in_kernel'], d['clocks'])
"
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1
tail -3 $O/${TAG}_gpu_tests.log
