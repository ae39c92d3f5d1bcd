#!/bin/bash
TAG=${1:-r03g}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "gemm" > $O/${TAG}_gemm_tests.log 2>&1
tail -3 $O/${TAG}_gemm_tests.log
for e in gelu dgelu; do
  K3M_BENCH_EPI=$e python tools/bench_gemm.py "text ffn1" 2>&1 | head -1 | sed "s/^/$e: /" >> $O/${TAG}_epi.txt
done
cat $O/${TAG}_epi.txt
for v in 0 1 0 1; do
  echo "gelu_grad=$v" >> $O/${TAG}_ab.txt
  K3M_GELU_GRAD=$v python bench.py --steps 20 --warmup 3 --no-breakdown --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_per_step %.3f  e2e %.3f  launches %d' % (d['ms_per_step'], d['e2e_ms_per_step'], d['launches_per_step']))" >> $O/${TAG}_ab.txt 2>&1
done
cat $O/${TAG}_ab.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1
tail -3 $O/${TAG}_gpu_tests.log
python bench.py --steps 20 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
python -c "
import json
d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['e2e']['ms_per_step'], d['launches_per_step'], d['roofline']['achieved'], d['roofline']['frac'], d['roofline']['ms_per_step_in_kernel'], d['clocks'])
"
