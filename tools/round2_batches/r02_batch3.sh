#!/bin/bash
TAG=${1:-r03c}
O=gpurun_out
mkdir -p $O
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/launch_floor tools/micro/launch_floor.cu 2>/dev/null && /tmp/launch_floor > $O/${TAG}_launch_floor.txt 2>&1
cat $O/${TAG}_launch_floor.txt
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_step_gpu.py -x -q > $O/${TAG}_tests.log 2>&1
tail -3 $O/${TAG}_tests.log
python tools/bench_gemm_group.py "w" > $O/${TAG}_group_w.txt 2>&1
K3M_GEMM_SPLIT_MODEL=0 python tools/bench_gemm_group.py "w" > $O/${TAG}_group_w_oldsplit.txt 2>&1
cat $O/${TAG}_group_w.txt $O/${TAG}_group_w_oldsplit.txt | cut -c1-150
for v in 0 1 0 1; do
  echo "split_model=$v" >> $O/${TAG}_ab.txt
  K3M_GEMM_SPLIT_MODEL=$v python bench.py --steps 20 --warmup 3 --no-breakdown --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_per_step %.3f  e2e %.3f  host %.3f launches %d' % (d['ms_per_step'], d['e2e_ms_per_step'], d['e2e_host_ms_per_step'], d['launches_per_step']))" >> $O/${TAG}_ab.txt 2>&1
done
cat $O/${TAG}_ab.txt
python bench.py --steps 20 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
python -c "
import json
d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['e2e'], d['launches_per_step'], d['roofline']['achieved'], d['roofline']['frac'], d['roofline']['ms_per_step_in_kernel'], d['clocks'])
"
