#!/bin/bash
TAG=${1:-r03f}
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -k "relu or fuse or pool" > $O/${TAG}_tests.log 2>&1
tail -2 $O/${TAG}_tests.log
run() {
  echo "$*" >> $O/${TAG}_sweep.txt
  env "$@" python bench.py --steps 20 --warmup 3 --no-breakdown --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_per_step %.3f  e2e %.3f  launches %d' % (d['ms_per_step'], d['e2e_ms_per_step'], d['launches_per_step']))" >> $O/${TAG}_sweep.txt 2>&1
}
run K3M_BUCKET_MB=64
run K3M_BUCKET_MB=32
run K3M_BUCKET_MB=32 K3M_READY_PER_LAYER=1
run K3M_BUCKET_MB=16 K3M_READY_PER_LAYER=1
run K3M_BUCKET_MB=128
run K3M_BUCKET_MB=4000
run K3M_BUCKET_MB=64 K3M_GEMM_PDL=1
run K3M_BUCKET_MB=64
cat $O/${TAG}_sweep.txt
