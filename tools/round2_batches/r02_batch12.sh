#!/bin/bash
TAG=${1:-r03l}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "ln or layer_norm or LayerNorm" > $O/${TAG}_ln_tests.log 2>&1
tail -4 $O/${TAG}_ln_tests.log
for v in 0 1 0 1; do
  echo "ln_bwd_tile=$v" >> $O/${TAG}_ab.txt
  K3M_LN_BWD_TILE=$v python bench.py --steps 20 --warmup 3 --no-breakdown --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_per_step %.3f  e2e %.3f  launches %d' % (d['ms_per_step'], d['e2e_ms_per_step'], d['launches_per_step']))" >> $O/${TAG}_ab.txt 2>&1
done
cat $O/${TAG}_ab.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1
tail -3 $O/${TAG}_gpu_tests.log
