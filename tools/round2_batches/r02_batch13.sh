#!/bin/bash
TAG=${1:-r03m}
O=gpurun_out
mkdir -p $O
K3M_LN_BWD_TILE=1 K3M_LN_BWD_TILE_CTAS=3 timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "ln" > $O/${TAG}_ln_tests.log 2>&1
tail -2 $O/${TAG}_ln_tests.log
run() {
  echo "$*" >> $O/${TAG}_ab.txt
  env "$@" python bench.py --steps 20 --warmup 3 --no-breakdown --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_per_step %.3f  e2e %.3f  launches %d' % (d['ms_per_step'], d['e2e_ms_per_step'], d['launches_per_step']))" >> $O/${TAG}_ab.txt 2>&1
}
run K3M_LN_BWD_TILE=0
run K3M_LN_BWD_TILE=1 K3M_LN_BWD_TILE_CTAS=3
run K3M_LN_BWD_TILE=1 K3M_LN_BWD_TILE_CTAS=2
run K3M_LN_BWD_TILE=0
run K3M_LN_BWD_TILE=1 K3M_LN_BWD_TILE_CTAS=3
run K3M_LN_BWD_TILE=1 K3M_LN_BWD_TILE_CTAS=6
cat $O/${TAG}_ab.txt
