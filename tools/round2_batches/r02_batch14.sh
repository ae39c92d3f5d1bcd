#!/bin/bash
TAG=${1:-r03n}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "gemm" > $O/${TAG}_gemm_tests.log 2>&1
tail -2 $O/${TAG}_gemm_tests.log
python tools/bench_gemm_group.py > $O/${TAG}_group_facc.txt 2>&1
K3M_GEMM_NO_FACC=1 python tools/bench_gemm_group.py > $O/${TAG}_group_generic.txt 2>&1
paste -d'\n' $O/${TAG}_group_facc.txt $O/${TAG}_group_generic.txt | cut -c1-62 
for v in 1 0 1 0; do
  echo "no_facc=$v" >> $O/${TAG}_ab.txt
  K3M_GEMM_NO_FACC=$v python bench.py --steps 20 --warmup 3 --no-breakdown --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_per_step %.3f  e2e %.3f  launches %d' % (d['ms_per_step'], d['e2e_ms_per_step'], d['launches_per_step']))" >> $O/${TAG}_ab.txt 2>&1
done
cat $O/${TAG}_ab.txt
