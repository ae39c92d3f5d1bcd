#!/bin/bash
TAG=${1:-r03d}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "gemm or epilogue or group" > $O/${TAG}_gemm_tests.log 2>&1
tail -3 $O/${TAG}_gemm_tests.log
python tools/bench_gemm_group.py > $O/${TAG}_group_3d.txt 2>&1
K3M_GEMM_NO_3D=1 python tools/bench_gemm_group.py > $O/${TAG}_group_2d.txt 2>&1
paste -d'\n' $O/${TAG}_group_3d.txt $O/${TAG}_group_2d.txt | cut -c1-100
grep -o "mn3d=[01]*" $O/${TAG}_group_3d.txt | sort | uniq -c
python tools/bench_gemm.py > $O/${TAG}_gemm_3d.txt 2>&1
K3M_GEMM_NO_3D=1 python tools/bench_gemm.py > $O/${TAG}_gemm_2d.txt 2>&1
paste -d'\n' $O/${TAG}_gemm_3d.txt $O/${TAG}_gemm_2d.txt | cut -c1-110
for v in 1 0 1 0; do
  echo "no_3d=$v" >> $O/${TAG}_ab.txt
  K3M_GEMM_NO_3D=$v python bench.py --steps 20 --warmup 3 --no-breakdown --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_per_step %.3f  e2e %.3f  launches %d' % (d['ms_per_step'], d['e2e_ms_per_step'], d['launches_per_step']))" >> $O/${TAG}_ab.txt 2>&1
done
cat $O/${TAG}_ab.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1
tail -3 $O/${TAG}_gpu_tests.log
