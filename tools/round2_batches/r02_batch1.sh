#!/bin/bash
# Round-2 measurement batch (one gpurun call): mirrored tile walk A/B on the step's GEMM groups, epilogue cost per
# variant on the K = 768 FFN shape, a source-level ncu capture of the GELU epilogue, ncu --set full of the row-wise
# backward kernels of one step, the bench line and the GPU suite.  Writes gpurun_out/${TAG}_*.
TAG=${1:-r03a}
O=gpurun_out
mkdir -p $O
K3M_GEMM_SNAKE=0 python tools/bench_gemm_group.py > $O/${TAG}_group_snake0.txt 2>&1
K3M_GEMM_SNAKE=1 python tools/bench_gemm_group.py > $O/${TAG}_group_snake1.txt 2>&1
tail -1 $O/${TAG}_group_snake0.txt $O/${TAG}_group_snake1.txt
for e in plain gelu dgelu drop_add; do
  epi=$e; [ $e = plain ] && epi=""
  for s in "text ffn1" "text attn-out" "co dffn2"; do
    K3M_BENCH_EPI=$epi python tools/bench_gemm.py "$s" 2>&1 | head -1 | sed "s/^/$e: /" >> $O/${TAG}_epi.txt
  done
done
cat $O/${TAG}_epi.txt
for e in gelu drop_add; do
  K3M_BENCH_EPI=$e timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 1 -c 1 \
      -f -o $O/${TAG}_gemm_$e python tools/bench_gemm.py "text ffn1" > $O/${TAG}_ncu_$e.log 2>&1
  ncu -i $O/${TAG}_gemm_$e.ncu-rep --page source --csv > $O/${TAG}_gemm_${e}_source.csv 2>/dev/null
  ncu -i $O/${TAG}_gemm_$e.ncu-rep --page raw --csv > $O/${TAG}_gemm_${e}_raw.csv 2>/dev/null
  python tools/ncu_sass_hot.py $O/${TAG}_gemm_${e}_source.csv 60 > $O/${TAG}_gemm_${e}_hot.txt 2>&1
  rm -f $O/${TAG}_gemm_$e.ncu-rep $O/${TAG}_gemm_${e}_source.csv
done
for v in "0 0" "1 0" "1 1"; do
  set -- $v
  echo "snake=$1 defer=$2" >> $O/${TAG}_ab.txt
  K3M_GEMM_SNAKE=$1 K3M_DEFER_GRADS=$2 python bench.py --steps 20 --warmup 3 --no-breakdown --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_per_step %.3f  e2e %.3f  launches %d' % (d['ms_per_step'], d['e2e_ms_per_step'], d['launches_per_step']))" >> $O/${TAG}_ab.txt 2>&1
done
cat $O/${TAG}_ab.txt
python bench.py --steps 20 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
tail -c 600 $O/${TAG}_bench.json
CMD="python bench.py --steps 1 --warmup 3 --no-breakdown --no-cpu-baseline"
for k in ln_bwd attn_bwd; do
  timeout 300 ncu --set full --clock-control none --import-source on --graph-profiling node -k "regex:$k" -s 60 -c 2 \
      -f -o $O/${TAG}_$k $CMD > $O/${TAG}_ncu_$k.log 2>&1
  ncu -i $O/${TAG}_$k.ncu-rep --page raw --csv > $O/${TAG}_${k}_raw.csv 2>/dev/null
  ncu -i $O/${TAG}_$k.ncu-rep --page source --csv > $O/${TAG}_${k}_source.csv 2>/dev/null
  python tools/ncu_sass_hot.py $O/${TAG}_${k}_source.csv 40 > $O/${TAG}_${k}_hot.txt 2>&1
  rm -f $O/${TAG}_$k.ncu-rep $O/${TAG}_${k}_source.csv
done
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1
tail -5 $O/${TAG}_gpu_tests.log
ls -la $O | tail -20
