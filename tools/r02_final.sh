#!/bin/bash
# final validation of the round: GPU suite, smoke(), the bench line, and a source-level ncu capture of the GELU epilogue
TAG=${1:-r03z}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1
tail -3 $O/${TAG}_gpu_tests.log
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > $O/${TAG}_smoke.log 2>&1
tail -1 $O/${TAG}_smoke.log
python bench.py --steps 20 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
python -c "
import json
d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], d['launches_per_step'], d['roofline']['achieved'], d['roofline']['frac'], d['roofline']['ms_per_step_in_kernel'], d['roofline']['traffic_note'], d['clocks'])
print(d['step_roofline'])
print(d.get('e2e_class_api'))
print(d.get('cpu_baseline'))
"
python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_reference.json 2>/dev/null
tail -c 600 $O/${TAG}_bench_reference.json
for e in gelu drop_add; do
  K3M_BENCH_EPI=$e timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 1 -c 1 \
      -f -o $O/${TAG}_gemm_$e python tools/bench_gemm.py "text ffn1" > $O/${TAG}_ncu_$e.log 2>&1
  ncu -i $O/${TAG}_gemm_$e.ncu-rep --page source --csv > $O/${TAG}_gemm_${e}_source.csv 2>/dev/null
  python tools/ncu_sass_hot.py $O/${TAG}_gemm_${e}_source.csv 40 > $O/${TAG}_gemm_${e}_hot.txt 2>&1
  rm -f $O/${TAG}_gemm_$e.ncu-rep $O/${TAG}_gemm_${e}_source.csv
  head -30 $O/${TAG}_gemm_${e}_hot.txt
done
