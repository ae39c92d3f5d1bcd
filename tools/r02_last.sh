#!/bin/bash
# last GPU call of round 2 (about 10 GPU-minutes were left): GPU suite, bench line, the grouped-fp32-store reproducer, smoke()
TAG=${1:-r03last}
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1
tail -2 $O/${TAG}_gpu_tests.log
timeout 150 python bench.py --steps 20 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
python -c "
import json
d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], d['launches_per_step'], d['roofline']['achieved'], d['roofline']['frac'], d['clocks'])
"
K3M_GEMM_GROUPED_F32_STORE=1 timeout 60 python tools/debug_group_f32.py > $O/${TAG}_group_f32.txt 2>&1
K3M_GEMM_GROUPED_F32_STORE=1 K3M_GEMM_NO_TMA_STORE=1 timeout 60 python tools/debug_group_f32.py tiny two_equal > $O/${TAG}_group_f32_direct.txt 2>&1
K3M_GEMM_GROUPED_F32_STORE=1 K3M_GEMM_SNAKE=0 timeout 60 python tools/debug_group_f32.py tiny two_equal > $O/${TAG}_group_f32_nosnake.txt 2>&1
cat $O/${TAG}_group_f32.txt | head -60
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1
tail -1 $O/${TAG}_smoke.log
