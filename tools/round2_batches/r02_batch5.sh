#!/bin/bash
TAG=${1:-r03e}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1
tail -3 $O/${TAG}_gpu_tests.log
python bench.py --steps 20 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
python -c "
import json
d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['e2e']['ms_per_step'], d['launches_per_step'], d['roofline']['achieved'], d['roofline']['frac'], d['roofline']['ms_per_step_in_kernel'], d['clocks'])
for r in d['hbm_kernels']['rows']: print(r['kernel'], r['launches'], r['ms_serial'], r['frac_of_hbm_peak'])
"
