# last call of round 2: GPU suite + default bench line after the loss-ring change (trainer.py) and the bench warm-up of the lagged read
TAG=r03end2
O=gpurun_out
mkdir -p $O
timeout 200 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1
tail -1 $O/${TAG}_gpu_tests.log
timeout 120 python bench.py --steps 20 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
python -c "
import json
d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], d['launches_per_step'], d['roofline']['achieved'], d['roofline']['frac'], d['clocks'], d.get('e2e_class_api'))
"
