TAG=r03end
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -m gpu -x -q > $O/${TAG}_gpu_tests.log 2>&1
tail -2 $O/${TAG}_gpu_tests.log
grep -h "resume:\|group fp32 out" $O/${TAG}_gpu_tests.log | head
timeout 150 python bench.py --steps 20 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
python -c "
import json
d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], d['launches_per_step'], d['roofline']['achieved'], d['roofline']['frac'], d['clocks'])
"
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1
tail -1 $O/${TAG}_smoke.log
