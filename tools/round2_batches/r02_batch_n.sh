#!/bin/bash
# multi-GPU check of the final build: data-parallel parity tool, then the bench line at N ranks
N=${1:-2}
TAG=${2:-r03n$N}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_dp_gpu.py -x -q -s > $O/${TAG}_check_dp.log 2>&1
tail -15 $O/${TAG}_check_dp.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
tail -c 1500 $O/${TAG}_bench.json | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], d.get('comm'))
" 2>&1 || tail -5 $O/${TAG}_bench.err
