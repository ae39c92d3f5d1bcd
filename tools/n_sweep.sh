#!/bin/bash
# data-parallel variants of the step at N GPUs (timing only): tools/n_sweep.sh N out.txt [variant ...]
N=${1:-2}
out=${2:-gpurun_out/n_sweep.txt}
shift; shift
: > $out
run() {
  echo "== $*" >> $out
  timeout 300 env $* python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 3 --no-breakdown 2>>$out.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_per_step %.3f  e2e %.3f  value %.0f  comm %s' % (d['ms_per_step'], d['e2e_ms_per_step'], d['value'], json.dumps({k: v for k, v in (d.get('comm') or {}).items() if k not in ('note',)})))" >> $out 2>&1
}
if [ $# -eq 0 ]; then
  run K3M_DP_FUSED=1
  run K3M_DP_FUSED=1 K3M_BUCKET_MB=64
  run K3M_DP_FUSED=1 K3M_BUCKET_MB=128
  run K3M_DP_FUSED=0 K3M_COMM_BF16=1
else
  for v in "$@"; do run $v; done
fi
cat $out
