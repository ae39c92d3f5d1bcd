run() { name=$1; shift; env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('$name', d['value'], d['ms_per_step'], d['e2e']['ms_per_step'])" ; }
run b512 K3M_BUCKET_MB=512
run b256prio K3M_BUCKET_MB=256 K3M_COMM_PRIO=1
run b256ctas32 K3M_BUCKET_MB=256 NCCL_MIN_CTAS=32
