#!/bin/bash
# Profile of the captured step of the final build (one GPU): ncu launch list of one replayed step + summary, one
# `--set full` capture of eight GEMM launches (DRAM traffic, tensor-pipe activity), then the exposed-time table.
TAG=${1:-r03}
O=gpurun_out
mkdir -p $O
CMD="python bench.py --steps 1 --warmup 3 --no-breakdown --no-cpu-baseline"
KERN='regex:gemm_kernel|ln_fwd|ln_bwd|attn_|adamw|colsum|xent_|kl_kernel|embed_|fuse_|relu_|pool_|pair_gather|struct_|lpm_|itm_|cast_|axpy|lincomb|scale_|mask_|loc_|count_|dropout_|gate_|dgelu|gather_rows|scatter_add'
$CMD > $O/${TAG}_plain.log 2>&1 || { tail -5 $O/${TAG}_plain.log; exit 1; }
L=$(python -c "import json;print(json.loads(open('$O/${TAG}_plain.log').read().strip().splitlines()[-1])['launches_per_step'])")
echo "launches per step: $L"
# warm-up: 1 eager + 1 capture + replays; skip generously, then take two steps' worth and summarise the last one
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --graph-profiling node -k "$KERN" -s $((L * 2 + 100)) -c $((L * 2)) \
    --csv --log-file $O/${TAG}_ncu_launches.csv $CMD > $O/${TAG}_ncu_launches.log 2>&1
tail -1 $O/${TAG}_ncu_launches.log
python tools/ncu_summary.py $O/${TAG}_ncu_launches.csv > $O/${TAG}_ncu_summary.csv 2>&1
head -12 $O/${TAG}_ncu_summary.csv
timeout 600 ncu --set full --clock-control none --import-source on --graph-profiling node -k regex:gemm_kernel -s $((L)) -c 8 \
    -f -o $O/${TAG}_gemm_full $CMD > $O/${TAG}_ncu_full.log 2>&1
ncu -i $O/${TAG}_gemm_full.ncu-rep --page raw --csv > $O/${TAG}_gemm_full_raw.csv 2>/dev/null
python tools/ncu_full_summary.py $O/${TAG}_gemm_full_raw.csv > $O/${TAG}_ncu_gemm_full.csv 2>&1
cat $O/${TAG}_ncu_gemm_full.csv
rm -f $O/${TAG}_gemm_full.ncu-rep
bash tools/exposed_time.sh $O/${TAG}_exposed_time.txt
