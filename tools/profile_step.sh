#!/bin/bash
# Round-2 profile of the captured step (one GPU): plain run, then the ncu launch list of one replayed step and one
# `--set full` capture of the dominant GEMM kernel.  Writes gpurun_out/${TAG}_*.
TAG=${1:-r02}
CMD="python bench.py --steps 1 --warmup 3 --no-breakdown --no-cpu-baseline"
KERN='regex:gemm_kernel|ln_fwd|ln_bwd|attn_|adamw|colsum|xent_kernel|kl_kernel|embed_|fuse_|relu_|pool_|pair_gather|struct_|lpm_|itm_|cast_|axpy_|lincomb|scale_|mask_|loc_|count_|dropout_|gate_|dgelu'
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --graph-profiling node -k "$KERN" -s 1300 -c 660 \
    --csv --log-file gpurun_out/${TAG}_ncu_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_launches.log
ncu --set full --clock-control none --import-source on --graph-profiling node -k regex:gemm_kernel -s 1400 -c 8 \
    -f -o gpurun_out/${TAG}_gemm_full $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_full.log
ls -la gpurun_out/${TAG}_*
