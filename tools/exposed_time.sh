#!/bin/bash
# Exposed (critical-path) time of each kernel family of the captured step: the step is re-timed with that family's
# launches skipped (results are garbage, timing is not).  Output: $1
out=${1:-gpurun_out/exposed.txt}
shift
fams=("$@")
if [ ${#fams[@]} -eq 0 ]; then
  fams=("" k3m_ln_fwd k3m_ln_bwd k3m_attn_fwd k3m_attn_bwd k3m_colsum "k3m_adamw,k3m_adamw_bg" "k3m_ln_fwd,k3m_ln_bwd,k3m_attn_fwd,k3m_attn_bwd,k3m_colsum,k3m_adamw,k3m_adamw_bg" "k3m_gemm_bf16,k3m_gemm_grouped_bf16")
fi
: > $out
for skip in "${fams[@]}"; do
  echo "== skip: $skip" >> $out
  K3M_BENCH_SKIP="$skip" python bench.py --steps 10 --warmup 3 --no-breakdown 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ms_per_step %.3f  e2e %.3f  launches %d' % (d['ms_per_step'], d['e2e_ms_per_step'], d['launches_per_step']))" >> $out 2>&1
done
cat $out
