#!/bin/bash
# One `ncu --set full` capture per family kernel (run under gpurun, 1 GPU): each workload of bench.py is run in
# --profile mode (3 warm-up steps, then 1 plain step); ncu skips the warm-up launches and captures the step's kernels.
# Reports land in gpurun_out/ncu_<workload>.ncu-rep; tools/ncu_summary.py turns them into the profiles/ summaries.
set -u
out=${1:-gpurun_out}
for wl in rmsnorm swiglu cross_entropy qkv_rope bias_gelu layernorm softmax; do
  python bench.py --workload $wl --profile --steps 1 --warmup 3 > $out/plain_$wl.log 2>&1 || { echo "plain run of $wl failed"; continue; }
  n=$(python -c "import json,sys; print(json.loads(open('$out/plain_$wl.log').read().strip().splitlines()[-1])['launches_per_step'])")
  ncu --set full --clock-control none --import-source on -s $((3 * n)) -c $n -o $out/ncu_$wl -f python bench.py --workload $wl --profile --steps 1 --warmup 3 > $out/ncu_$wl.log 2>&1
done
