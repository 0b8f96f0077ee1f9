#!/bin/bash
# One `ncu --set full` capture per family kernel (run under gpurun, 1 GPU): each workload of bench.py is run in
# --profile mode (3 warm-up steps, then 1 plain step between cudaProfilerStart/Stop); ncu captures that step's kernels.
# Each report is summarised on the box (tools/ncu_summary.py -> $out/ncu_<workload>_summary.txt) and then deleted:
# seven full reports exceed what gpurun brings back.
set -u
out=${1:-gpurun_out}
for wl in rmsnorm swiglu cross_entropy qkv_rope bias_gelu layernorm softmax; do
  python bench.py --workload $wl --profile --steps 1 --warmup 3 > $out/plain_$wl.log 2>&1 || { echo "plain run of $wl failed"; continue; }
  ncu --set full --clock-control none --import-source on --profile-from-start off -o $out/ncu_$wl -f python bench.py --workload $wl --profile --steps 1 --warmup 3 > $out/ncu_$wl.log 2>&1
  python tools/ncu_summary.py $out/ncu_$wl.ncu-rep > $out/ncu_${wl}_summary.txt 2>&1
  rm -f $out/ncu_$wl.ncu-rep
done
