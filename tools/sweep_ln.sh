#!/bin/bash
# A/B sweep of the LayerNorm knobs (layernorm.cu) on the bench workload.
run() { echo "== $*"; env "$@" python bench.py --workload layernorm --steps 300 --warmup 20 --no-cpu-baseline --no-e2e > /tmp/ln.json 2>/tmp/ln.err || tail -3 /tmp/ln.err; python tools/show_bench.py /tmp/ln.json | cut -c1-130; }
run B200K_LN_BWD_VPL=0
for shape in 32768,768 32768,1024 16384,2048 4096,8192; do
  run B200_BENCH_SHAPE=$shape
  run B200_BENCH_SHAPE=$shape B200K_LN_BWD_VPL=1
  run B200_BENCH_SHAPE=$shape B200K_LN_BWD_VPL=2
done
