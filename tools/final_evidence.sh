#!/bin/bash
# Round-end evidence on ONE B200 (run under gpurun): the full GPU test-suite, every bench table, both bench arms, the
# Llama step in both spellings, the ncu launch list of the bench command and full captures of the kernels that changed.
# Everything lands in gpurun_out/; the summaries that are to be judged are then copied into profiles/ by hand.
set -u
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -q --timeout 900 > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/pytest_gpu.log; tail -3 $out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/smoke.log 2>&1; tail -1 $out/smoke.log
python tools/bench_all.py --no-train-step > $out/bench_all.txt 2>&1
python tools/bench_regionc.py > $out/bench_regionc.txt 2>&1
python bench.py --impl reference > $out/bench_reference_arm.json 2> $out/bench_reference_arm.err
python bench.py > $out/bench_default.json 2> $out/bench_default.err
python bench.py --workload llama_step --spelling litgpt > $out/llama_n1_litgpt.json 2> $out/llama_n1_litgpt.err
python bench.py --workload llama_step --spelling ops > $out/llama_n1_ops.json 2> $out/llama_n1_ops.err
# ncu only after the plain commands above exited: launch list of the bench command, then full captures
python bench.py --steps 2 --warmup 3 --no-train-step --no-cpu-baseline > /dev/null 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/launches_default.csv \
      python bench.py --steps 2 --warmup 3 --no-train-step --no-cpu-baseline > $out/ncu_launches.log 2>&1
for wl in rmsnorm layernorm bias_gelu; do
  python bench.py --workload $wl --profile --steps 1 --warmup 3 > $out/plain_$wl.log 2>&1 || continue
  ncu --set full --clock-control none --import-source on --profile-from-start off -o $out/ncu_$wl -f python bench.py --workload $wl --profile --steps 1 --warmup 3 > $out/ncu_$wl.log 2>&1
  python tools/ncu_summary.py $out/ncu_$wl.ncu-rep > $out/ncu_${wl}_summary.txt 2>&1
  rm -f $out/ncu_$wl.ncu-rep
done
B200_PROFILE=1 python tools/bench_regionc.py > /dev/null 2>&1 && \
  B200_PROFILE=1 ncu --set full --clock-control none --kernel-name regex:rk_ -o $out/ncu_regionc -f python tools/bench_regionc.py > $out/ncu_regionc.log 2>&1
python tools/ncu_summary.py $out/ncu_regionc.ncu-rep > $out/ncu_regionc_summary.txt 2>&1
rm -f $out/ncu_regionc.ncu-rep
ls -la $out | tail -40
