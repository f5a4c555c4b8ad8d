#!/bin/bash
# Round-end evidence on ONE GPU: bench lines (no profiler), then the ncu launch list of the same command, then one
# `ncu --set full` capture per kernel of a 4-frame group launch (tools/prof_batch.py).  Outputs in gpurun_out/.
set -x
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err || exit 1
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2>> gpurun_out/bench_n1.err
for w in vga kitti 4k dense; do
  timeout 300 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/bench_${w}_n1.json 2> gpurun_out/bench_${w}_n1.err
done
python bench.py --steps 2 --warmup 1 --no-cpu > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_launches.log 2>&1
python tools/prof_batch.py 3 > /dev/null 2>&1 || exit 1
for k in hist_kernel resize_c4_kernel svm_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 1 --launch-count 1 \
      -o gpurun_out/prof_$k -f python tools/prof_batch.py 3 > gpurun_out/ncu_$k.log 2>&1
done
ls -la gpurun_out
