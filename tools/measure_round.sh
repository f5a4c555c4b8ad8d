#!/bin/bash
# Round-end evidence on ONE GPU (gpurun -- 'bash tools/measure_round.sh r02f'): the GPU tests, the bench lines (no
# profiler) of every workload + the reference arm, then the ncu launch list of the same bench command, then one
# `ncu --set full` capture per kernel of a 4-frame group launch (tools/prof_batch.py).  Outputs in gpurun_out/<tag>_*.
tag=${1:-r02f}
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/${tag}_pytest_gpu.txt
python bench.py --steps 20 --warmup 5 > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench_n1.err || exit 1
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2>> gpurun_out/${tag}_bench_n1.err
for w in vga kitti 4k dense; do
  steps=10; [ $w = 4k ] && steps=4
  timeout 400 python bench.py --workload $w --steps $steps --warmup 3 > gpurun_out/${tag}_bench_${w}_n1.json 2> gpurun_out/${tag}_bench_${w}_n1.err
done
for g in 2 8; do echo "frames per launch $g"; HOGB200_GROUP=$g python tools/prof_batch.py 6; done 2>&1 | tee gpurun_out/${tag}_group_probe.txt
python bench.py --steps 2 --warmup 3 --quick --no-cpu > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --quick --no-cpu > gpurun_out/${tag}_ncu_launches.log 2>&1
python tools/prof_batch.py 3 > /dev/null 2>&1 || exit 1
for k in hist_kernel resize_c4_kernel svm_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 1 --launch-count 1 \
      -o gpurun_out/${tag}_prof_$k -f python tools/prof_batch.py 3 > gpurun_out/${tag}_ncu_$k.log 2>&1
done
ls -la gpurun_out | tail -20
