#!/bin/bash
# The bench line of every BASELINE workload at N=1 (plus the reference arm), into gpurun_out/.
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2>> gpurun_out/bench_n1.err
for w in vga kitti 4k dense; do
  timeout 300 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/bench_${w}_n1.json 2> gpurun_out/bench_${w}_n1.err
done
for f in gpurun_out/bench_n1.json gpurun_out/bench_vga_n1.json gpurun_out/bench_kitti_n1.json gpurun_out/bench_4k_n1.json gpurun_out/bench_dense_n1.json; do python -c "
import json
d=json.load(open('$f')); r=d['roofline']
print('$f', round(d['value']), round(d['e2e']['value']), round(d.get('cpu_baseline',{}).get('value',0),1), r['frames_per_launch'], {k: round(v,3) for k,v in r['stage_ms_per_launch'].items()}, round(r['frac'],3))"; done
