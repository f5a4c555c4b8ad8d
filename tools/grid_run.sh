#!/bin/bash
# usage: gpurun --gpus N -- 'bash tools/grid_run.sh N'
# The bench line of every BASELINE workload at N GPUs (north_star: "frames/s on synthetic frames of each named shape at
# 1, 2, 4 and 8 GPUs") plus the copy-only H2D ceiling of the box at N ranks -> gpurun_out/r02_grid_<workload>_n<N>.json
N=${1:-1}
mkdir -p gpurun_out
port=29600
run() {   # run <outfile> <script args...>
  out=$1; shift
  if [ $N = 1 ]; then timeout 600 python "$@" > $out 2> ${out%.json}.err
  else port=$((port+1)); timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@" > $out 2> ${out%.json}.err; fi
}
run gpurun_out/r02_h2d_ceiling_n$N.json tools/micro/h2d_concurrent.py
cat gpurun_out/r02_h2d_ceiling_n$N.json
for w in 1080p vga kitti 4k dense; do
  steps=10; [ $w = 4k ] && steps=4
  run gpurun_out/r02_grid_${w}_n$N.json bench.py --gpus $N --workload $w --steps $steps --warmup 3 --quick --no-cpu
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/r02_grid_${w}_n$N.json") if l.startswith("{")][-1]); r = d["roofline"]
    print("$w", "N=%d" % d["n_gpus"], "resident %.0f" % d["value"], "e2e %.0f" % d["e2e"]["value"], "hist frac %.3f" % r["frac"],
          {k: round(v, 3) for k, v in r["stage_ms_per_launch"].items()}, d["clocks"])
except Exception as e:
    print("$w", "FAILED", e); print(open("gpurun_out/r02_grid_${w}_n$N.err").read()[-1500:])
PY
done
