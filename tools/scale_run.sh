#!/bin/bash
# Scaling evidence on one 8-GPU box: the bench line at N = 1, 2, 4, 8 (1080p), plus the 4K stream at N = 8.
# usage: tools/scale_run.sh ["1 2 4 8" [with4k]]   (an 8-GPU box is charged 8x: pass fewer N when minutes are short)
mkdir -p gpurun_out
NS=${1:-"1 2 4 8"}
port=29500
for n in $NS; do
  if [ $n = 1 ]; then python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu > gpurun_out/scale_1080p_n1.json 2> gpurun_out/scale_n1.err; continue; fi
  port=$((port+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/scale_1080p_n$n.json 2> gpurun_out/scale_n$n.err
done
port=$((port+1))
[ -n "$2" ] && timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus 8 --steps 6 --warmup 3 --workload 4k > gpurun_out/scale_4k_n8.json 2> gpurun_out/scale_4k_n8.err
for f in gpurun_out/scale_*.json; do python -c "
import json,sys
try:
    d=json.loads([l for l in open('$f') if l.startswith('{')][-1]); print('$f', d['n_gpus'], round(d['value'],1), round(d['e2e']['value'],1), d['clocks'])
except Exception as e: print('$f', 'FAILED', e)
"; done
tail -3 gpurun_out/scale_n8.err 2>/dev/null
