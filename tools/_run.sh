python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for l in 2 3 4; do for w in 1080p vga kitti 4k dense; do steps=10; [ $w = 4k ] && steps=4; HOGB200_LANES=$l python bench.py --workload $w --quick --no-cpu --steps $steps --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('lanes $l $w', round(d['value']), round(d['e2e']['value']))"; done; done
