python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/prof_batch.py 6
python bench.py --workload vga --quick --no-cpu --steps 10 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('vga', d['value'], d['roofline']['stage_ms_per_launch'])"
HOGB200_LIB=build_variants/libhogb200_r1c2.so python bench.py --workload vga --quick --no-cpu --steps 10 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('vga chunks=2', d['value'], d['roofline']['stage_ms_per_launch'])"
python bench.py --workload kitti --quick --no-cpu --steps 10 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('kitti', d['value'], d['roofline']['stage_ms_per_launch'])"
