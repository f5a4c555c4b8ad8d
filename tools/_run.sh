python -m pytest tests -m gpu -x -q 2>&1 | tail -4
echo "== checks variant (device-side bounds traps)"; HOGB200_LIB=build_variants/libhogb200_checks.so python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/race_repeat.py 25 2>&1 | tee gpurun_out/r02_race_repeat.txt
python tools/prof_batch.py 6
python bench.py --workload vga --quick --no-cpu --steps 10 --warmup 3 | cut -c1-1400
