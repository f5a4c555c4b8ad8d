mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu -k "tail or group or dense or submit or hits or edge" 2>&1 | tail -3
python tools/prof_tail.py 640 480 1 256 4 | tail -1
python tools/prof_group.py 2>&1 | tail -4
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/prof_tail_vga_launches6.csv python tools/prof_tail.py 640 480 1 256 1 > gpurun_out/prof_tail_ncu.log 2>&1
