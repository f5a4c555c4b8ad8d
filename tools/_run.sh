HOGB200_TRACE=1 python tools/trace_batch.py 2> gpurun_out/r02d_trace.txt; tail -12 gpurun_out/r02d_trace.txt
sed -i 's/r02a_prof_/r02d_prof_/' tools/prof_three_kernels.sh
bash tools/prof_three_kernels.sh 2>&1 | tail -3
