baseline/_ref/ref_cuda_bench 2>&1 | tee gpurun_out/r02_ref_cuda_kernels.txt
echo "--- ours, same GPU: one 1080p frame per launch (all 44 levels in 3 launches), then 4 frames per launch" | tee -a gpurun_out/r02_ref_cuda_kernels.txt
python tools/prof_one.py 6 2>&1 | tee -a gpurun_out/r02_ref_cuda_kernels.txt
python tools/prof_batch.py 6 2>&1 | tee -a gpurun_out/r02_ref_cuda_kernels.txt
