set -x
mkdir -p gpurun_out
python tools/prof_batch.py 3 > gpurun_out/prof_batch_plain.log 2>&1 || exit 1
for k in hist_kernel resize_c4_kernel svm_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 1 --launch-count 1 \
      -o gpurun_out/r02a_prof_$k -f python tools/prof_batch.py 3 > gpurun_out/ncu_$k.log 2>&1
done
cat gpurun_out/prof_batch_plain.log
ls -la gpurun_out
