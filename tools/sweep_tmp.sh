for v in nolds nofma; do
  export HOGB200_LIB=build_variants/libhogb200_$v.so
  echo "== $v"; timeout -s KILL 120 python tools/prof_batch.py 6 | grep -o "svm_ms.: [0-9.]*"
done
