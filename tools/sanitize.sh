#!/bin/bash
# usage (ONE tool per gpurun call -- see /opt/skills/guides/B200_PROFILING.md):
#   gpurun -- 'bash tools/sanitize.sh memcheck|racecheck|synccheck|initcheck'
# Runs tools/sanitize_case.py plainly first (it must exit 0), then under compute-sanitizer; the condensed log goes to
# gpurun_out/sanitizer_<tool>.txt (copied to profiles/ by hand).
set -x
tool=${1:-memcheck}
size=full; [ "$tool" = racecheck ] && size=small
mkdir -p gpurun_out
python tools/sanitize_case.py $size > gpurun_out/sanitize_plain.log 2>&1 || { tail -20 gpurun_out/sanitize_plain.log; exit 1; }
extra=""
[ "$tool" = racecheck ] && extra="--racecheck-report all"
[ "$tool" = memcheck ] && extra="--leak-check no"
timeout 1500 compute-sanitizer --tool $tool $extra --print-limit 50 python tools/sanitize_case.py $size > gpurun_out/sanitizer_${tool}_full.log 2>&1
rc=$?
{ echo "# compute-sanitizer --tool $tool $extra  python tools/sanitize_case.py $size   (exit code $rc)";
  grep -v "^=========     at \|^=========     by \|^=========         in \|Host Frame\|^=========$" gpurun_out/sanitizer_${tool}_full.log | tail -60; } > gpurun_out/sanitizer_${tool}.txt
tail -15 gpurun_out/sanitizer_${tool}.txt
exit $rc
