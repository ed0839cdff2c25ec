#!/usr/bin/env bash
# Runs every GPU test file in its own process (a trapped kernel kills only that file's CUDA context),
# full logs under gpurun_out/. Usage on the box:  bash tools/gpu_tests.sh [extra pytest args]
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.jsonl
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
rc=0
FILES=${GPU_TEST_FILES:-$(ls tests/test_*_gpu.py tests/test_zz_no_alloc.py)}
for f in $FILES; do
  n=$(basename "$f" .py)
  timeout 900 python -m pytest "$f" -q -m gpu -p no:cacheprovider "$@" > "gpurun_out/$n.log" 2>&1
  r=$?
  echo "== $f exit $r"; tail -n 25 "gpurun_out/$n.log" | cut -c1-300
  [ $r -ne 0 ] && rc=$r
done
exit $rc
