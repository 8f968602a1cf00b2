#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_tests1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests1.log
for args in "4096 1048576 16 --check 64" "4096 1048576 16 --no-pair" "4096 1048576 16 --no-order" "4096 1048576 16 --no-pair --no-order" "16384 8388608 16 --seed 4 --check 16" "16384 8388608 16 --seed 4 --no-order" "1 1048576 50" "128 1048576 50" "1 10000 50"; do
  echo "== $args" >> gpurun_out/r2_tm1.log
  timeout 600 python tools/time_match.py $args >> gpurun_out/r2_tm1.log 2>&1
done
tail -5 gpurun_out/r2_tests1.log; cat gpurun_out/r2_tm1.log
