#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
for args in "4096 1048576 16 --check 64" "4096 1048576 16 --no-pair" "16384 8388608 16 --seed 4 --check 16" "1 1048576 50" "1 10000 50"; do
  echo "== $args" >> gpurun_out/r2_tm3.log
  timeout 600 python tools/time_match.py $args >> gpurun_out/r2_tm3.log 2>&1
done
cat gpurun_out/r2_tm3.log
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r2_tests3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests3.log
tail -5 gpurun_out/r2_tests3.log
