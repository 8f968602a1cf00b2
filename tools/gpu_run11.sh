#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
L=gpurun_out/r2_tm11.log
for args in "4096 1048576 16 --check 64" "16384 8388608 16 --seed 4 --check 16"; do
  echo "== $args" >> $L
  timeout 600 python tools/time_match.py $args >> $L 2>&1
done
grep -v "^columns" $L
timeout 1500 python -m pytest tests/test_gpu_topk.py tests/test_abi_gpu.py tests/test_device_pack.py -m gpu -q -x > gpurun_out/r2_tests11.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests11.log
tail -5 gpurun_out/r2_tests11.log
