#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
L=gpurun_out/r2_tm18.log; rm -f $L
for args in "4096 1048576 16 --check 64" "16384 8388608 16 --seed 4 --check 16" "16384 8388608 16 --seed 4 --no-pair"; do
  echo "== $args" >> $L
  timeout 600 python tools/time_match.py $args >> $L 2>&1
done
for lib in libdsm_lr1.so libdsm_lr3.so libdsm_lr4.so; do
  echo "== $lib" >> $L; DSM_LIB=dstack_b200/$lib timeout 600 python tools/time_match.py 4096 1048576 16 2>&1 | grep "^build" >> $L
done
grep -v "^columns" $L
timeout 1500 python -m pytest tests/test_gpu_topk.py tests/test_abi_gpu.py tests/test_device_pack.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r2_tests18.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests18.log
tail -5 gpurun_out/r2_tests18.log
