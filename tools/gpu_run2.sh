#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2_tests2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests2.log
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_match --launch-skip 3 --launch-count 1 -o gpurun_out/r2_kmatch_a -f python tools/time_match.py 4096 1048576 16 > gpurun_out/r2_ncu_a.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:k_extract --launch-skip 3 --launch-count 1 -o gpurun_out/r2_kextract_a -f python tools/time_match.py 4096 1048576 16 > gpurun_out/r2_ncu_b.log 2>&1
tail -30 gpurun_out/r2_tests2.log; tail -3 gpurun_out/r2_ncu_a.log
