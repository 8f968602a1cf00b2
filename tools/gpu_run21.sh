#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/r2_tests21.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests21.log
tail -6 gpurun_out/r2_tests21.log
