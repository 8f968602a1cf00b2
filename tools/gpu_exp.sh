#!/bin/bash
cd "$GRAFT_REPO_ROOT"
echo "== affine (default)"; python tools/time_match.py 4096 1048576 16 --check 64 2>&1 | grep -v "^columns" | tail -3
echo "== DSM_AFFINE=0"; DSM_AFFINE=0 python tools/time_match.py 4096 1048576 16 2>&1 | grep -v "^columns" | tail -2
for w in 2 3 4 8; do echo "== affine DSM_WPW=$w"; DSM_WPW=$w python tools/time_match.py 4096 1048576 16 2>&1 | grep -v "^columns" | tail -2; done
echo "== 1024 jobs x 2M"; python tools/time_match.py 1024 2097152 16 --check 32 2>&1 | grep -v "^columns" | tail -3
echo "== 1024 jobs x 2M DSM_AFFINE=0"; DSM_AFFINE=0 python tools/time_match.py 1024 2097152 16 2>&1 | grep -v "^columns" | tail -2
timeout 600 python -m pytest tests/test_gpu_topk.py tests/test_abi_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
