#!/bin/bash
# scratch: k_extract with counts held 16 consecutive slots per lane
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
python tools/time_match.py 4096 1048576 16 --check 64 2>&1 | tail -2
python tools/time_match.py 16384 8388608 16 --check 16 2>&1 | tail -2
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
