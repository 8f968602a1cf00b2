#!/bin/bash
cd "$GRAFT_REPO_ROOT"
run() { echo "== $*"; env "${@:4}" python tools/time_match.py $1 $2 $3 --check 16 2>&1 | grep -v "^columns" | tail -3; }
run 4096 1048576 16 X=1
run 4096 4194304 16 X=1
run 16384 8388608 16 X=1
