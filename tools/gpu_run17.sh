#!/bin/bash
cd "$GRAFT_REPO_ROOT"
for lib in libdsmatch.so libdsm_lr1.so libdsm_lr3.so libdsm_lr4.so; do
  echo "== $lib"; DSM_LIB=dstack_b200/$lib timeout 600 python tools/time_match.py 4096 1048576 16 2>&1 | grep "^build"
  DSM_LIB=dstack_b200/$lib timeout 600 python tools/time_match.py 16384 8388608 16 --seed 4 2>&1 | grep "^build"
done
