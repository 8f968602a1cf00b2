#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
L=gpurun_out/r2_tm12.log
for rep in 1 2 3; do
for lib in libdsmatch.so libdsm_noev.so; do
for args in "4096 1048576 16" "16384 8388608 16 --seed 4"; do
  echo "== $lib $args" >> $L
  DSM_LIB=dstack_b200/$lib timeout 600 python tools/time_match.py $args 2>&1 | grep "^build" >> $L
done; done; done
cat $L
