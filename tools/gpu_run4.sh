#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
L=gpurun_out/r2_tm4.log
for u in 0 2 3 6 8; do
  echo "== units_per_cta $u: 4096 1048576 16" >> $L
  if [ $u = 0 ]; then timeout 600 python tools/time_match.py 4096 1048576 16 --check 32 >> $L 2>&1; else DSM_UNITS_PER_CTA=$u timeout 600 python tools/time_match.py 4096 1048576 16 >> $L 2>&1; fi
done
echo "== no-pair" >> $L; timeout 600 python tools/time_match.py 4096 1048576 16 --no-pair >> $L 2>&1
for u in 0 2 8; do
  echo "== units_per_cta $u: 16384 8388608 16 --seed 4" >> $L
  if [ $u = 0 ]; then timeout 600 python tools/time_match.py 16384 8388608 16 --seed 4 --check 8 >> $L 2>&1; else DSM_UNITS_PER_CTA=$u timeout 600 python tools/time_match.py 16384 8388608 16 --seed 4 >> $L 2>&1; fi
done
grep -v "^columns\|^verified" $L
