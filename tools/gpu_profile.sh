#!/bin/bash
# ncu evidence for profiles/ (run under gpurun on one GPU): the launch list of a bench run, full
# captures of k_match (config 3 and config 4) and, with "all", of k_extract and k_build_tables.
# Every command is run without ncu first; numbers printed under ncu are never bench values.
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline"
timeout 600 $B > gpurun_out/prof_bench.json 2> gpurun_out/prof_bench.err || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_cfg3.csv $B > gpurun_out/prof_ncu1.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_match --launch-skip 4 --launch-count 1 -o gpurun_out/r02_k_match -f $B > gpurun_out/prof_ncu2.log 2>&1
if [ "$1" = "all" ]; then
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_extract --launch-skip 4 --launch-count 1 -o gpurun_out/r02_k_extract -f $B > gpurun_out/prof_ncu3.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:k_build_tables --launch-skip 4 --launch-count 1 -o gpurun_out/r02_k_build -f $B > gpurun_out/prof_ncu5.log 2>&1
fi
timeout 900 $B --workload cfg4 > gpurun_out/prof_bench4.json 2> gpurun_out/prof_bench4.err || exit 1
timeout 1200 ncu --set full --clock-control none -k regex:k_match --launch-skip 4 --launch-count 1 -o gpurun_out/r02_k_match_cfg4 -f $B --workload cfg4 > gpurun_out/prof_ncu4.log 2>&1
ls -la gpurun_out | grep r02_
