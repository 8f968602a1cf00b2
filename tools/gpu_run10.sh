#!/bin/bash
# ncu evidence for profiles/: launch list of a bench run + full captures of k_match and k_extract
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 600 python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2_prof_bench.json 2> gpurun_out/r2_prof_bench.err || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_cfg3.csv python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2_prof_ncu1.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_match --launch-skip 4 --launch-count 1 -o gpurun_out/r02_k_match -f python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2_prof_ncu2.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_extract --launch-skip 4 --launch-count 1 -o gpurun_out/r02_k_extract -f python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2_prof_ncu3.log 2>&1
ls -la gpurun_out | tail -8
