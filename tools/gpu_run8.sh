#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/r2_tests8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests8.log
tail -30 gpurun_out/r2_tests8.log
timeout 900 python bench.py --no-extras --no-cpu-baseline > gpurun_out/r2_bench8.json 2> gpurun_out/r2_bench8.err; tail -3 gpurun_out/r2_bench8.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench8.json').read().strip().splitlines()[-1])
print(json.dumps({k:d[k] for k in ('value','ms_per_step','e2e','e2e_api','step_breakdown_ms')}, indent=1))
PY
timeout 600 python bench.py --workload cfg2 2>&1 | tail -1 | cut -c1-300
timeout 600 python bench.py --workload cfg1 2>&1 | tail -1 | cut -c1-300
