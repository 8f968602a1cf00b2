#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/r2_tests13.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests13.log
tail -4 gpurun_out/r2_tests13.log
( time timeout 1500 python bench.py ) > gpurun_out/r2_bench13.json 2> gpurun_out/r2_bench13.err; tail -4 gpurun_out/r2_bench13.err
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/r2_bench13_ref.json 2> gpurun_out/r2_bench13_ref.err; tail -4 gpurun_out/r2_bench13_ref.err; cut -c1-400 gpurun_out/r2_bench13_ref.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench13.json').read().strip().splitlines()[-1])
print(json.dumps({k:d[k] for k in ('value','ms_per_step','e2e','e2e_api','step_breakdown_ms','parity','gpu_launches','clocks')}, indent=1))
r=d['roofline']; print({k:r[k] for k in ('frac','kernel_ms','traffic','traffic_source') if k in r})
a=d['strong_scaling_anchor']; print({k:a[k] for k in ('value','ms_per_step','kernel_ms','roofline_hbm_fraction','wall_seconds')})
for k,v in d['other_configs'].items(): print(k, {x:y for x,y in v.items() if x.startswith(('p50','parity','wall','error','evals'))})
PY
