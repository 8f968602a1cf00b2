#!/bin/bash
# final single-GPU validation: the GPU test suite, smoke(), the default bench line
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/final_tests.log 2>&1
tail -4 gpurun_out/final_tests.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
( time timeout 1200 python bench.py ) > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
tail -4 gpurun_out/final_bench.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/final_bench.json').read().strip().splitlines() if l.startswith('{')][-1])
keep=('value','ms_per_step','e2e','e2e_api','step_breakdown_ms','gpu_launches','clocks','parity')
print(json.dumps({k:d.get(k) for k in keep}, indent=1)[:3000])
r=d['roofline']; print({k:r.get(k) for k in ('frac','kernel_ms','traffic','lsu_fraction','issue_fraction','dram_fraction','traffic_source')})
print(json.dumps(d.get('strong_scaling_anchor'),indent=1)[:1500])
print(json.dumps(d.get('other_configs'),indent=1)[:4000])
print(json.dumps(d.get('cpu_baseline'),indent=1)[:1500])
PY
