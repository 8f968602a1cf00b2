#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r2_tests5.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests5.log
tail -4 gpurun_out/r2_tests5.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke5.log 2>&1; tail -2 gpurun_out/r2_smoke5.log
( time timeout 1500 python bench.py ) > gpurun_out/r2_bench5.json 2> gpurun_out/r2_bench5.err; tail -5 gpurun_out/r2_bench5.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r2_bench5.json').read().strip().splitlines()[-1])
    print(json.dumps({k:d[k] for k in ('value','ms_per_step','e2e','e2e_api','step_breakdown_ms','parity')}, indent=1))
    print(json.dumps(d.get('roofline'), indent=1)[:1500])
    print(json.dumps(d.get('strong_scaling_anchor'), indent=1)[:2500])
    print(json.dumps(d.get('other_configs'), indent=1)[:4000])
    print(json.dumps(d.get('cpu_baseline'), indent=1)[:2000])
except Exception as e:
    print('parse failed', e); print(open('gpurun_out/r2_bench5.json').read()[-2000:])
PY
